// calib_linear on one minimal sample (reference calib_ransac.py:51-131), executed by one thread group.
#pragma once
#include "rsx_internal.cuh"
#include "rsx_arrow.cuh"

// Workspace of one hypothesis (shared memory; H and V may live in global memory for many cameras).
struct FitWs {
  double* vs;    // [C][10][3] sampled unit bone directions
  double* ns;    // [C][10][3] sampled rays
  double* w;     // [C][10][3] R_c^T n_ci
  double* R;     // [C][9]
  double* G;     // [10][11]   Gram matrix of the stacked directions, then scratch
  double* GV;    // [10][11]   its eigenvectors
  double* H;     // [nu][ld]   normal matrix of the translation system, nu = 3 (10 + C)
  double* V;     // [nu][ld]   its eigenvectors
  double* cs;    // [nu/2 + 1][2] rotation pairs; later u [nu]
  double* misc;  // [64]
  double* kb;    // [nu][4]   basis of the 4 smallest eigen-directions of H
  int* iflag;    // [8]
  size_t v_doubles;   // capacity of V (nu * ld, or the arrow workspace in the compact layout)
};
__host__ __device__ inline size_t fit_ws_small_doubles(int C) {      // everything but H, V
  const int nu = 3 * (RSX_S + C);
  return (size_t)C * 30 * 3 + 9 * C + 2 * 110 + 2 * (nu / 2 + 1) + nu + 64 + 4 * nu + 8;
}
// Compact layout of the warp kernel with deferred fall-backs (no Jacobi runs in it, H lives in global memory): G, GV
// and kb alias the solver workspace V — G / GV behind the scratch of the orthogonal iteration (dead before the arrow
// solver starts), kb on the arrow solver's Q (its result) — and V holds the arrow workspace only: 12 instead of 22 KB
// per warp at 4 cameras.
__host__ __device__ inline size_t fit_ws_compact_small_doubles(int C) {
  const int nu = 3 * (RSX_S + C);
  return (size_t)C * 30 * 3 + 9 * C + 2 * (nu / 2 + 1) + nu + 64 + 8;
}
__host__ __device__ inline size_t fit_ws_compact_v_doubles(int C) {
  const size_t a = arrow_ws_doubles(C, 32), b = 320 + 220;
  return ((a > b ? a : b) + 1) & ~(size_t)1;
}
__host__ __device__ inline void fit_ws_carve(double* base, double* HV /*null: carve from base*/, int C, FitWs& ws,
                                             double* Hg = nullptr /*H alone in global memory, V carved from base*/,
                                             bool compact = false) {
  const int nu = 3 * (RSX_S + C), ld = nu | 1;
  double* p = base;
  ws.vs = p; p += C * 30; ws.ns = p; p += C * 30; ws.w = p; p += C * 30; ws.R = p; p += 9 * C;
  if (compact) {
    ws.cs = p; p += 2 * (nu / 2 + 1) + nu; ws.misc = p; p += 64;
    ws.iflag = (int*)p; p += 8;
    ws.H = Hg; ws.V = p;
    ws.G = ws.V + 320; ws.GV = ws.V + 430;
    ArrowWs aw; arrow_ws_carve(ws.V, C, 32, aw);
    ws.kb = aw.Q;
    ws.v_doubles = fit_ws_compact_v_doubles(C);
    return;
  }
  ws.G = p; p += 110; ws.GV = p; p += 110; ws.cs = p; p += 2 * (nu / 2 + 1) + nu; ws.misc = p; p += 64; ws.kb = p; p += 4 * nu;
  ws.iflag = (int*)p; p += 8;
  if (HV) { ws.H = HV; ws.V = HV + (size_t)nu * ld; }
  else if (Hg) { ws.H = Hg; ws.V = p; }
  else { ws.H = p; ws.V = p + (size_t)nu * ld; }
  ws.v_doubles = (size_t)nu * ld;
}

// v: [C][N][3], n: [C][M][3] (global); sv, sm: the 10 + 10 sample indices; Rt_out: [C][12] rows [R_c | t_c].
// defer = true: when an iterative solver does not settle the function returns false WITHOUT running the Jacobi
// fall-back (which costs ~50 normal fits); the caller queues the hypothesis for a second pass.
// DEFER_ONLY (compile time): the iterative solvers or nothing — no Jacobi code is generated at all.  The warp kernel
// with a second pass uses it: with the three Jacobi solvers inlined it was 19 k SASS instructions (300 KB) against a
// 32 KB instruction cache, and instruction fetch was its largest stall.
template <bool DEFER_ONLY, class G>
__host__ __device__ inline bool fit_hypothesis_t(const G& g, int C, int64_t N, int64_t M, const double* __restrict__ v,
                                                 const double* __restrict__ n, const int32_t* __restrict__ sv,
                                                 const int32_t* __restrict__ sm, const FitWs& ws, double* Rt_out,
                                                 bool fast, bool defer) {
  if (DEFER_ONLY) { fast = true; defer = true; }
  const int nu = 3 * (RSX_S + C), ld = nu | 1, T0 = 3 * RSX_S;
  // ---- 1. gather the sample
  for (int idx = g.tid; idx < C * 30; idx += g.nt) {
    const int c = idx / 30, k = (idx % 30) / 3, d = idx % 3;
    ws.vs[idx] = v[((size_t)c * N + sv[k]) * 3 + d];
    ws.ns[idx] = n[((size_t)c * M + sm[k]) * 3 + d];
  }
  g.sync();
  // ---- 2. rotations (calib_ransac.py:59-70).  The top-3 right singular vectors of the 10 x 3C stack V are
  // V^T u_k / sigma_k with u_k the top-3 eigenvectors of the 10 x 10 Gram matrix V V^T; any invertible 3x3
  // mixing (scale, sign, rotation) of those rows cancels in  R_all <- inv(R_all[:, :3]) R_all.
  for (int idx = g.tid; idx < 100; idx += g.nt) {
    const int a = idx / 10, b = idx % 10;
    double acc = 0.0;
    for (int c = 0; c < C; ++c)
      for (int d = 0; d < 3; ++d) acc += ws.vs[c * 30 + a * 3 + d] * ws.vs[c * 30 + b * 3 + d];
    ws.G[a * 11 + b] = acc;
  }
  g.sync();
  // small groups first try the orthogonal iteration for the dominant 3-dimensional subspace (rsx_arrow.cuh); the
  // Jacobi diagonalisation of the whole Gram matrix remains the general path and the fall-back
  bool have_u = false;
  if (fast && g.nt <= 32 && (size_t)TOP3_SCR_DOUBLES(g.nt) <= ws.v_doubles) {
    have_u = top3_subspace(g, ws.G, ws.V, ws.V + 220, ws.iflag + 9);
    if (have_u) {
      for (int idx = g.tid; idx < 30; idx += g.nt) ws.GV[(idx / 3) * 11 + idx % 3] = ws.V[220 + idx];
      if (g.tid == 0) { ws.iflag[1] = 0; ws.iflag[2] = 1; ws.iflag[3] = 2; }
    } else if (g.tid == 0) ws.iflag[8] = -1;
    g.sync();
    if (!have_u && defer) return false;          // uniform over the group
  }
  if (DEFER_ONLY && !have_u) return false;
  if (!DEFER_ONLY && !have_u) {
    cta_jacobi(g, ws.G, ws.GV, 10, 11, ws.cs, ws.iflag, 30);
    if (g.tid == 0) {                     // indices of the three largest eigenvalues
      int i0 = -1, i1 = -1, i2 = -1;
      for (int pass = 0; pass < 3; ++pass) {
        int bi = -1; double bv = 0;
        for (int i = 0; i < 10; ++i) {
          if (i == i0 || i == i1) continue;
          const double e = ws.G[i * 11 + i];
          if (bi < 0 || e > bv) { bi = i; bv = e; }
        }
        if (pass == 0) i0 = bi; else if (pass == 1) i1 = bi; else i2 = bi;
      }
      ws.iflag[1] = i0; ws.iflag[2] = i1; ws.iflag[3] = i2;
    }
    g.sync();
  }
  // A = U3^T V  (3 x 3C), stored per camera as A_c (3x3 row-major) in ws.w (free until step 3)
  for (int idx = g.tid; idx < 9 * C; idx += g.nt) {
    const int c = idx / 9, r = (idx % 9) / 3, d = idx % 3;
    const int col = ws.iflag[1 + r];
    double acc = 0.0;
    for (int a = 0; a < 10; ++a) acc += ws.GV[a * 11 + col] * ws.vs[c * 30 + a * 3 + d];
    ws.w[idx] = acc;
  }
  g.sync();
  for (int c = g.tid; c < C; c += g.nt) {        // R_c = (inv(A_0) A_c)^T
    double inv0[9];
    inv3(ws.w, inv0);
    const double* Ac = ws.w + 9 * c;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        double acc = 0.0;
        for (int k = 0; k < 3; ++k) acc += inv0[3 * j + k] * Ac[3 * k + i];
        ws.R[9 * c + 3 * i + j] = acc;            // R_c[i][j] = R_all[j][3c + i]
      }
  }
  g.sync();
  // ---- 3. translation system (calib_ransac.py:72-98), accumulated directly as H = C^T C
  for (int idx = g.tid; idx < C * 10; idx += g.nt) {          // w_ci = R_c^T n_ci   ("na @ Ra")
    const int c = idx / 10;
    const double* nn = ws.ns + idx * 3; const double* R = ws.R + 9 * c;
    for (int j = 0; j < 3; ++j) ws.w[idx * 3 + j] = R[j] * nn[0] + R[3 + j] * nn[1] + R[6 + j] * nn[2];
  }
  for (int idx = g.tid; idx < nu * ld; idx += g.nt) ws.H[idx] = 0.0;
  g.sync();
  const int n_xx = RSX_S, n_xt = RSX_S * C, n_tt = C * (C + 1) / 2;
  for (int task = g.tid; task < n_xx + n_xt + n_tt; task += g.nt) {
    double blk[9];
    for (int e = 0; e < 9; ++e) blk[e] = 0.0;
    int r0, c0;
    if (task < n_xx) {                                        // H[X_i, X_i] = sum_c Q^T Q,  Q = hat(n_ci) R_c
      const int i = task;
#pragma unroll 1
      for (int c = 0; c < C; ++c) {
        double Q[9]; hatmul(ws.ns + (c * 10 + i) * 3, ws.R + 9 * c, Q);
        for (int a = 0; a < 3; ++a)
          for (int b = 0; b < 3; ++b) blk[3 * a + b] += Q[a] * Q[b] + Q[3 + a] * Q[3 + b] + Q[6 + a] * Q[6 + b];
      }
      r0 = 3 * i; c0 = 3 * i;
    } else if (task < n_xx + n_xt) {                          // H[X_i, t_c] = Q^T hat(n_ci)
      const int i = (task - n_xx) / C, c = (task - n_xx) % C;
      const double* nn = ws.ns + (c * 10 + i) * 3;
      double Q[9]; hatmul(nn, ws.R + 9 * c, Q);
      const double Nx[9] = {0, -nn[2], nn[1], nn[2], 0, -nn[0], -nn[1], nn[0], 0};
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) blk[3 * a + b] = Q[a] * Nx[b] + Q[3 + a] * Nx[3 + b] + Q[6 + a] * Nx[6 + b];
      r0 = 3 * i; c0 = T0 + 3 * c;
    } else {                                                  // H[t_a, t_b], a <= b
      int t = task - n_xx - n_xt, a = 0;
      while (t >= C - a) { t -= C - a; ++a; }
      const int b = a + t;
      if (a == b) {
#pragma unroll 1
        for (int i = 0; i < 10; ++i) {                        // collinearity: hat(n)^T hat(n)
          const double* nn = ws.ns + (a * 10 + i) * 3;
          const double Nx[9] = {0, -nn[2], nn[1], nn[2], 0, -nn[0], -nn[1], nn[0], 0};
          for (int p = 0; p < 3; ++p)
            for (int q = 0; q < 3; ++q) blk[3 * p + q] += Nx[p] * Nx[q] + Nx[3 + p] * Nx[3 + q] + Nx[6 + p] * Nx[6 + q];
        }
        for (int o = 0; o < C; ++o) {                         // coplanarity rows of every pair containing a
          if (o == a) continue;
#pragma unroll 1
          for (int i = 0; i < 10; ++i) {
            double m[3], ra[3];
            cross3(ws.w + (a * 10 + i) * 3, ws.w + (o * 10 + i) * 3, m);
            const double* R = ws.R + 9 * a;
            for (int p = 0; p < 3; ++p) ra[p] = R[3 * p] * m[0] + R[3 * p + 1] * m[1] + R[3 * p + 2] * m[2];
            for (int p = 0; p < 3; ++p)
              for (int q = 0; q < 3; ++q) blk[3 * p + q] += ra[p] * ra[q];
          }
        }
      } else {
#pragma unroll 1
        for (int i = 0; i < 10; ++i) {                        // (R_a m)(-R_b m)^T,  m = w_ai x w_bi
          double m[3], ra[3], rb[3];
          cross3(ws.w + (a * 10 + i) * 3, ws.w + (b * 10 + i) * 3, m);
          const double* Ra = ws.R + 9 * a; const double* Rb = ws.R + 9 * b;
          for (int p = 0; p < 3; ++p) {
            ra[p] = Ra[3 * p] * m[0] + Ra[3 * p + 1] * m[1] + Ra[3 * p + 2] * m[2];
            rb[p] = -(Rb[3 * p] * m[0] + Rb[3 * p + 1] * m[1] + Rb[3 * p + 2] * m[2]);
          }
          for (int p = 0; p < 3; ++p)
            for (int q = 0; q < 3; ++q) blk[3 * p + q] += ra[p] * rb[q];
        }
      }
      r0 = T0 + 3 * a; c0 = T0 + 3 * b;
    }
    for (int p = 0; p < 3; ++p)
      for (int q = 0; q < 3; ++q) {
        ws.H[(r0 + p) * ld + c0 + q] = blk[3 * p + q];
        if (r0 != c0) ws.H[(c0 + q) * ld + r0 + p] = blk[3 * p + q];
      }
  }
  g.sync();
  // ---- 4. the four eigenvectors of the smallest eigenvalues (calib_ransac.py:96-107): only their span matters below.
  // Small groups (one warp / the host) first try the arrow-structured inverse subspace iteration, whose workspace
  // lives in V; the Jacobi diagonalisation of H remains the general path and the fall-back.
  bool have_kb = false;
  if (fast && g.nt <= 32 && arrow_ws_doubles(C, g.nt) <= ws.v_doubles) {
    ArrowWs aw; arrow_ws_carve(ws.V, C, g.nt, aw);
    // where a second pass exists (a deferred Jacobi fit costs ~50 fits and a kernel of its own) slow runs may go on
    have_kb = arrow_null4(g, ws.H, ld, C, ws.R, aw, ws.kb, ws.iflag + 8, defer ? 200 : ARW_MAX_SWEEPS);
    if (!have_kb && g.tid == 0) ws.iflag[8] = -1;
  }
  if (!have_kb && defer) return false;           // uniform over the group
  if (!DEFER_ONLY && !have_kb) {
    cta_jacobi(g, ws.H, ws.V, nu, ld, ws.cs, ws.iflag, 40);
    if (g.tid == 0) {
      int sel[4] = {-1, -1, -1, -1};
      for (int pass = 0; pass < 4; ++pass) {
        int bi = -1; double bv = 0;
        for (int i = 0; i < nu; ++i) {
          bool used = false;
          for (int q = 0; q < pass; ++q) used |= (sel[q] == i);
          if (used) continue;
          const double e = ws.H[i * ld + i];
          if (bi < 0 || e < bv) { bi = i; bv = e; }
        }
        sel[pass] = bi;
      }
      for (int j = 0; j < 4; ++j) ws.iflag[4 + j] = sel[j];
    }
    g.sync();
    for (int idx = g.tid; idx < nu * 4; idx += g.nt) ws.kb[idx] = ws.V[(idx >> 2) * ld + ws.iflag[4 + (idx & 3)]];
    g.sync();
  }
  if (g.tid == 0) {
    // coefficients that make t_0 = 0: null vector of the 3x4 block k[t_0 rows, :]  (calib_ransac.py:110-111);
    // the SVD's unit norm and sign cancel in the scaling by |t_1| and the z-test below.
    double Kt[3][4];
    for (int r = 0; r < 3; ++r)
      for (int j = 0; j < 4; ++j) Kt[r][j] = ws.kb[(T0 + r) * 4 + j];
    double cf[4];
    for (int j = 0; j < 4; ++j) {
      double mnr[9]; int cc = 0;
      for (int q = 0; q < 4; ++q) {
        if (q == j) continue;
        for (int r = 0; r < 3; ++r) mnr[3 * r + cc] = Kt[r][q];
        ++cc;
      }
      cf[j] = ((j & 1) ? -1.0 : 1.0) * det3(mnr);
    }
    for (int j = 0; j < 4; ++j) ws.misc[j] = cf[j];
  }
  g.sync();
  double* u = ws.cs;                                           // [nu] (rotation pairs no longer needed)
  for (int i = g.tid; i < nu; i += g.nt) {
    double acc = 0.0;
    for (int j = 0; j < 4; ++j) acc += ws.kb[i * 4 + j] * ws.misc[j];
    u[i] = acc;
  }
  g.sync();
  if (g.tid == 0) {
    const double* t1 = u + T0 + 3;
    ws.misc[8] = sqrt(t1[0] * t1[0] + t1[1] * t1[1] + t1[2] * t1[2]);      // s = |t_1|   (calib_ransac.py:114)
    ws.iflag[0] = 0; ws.iflag[1] = 0; ws.iflag[10] = 0;
  }
  g.sync();
  // ---- 5. z-test on cameras 0, 1 with both signs (util.py:161-180); counts with +t, as the reference does
  for (int task = g.tid; task < 20; task += g.nt) {
    const int k = task % 10; const double sg = task < 10 ? 1.0 : -1.0;
    const double s = ws.misc[8];
    double P0[12], P1[12], t0[3], t1[3];
    for (int r = 0; r < 3; ++r) {
      t0[r] = u[T0 + r] / s; t1[r] = u[T0 + 3 + r] / s;
      for (int j = 0; j < 3; ++j) { P0[4 * r + j] = ws.R[3 * r + j]; P1[4 * r + j] = ws.R[9 + 3 * r + j]; }
      P0[4 * r + 3] = sg * t0[r]; P1[4 * r + 3] = sg * t1[r];
    }
    double m[10];
    for (int e = 0; e < 10; ++e) m[e] = 0.0;
    dlt_accumulate(P0, ws.ns[k * 3], ws.ns[k * 3 + 1], m);
    dlt_accumulate(P1, ws.ns[30 + k * 3], ws.ns[30 + k * 3 + 1], m);
    // Rayleigh-quotient iteration, Jacobi only as its last resort (the Jacobi solver here cost 20 % of the warp kernel's
    // time in instruction fetch alone).  The sign decided below is a discontinuous function of the data (z ~ 0, a point
    // near infinity): for a degenerate sample it hangs on the last bits of R and X whatever the solver — one garbage
    // hypothesis of the 1500-iteration golden comes out with -t (its score, invariant to the sign, agrees to 1e-12).
    double X[4];
    if (DEFER_ONLY) { if (!eig4_smallest_rqi<true, false>(m, X)) ws.iflag[10] = 1; }
    else eig4_smallest_fast(m, X);
    const double x0 = X[0] / X[3], x1 = X[1] / X[3], x2 = X[2] / X[3];
    const double z0 = ws.R[6] * x0 + ws.R[7] * x1 + ws.R[8] * x2 + t0[2];
    const double z1 = ws.R[15] * x0 + ws.R[16] * x1 + ws.R[17] * x2 + t1[2];
    const int cnt = (z0 > 0 ? 1 : 0) + (z1 > 0 ? 1 : 0);
    ws.misc[16 + task] = (double)cnt;
  }
  g.sync();
  if (DEFER_ONLY && ws.iflag[10]) { if (g.tid == 0) ws.iflag[8] = -1; g.sync(); return false; }
  if (g.tid == 0) {
    double zp = 0, zn = 0;
    for (int k = 0; k < 10; ++k) { zp += ws.misc[16 + k]; zn += ws.misc[26 + k]; }
    ws.misc[9] = zp > zn ? 1.0 : -1.0;
  }
  g.sync();
  for (int idx = g.tid; idx < 12 * C; idx += g.nt) {
    const int c = idx / 12, r = (idx % 12) / 4, j = idx % 4;
    Rt_out[idx] = j < 3 ? ws.R[9 * c + 3 * r + j] : ws.misc[9] * u[T0 + 3 * c + r] / ws.misc[8];
  }
  g.sync();
  return true;
}

template <class G>
__host__ __device__ inline bool fit_hypothesis(const G& g, int C, int64_t N, int64_t M, const double* __restrict__ v,
                                               const double* __restrict__ n, const int32_t* __restrict__ sv,
                                               const int32_t* __restrict__ sm, const FitWs& ws, double* Rt_out,
                                               bool fast = true, bool defer = false) {
  return fit_hypothesis_t<false>(g, C, N, M, v, n, sv, sm, ws, Rt_out, fast, defer);
}

// ---------------------------------------------------------------------------------------------------
// Score pieces of one model (calib_ransac.py:154-180), per-thread partial sums over a strided range.
// reprojection error of the triangulated point X against ray i of every camera (calib_ransac.py:163-178), added to acc
__host__ __device__ inline void score_ray(int C, int64_t M, const double* __restrict__ n, const double* Rt, int64_t i,
                                          const double (&X)[4], double* e_out, const double* K00, double th2d,
                                          uint8_t* mask2d, double& acc) {
  bool inl = true;
  for (int c = 0; c < C; ++c) {
    const double* nn = n + ((size_t)c * M + i) * 3;
    const double* P = Rt + 12 * c;
    const double x0 = P[0] * X[0] + P[1] * X[1] + P[2] * X[2] + P[3] * X[3];
    const double x1 = P[4] * X[0] + P[5] * X[1] + P[6] * X[2] + P[7] * X[3];
    const double x2 = P[8] * X[0] + P[9] * X[1] + P[10] * X[2] + P[11] * X[3];
    // one reciprocal instead of the reference's three divisions (calib_ransac.py:163-165); x2 / x2 is 1 or NaN
    const double ix2 = 1.0 / x2;
    const double e0 = x0 * ix2 - nn[0], e1 = x1 * ix2 - nn[1], e2 = ((x2 - x2) == 0.0 && x2 != 0.0 ? 1.0 : x2 * ix2) - nn[2];
    acc += e0 * e0 + e1 * e1 + e2 * e2;
    if (e_out) { double* o = e_out + ((size_t)i * C + c) * 3; o[0] = e0; o[1] = e1; o[2] = e2; }
    if (mask2d) inl = inl && (K00[c] * sqrt(e0 * e0 + e1 * e1) <= th2d);
  }
  if (mask2d) mask2d[i] = inl ? 1 : 0;
}

// Sum-only variant for the RANSAC score pass (no error output, no masks): |e_ci|^2 = ((x0 - u x2)^2 + (x1 - v x2)^2) / x2^2
// (the third component x2/x2 - 1 of calib_ransac.py:163-165 is 0 or one rounding, ~1e-32 in the sum), and the
// per-camera quotients of up to four cameras are added over a common denominator, so a ray costs one FP64 division per
// four cameras instead of one per camera (~10 instructions each).  x2 = 0 gives inf / NaN, which never wins (:182).
__host__ __device__ inline void score_ray_sum(int C, int64_t M, const double* __restrict__ n, const double* Rt, int64_t i,
                                              const double (&X)[4], double& acc) {
  for (int c0 = 0; c0 < C; c0 += 4) {
    double num = 0.0, den = 1.0;
    const int c1 = c0 + 4 < C ? c0 + 4 : C;
    for (int c = c0; c < c1; ++c) {
      const double* nn = n + ((size_t)c * M + i) * 3;
      const double* P = Rt + 12 * c;
      const double x0 = P[0] * X[0] + P[1] * X[1] + P[2] * X[2] + P[3] * X[3];
      const double x1 = P[4] * X[0] + P[5] * X[1] + P[6] * X[2] + P[7] * X[3];
      const double x2 = P[8] * X[0] + P[9] * X[1] + P[10] * X[2] + P[11] * X[3];
      const double d0 = x0 - nn[0] * x2, d1 = x1 - nn[1] * x2;
      const double s = d0 * d0 + d1 * d1, q = x2 * x2;
      num = num * q + s * den;
      den = den * q;
    }
    acc += num / den;
  }
}

//   MODE 1 (default): Rayleigh-quotient eigen-solver + common-denominator sum; MODE 0: round-1 arithmetic
//   rays i0 + tid, i0 + tid + nt, ... < i1 of this thread added to acc (score_points_sum = the range [0, M); the
//   pruned score kernel walks the same sequence in pieces, so both give the same bits)
template <int MODE, class G>
__host__ __device__ inline void score_points_accum(const G& g, int C, int64_t M, const double* __restrict__ n, const double* Rt,
                                                   int64_t i0, int64_t i1, double& acc) {
  for (int64_t i = i0 + g.tid; i < i1; i += g.nt) {
    double m[10];
#pragma unroll
    for (int e = 0; e < 10; ++e) m[e] = 0.0;
    for (int c = 0; c < C; ++c) {
      const double* nn = n + ((size_t)c * M + i) * 3;
      dlt_accumulate(Rt + 12 * c, nn[0], nn[1], m);
    }
    double X[4];
    if (MODE == 1) {
      eig4_smallest_rqi<false>(m, X);
      score_ray_sum(C, M, n, Rt, i, X, acc);
    } else {
      eig4_smallest_invit(m, X);
      score_ray(C, M, n, Rt, i, X, nullptr, nullptr, 0.0, nullptr, acc);
    }
  }
}
template <int MODE, class G>
__host__ __device__ inline double score_points_sum(const G& g, int C, int64_t M, const double* __restrict__ n, const double* Rt) {
  double acc = 0.0;
  score_points_accum<MODE>(g, C, M, n, Rt, 0, M, acc);
  return acc;
}

//   Rt: [C][12] (shared);  returns sum over this thread's points of |e_ci|^2 (all cameras)
template <class G>
__host__ __device__ inline double score_points(const G& g, int C, int64_t M, const double* __restrict__ n,
                                               const double* Rt, double* e_out /*[M][C][3] or null*/,
                                               const double* K00 /*null unless masks*/, double th2d, uint8_t* mask2d) {
  double acc = 0.0;
  for (int64_t i = g.tid; i < M; i += g.nt) {
    double m[10];
#pragma unroll
    for (int e = 0; e < 10; ++e) m[e] = 0.0;
    for (int c = 0; c < C; ++c) {
      const double* nn = n + ((size_t)c * M + i) * 3;
      dlt_accumulate(Rt + 12 * c, nn[0], nn[1], m);
    }
    double X[4];
    eig4_smallest_fast(m, X);
    score_ray(C, M, n, Rt, i, X, e_out, K00, th2d, mask2d, acc);
  }
  return acc;
}

// Sum-only variant for the score pass: <R_0^T v_0, R_c^T v_c> = v_0^T (R_0 R_c^T) v_c, so with G_c = R_0 R_c^T formed once
// per hypothesis (Gc: [C][9] row-major, entry 0 unused) an orientation costs 9 (C - 1) + 3 FMA instead of 9 + 12 (C - 1).
template <class G>
__host__ __device__ inline double score_dirs_sum(const G& g, int C, int64_t N, const double* __restrict__ v, const double* Gc) {
  double acc = 0.0;
  for (int64_t i = g.tid; i < N; i += g.nt) {
    double w0 = 0.0, w1 = 0.0, w2 = 0.0;
    for (int c = 1; c < C; ++c) {
      const double* d = v + ((size_t)c * N + i) * 3;
      const double* Gm = Gc + 9 * c;
      w0 += Gm[0] * d[0] + Gm[1] * d[1] + Gm[2] * d[2];
      w1 += Gm[3] * d[0] + Gm[4] * d[1] + Gm[5] * d[2];
      w2 += Gm[6] * d[0] + Gm[7] * d[1] + Gm[8] * d[2];
    }
    const double* d0 = v + (size_t)i * 3;
    const double cbar = (d0[0] * w0 + d0[1] * w1 + d0[2] * w2) / (double)(C - 1);
    const double ang = acos(cbar > 1.0 ? 1.0 : cbar);        // NaN stays NaN, as np.min([x, 1]) does
    acc += ang * ang;
  }
  return acc;
}
// G_c = R_0 R_c^T from the models [C][12] (rows [R | t]); cooperative, the caller syncs afterwards
template <class G>
__host__ __device__ inline void score_dirs_prepare(const G& g, int C, const double* Rt, double* Gc) {
  for (int idx = g.tid; idx < 9 * C; idx += g.nt) {
    const int c = idx / 9, r = (idx % 9) / 3, q = idx % 3;
    // (R_0^T u) . (R_c^T w) = u^T R_0 R_c^T w: entry (r, q) = sum_j R_0[r][j] R_c[q][j]
    Gc[idx] = Rt[4 * r] * Rt[12 * c + 4 * q] + Rt[4 * r + 1] * Rt[12 * c + 4 * q + 1] + Rt[4 * r + 2] * Rt[12 * c + 4 * q + 2];
  }
}

//   returns sum over this thread's orientations of arccos(min(mean_{c>=1} <R_0^T v_0i, R_c^T v_ci>, 1))^2
template <class G>
__host__ __device__ inline double score_dirs(const G& g, int C, int64_t N, const double* __restrict__ v,
                                             const double* Rt, double* e3d_out, double th3d, uint8_t* mask3d) {
  double acc = 0.0;
  for (int64_t i = g.tid; i < N; i += g.nt) {
    const double* d0 = v + (size_t)i * 3;
    double a0[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) a0[j] = Rt[j] * d0[0] + Rt[4 + j] * d0[1] + Rt[8 + j] * d0[2];
    double s = 0.0;
    for (int c = 1; c < C; ++c) {
      const double* d = v + ((size_t)c * N + i) * 3;
      const double* P = Rt + 12 * c;
      double dot = 0.0;
#pragma unroll
      for (int j = 0; j < 3; ++j) dot += a0[j] * (P[j] * d[0] + P[4 + j] * d[1] + P[8 + j] * d[2]);
      s += dot;
    }
    const double cbar = s / (double)(C - 1);
    const double ang = acos(cbar > 1.0 ? 1.0 : cbar);        // NaN stays NaN, as np.min([x, 1]) does
    acc += ang * ang;
    if (e3d_out) e3d_out[i] = ang;
    if (mask3d) mask3d[i] = (ang <= th3d) ? 1 : 0;
  }
  return acc;
}
