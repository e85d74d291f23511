// The 4-dimensional invariant subspace of the smallest eigenvalues of the translation normal matrix H of one
// minimal sample (reference calib_ransac.py:96-107 takes it from a dense LAPACK eigh), by inverse subspace
// iteration that exploits H's arrow structure:
//
//   H = [ D  B ]   D = blockdiag(D_0..D_9), 3x3 (one per sampled ray),  B: 30 x 3C,  T: 3C x 3C
//       [ B' T ]
//   (H + s I)^-1 is applied through  D_i + s I = L_i L_i',  W_i = L_i^-1 B_i,  S = T + s I - sum_i W_i' W_i = Ls Ls'.
//
// For the minimal samples of this path lambda_4 / lambda_5 has median 0.013 and stays below ~0.2, so 6-20 sweeps
// reach round-off; a parallel Jacobi diagonalisation of the same 42 x 42 matrix costs ~50x the flops.  Any basis of
// the subspace serves the caller (it only needs the combination whose t_0 block vanishes).  Returns false when
// the sweep count runs out (pathological gap, NaN data): the caller then falls back to the Jacobi solver, so the
// result never depends on this shortcut converging.
//
// Written against the same cooperating-group context as rsx_fit.cuh (tid, nt, sync()); group-wide sums go through
// a [k][nt] scratch, so the code runs unchanged on one warp of the GPU and single-threaded on the host.
#pragma once
#include "rsx_internal.cuh"

#define ARW_K 4            // subspace dimension
#ifndef ARW_MAX_SWEEPS
#define ARW_MAX_SWEEPS 48
#endif
#ifndef ARW_LATE_TOL
#define ARW_LATE_TOL 1e-28      // settle threshold (squared movement x amplification) of slow runs (sweep >= ARW_LATE)
#endif
#ifndef ARW_LATE
#define ARW_LATE 32
#endif

struct ArrowWs {
  double* Ld;    // [10][9]  l00 l10 l11 l20 l21 l22 | 1/l00 1/l11 1/l22
  double* Wm;    // [30][3C]
  double* S;     // [3C][3C] lower Cholesky factor of the Schur complement
  double* rS;    // [3C]     reciprocal diagonal of that factor
  double* Q;     // [nu][4]
  double* Z;     // [nu][4]
  double* red;   // [16][nt] reduction scratch
  double* out;   // [32]     reduction results / small shared scalars
};
__host__ __device__ inline size_t arrow_ws_doubles(int C, int nt) {
  const int nu = 3 * (RSX_S + C), n3 = 3 * C;
  return 90 + (size_t)30 * n3 + (size_t)n3 * n3 + n3 + 2 * (size_t)nu * ARW_K + (nt == 32 ? 0 : 16 * (size_t)nt) + 32;   // a warp reduces by shuffles
}
__host__ __device__ inline void arrow_ws_carve(double* p, int C, int nt, ArrowWs& a) {
  const int nu = 3 * (RSX_S + C), n3 = 3 * C;
  a.Ld = p; p += 90; a.Wm = p; p += 30 * n3; a.S = p; p += n3 * n3; a.rS = p; p += n3;
  a.Q = p; p += nu * ARW_K; a.Z = p; p += nu * ARW_K; a.red = p; p += (nt == 32 ? 0 : 16 * nt); a.out = p;
}

// vals[j] <- sum over the group of vals[j], j < K <= 16, in every thread.  One warp on the device: xor-butterfly of
// shuffles (every lane ends with the bit-identical sum: a + b == b + a), no shared memory and no bank conflicts — the
// [16][nt] scratch of the generic version was 4 KB per warp and its second phase a 16-way conflict.  Any other group
// (CTA, host): through red[K][nt] / out[K] with two syncs.
template <int K, class G>
__host__ __device__ inline void group_allsum(const G& g, double (&vals)[K], double* red, double* out) {
  for (int j = 0; j < K; ++j) red[j * g.nt + g.tid] = vals[j];
  g.sync();
  for (int j = g.tid; j < K; j += g.nt) {
    double s = 0.0;
#pragma unroll 1
    for (int t = 0; t < g.nt; ++t) s += red[j * g.nt + t];
    out[j] = s;
  }
  g.sync();
  for (int j = 0; j < K; ++j) vals[j] = out[j];
  g.sync();                                              // out[] may be reused by the next call
}
#ifdef __CUDA_ARCH__
template <int K>
__device__ __forceinline__ void group_allsum(const WarpCtx& g, double (&vals)[K], double*, double*) {
  // the butterfly's five steps stay a LOOP: unrolled, the double-precision shuffles (2 SHFL + packing + the warp-sync
  // bracket each) were 7.2 k of the warp fit kernel's 15.4 k SASS instructions, and instruction fetch its top stall
#pragma unroll 1
  for (int m = 16; m >= 1; m >>= 1) {
#pragma unroll
    for (int j = 0; j < K; ++j) vals[j] += __shfl_xor_sync(0xffffffffu, vals[j], m);
  }
}
#endif

// Z <- orthonormal basis of span(Z) by Cholesky QR (Gram matrix -> 4x4 Cholesky -> Z R^-1); called twice per sweep,
// the second pass removes the O(eps cond^2) loss of orthogonality of the first.  Returns false on a non-positive pivot.
template <class G>
__host__ __device__ inline bool cholqr4(const G& g, double* Z, int nu, double* red, double* out) {
  double p[10];
  for (int e = 0; e < 10; ++e) p[e] = 0.0;
  for (int i = g.tid; i < nu; i += g.nt) {
    const double* z = Z + i * ARW_K;
    int e = 0;
    for (int a = 0; a < ARW_K; ++a)
      for (int b = a; b < ARW_K; ++b) p[e++] += z[a] * z[b];
  }
  group_allsum(g, p, red, out);
  // upper Cholesky factor R (Gram = R' R), computed redundantly by every thread
  const double m00 = p[0], m01 = p[1], m02 = p[2], m03 = p[3], m11 = p[4], m12 = p[5], m13 = p[6],
               m22 = p[7], m23 = p[8], m33 = p[9];
  if (!(m00 > 0)) return false;
  const double i0 = rsx_rsqrt_fast(m00);
  const double r01 = m01 * i0, r02 = m02 * i0, r03 = m03 * i0;
  const double d1 = m11 - r01 * r01;
  if (!(d1 > 0)) return false;
  const double i1 = rsx_rsqrt_fast(d1);
  const double r12 = (m12 - r01 * r02) * i1, r13 = (m13 - r01 * r03) * i1;
  const double d2 = m22 - r02 * r02 - r12 * r12;
  if (!(d2 > 0)) return false;
  const double i2 = rsx_rsqrt_fast(d2);
  const double r23 = (m23 - r02 * r03 - r12 * r13) * i2;
  const double d3 = m33 - r03 * r03 - r13 * r13 - r23 * r23;
  if (!(d3 > 0)) return false;
  const double i3 = rsx_rsqrt_fast(d3);
  for (int i = g.tid; i < nu; i += g.nt) {               // row z <- z R^-1 (forward substitution along the row)
    double* z = Z + i * ARW_K;
    const double q0 = z[0] * i0;
    const double q1 = (z[1] - q0 * r01) * i1;
    const double q2 = (z[2] - q0 * r02 - q1 * r12) * i2;
    const double q3 = (z[3] - q0 * r03 - q1 * r13 - q2 * r23) * i3;
    z[0] = q0; z[1] = q1; z[2] = q2; z[3] = q3;
  }
  g.sync();
  return true;
}

// Dominant 3-dimensional invariant subspace of the 10 x 10 Gram matrix of the stacked sample directions (the top-3
// right singular vectors of the reference's SVD at calib_ransac.py:59-63 are V^T u_k / sigma_k; any basis of the
// subspace serves, see fit_hypothesis).  Orthogonal iteration instead of the cyclic Jacobi diagonalisation of the whole
// matrix (~9 rounds x 3 synchronised phases x 6-8 sweeps, a fifth of the fit's instructions): six steps with G^4
// (contraction (s4/s3)^8 per step), then steps with G itself, Cholesky-QR twice, until the subspace moves by < 3e-15
// (or sits on its round-off floor)
// — G^4 only pre-conditions the start, the accuracy is that of iterating with G.  Returns false when it does not
// settle (s3 ~ s4: a degenerate sample); the caller then runs / defers to the Jacobi path, so the result never depends
// on this shortcut.  Gm: [10][11]; scr: 2 x 110 + 2 x 30 + 32 doubles (+ 16 nt unless one warp); U out: [10][3].
#define TOP3_SCR_DOUBLES(nt) (2 * 110 + 2 * 32 + 32 + ((nt) == 32 ? 0 : 16 * (nt)))
template <class G>
__host__ __device__ inline bool cholqr3(const G& g, double* Y /*[10][3]*/, double* red, double* out) {
  double p[6] = {0, 0, 0, 0, 0, 0};
  for (int a = g.tid; a < 10; a += g.nt) {
    const double y0 = Y[3 * a], y1 = Y[3 * a + 1], y2 = Y[3 * a + 2];
    p[0] += y0 * y0; p[1] += y0 * y1; p[2] += y0 * y2; p[3] += y1 * y1; p[4] += y1 * y2; p[5] += y2 * y2;
  }
  group_allsum(g, p, red, out);
  if (!(p[0] > 0)) return false;
  const double i0 = rsx_rsqrt_fast(p[0]);
  const double r01 = p[1] * i0, r02 = p[2] * i0;
  const double d1 = p[3] - r01 * r01;
  if (!(d1 > 0)) return false;
  const double i1 = rsx_rsqrt_fast(d1);
  const double r12 = (p[4] - r01 * r02) * i1;
  const double d2 = p[5] - r02 * r02 - r12 * r12;
  if (!(d2 > 0)) return false;
  const double i2 = rsx_rsqrt_fast(d2);
  for (int a = g.tid; a < 10; a += g.nt) {
    const double q0 = Y[3 * a] * i0;
    const double q1 = (Y[3 * a + 1] - q0 * r01) * i1;
    const double q2 = (Y[3 * a + 2] - q0 * r02 - q1 * r12) * i2;
    Y[3 * a] = q0; Y[3 * a + 1] = q1; Y[3 * a + 2] = q2;
  }
  g.sync();
  return true;
}
template <class G>
__host__ __device__ inline bool top3_subspace(const G& g, const double* Gm, double* scr, double* Uout, int* iters_out) {
  double* G2 = scr; double* G4 = scr + 110; double* U = scr + 220; double* Y = scr + 252; double* out = scr + 284;
  double* red = scr + 316;
  for (int idx = g.tid; idx < 100; idx += g.nt) {
    const int a = idx / 10, b = idx - 10 * a;
    double acc = 0.0;
    for (int k = 0; k < 10; ++k) acc += Gm[a * 11 + k] * Gm[k * 11 + b];
    G2[a * 11 + b] = acc;
  }
  for (int idx = g.tid; idx < 30; idx += g.nt) {
    const unsigned hsh = (unsigned)(idx + 1) * 2654435761u;
    U[idx] = (double)(hsh >> 8) * (1.0 / 16777216.0) - 0.5;
  }
  g.sync();
  for (int idx = g.tid; idx < 100; idx += g.nt) {
    const int a = idx / 10, b = idx - 10 * a;
    double acc = 0.0;
    for (int k = 0; k < 10; ++k) acc += G2[a * 11 + k] * G2[k * 11 + b];
    G4[a * 11 + b] = acc;
  }
  if (!cholqr3(g, U, red, out)) return false;          // (syncs inside)
  double prev = 1e300;
  int it = 0;
  for (; it < 48; ++it) {
    const bool pre = it < 6;
    const double* Mx = pre ? G4 : Gm;
    for (int idx = g.tid; idx < 30; idx += g.nt) {
      const int a = idx / 3, k = idx - 3 * a;
      double acc = 0.0;
      for (int b = 0; b < 10; ++b) acc += Mx[a * 11 + b] * U[3 * b + k];
      Y[idx] = acc;
    }
    g.sync();
    if (!cholqr3(g, Y, red, out)) return false;
    if (pre) {
      for (int idx = g.tid; idx < 30; idx += g.nt) U[idx] = Y[idx];
      g.sync();
      continue;
    }
    if (!cholqr3(g, Y, red, out)) return false;
    // change of the subspace: || Y - U (U' Y) ||_F^2
    double pm[9];
    for (int e = 0; e < 9; ++e) pm[e] = 0.0;
    for (int a = g.tid; a < 10; a += g.nt)
      for (int p = 0; p < 3; ++p)
        for (int q = 0; q < 3; ++q) pm[3 * p + q] += U[3 * a + p] * Y[3 * a + q];
    group_allsum(g, pm, red, out);
    double er[1] = {0.0};
    for (int a = g.tid; a < 10; a += g.nt)
      for (int q = 0; q < 3; ++q) {
        double d = Y[3 * a + q];
        for (int p = 0; p < 3; ++p) d -= U[3 * a + p] * pm[3 * p + q];
        er[0] += d * d;
      }
    group_allsum(g, er, red, out);
    const double err2 = er[0];
    g.sync();
    for (int idx = g.tid; idx < 30; idx += g.nt) U[idx] = Y[idx];
    g.sync();
    if (!(err2 == err2)) return false;
    if (err2 < 1e-29 || (err2 < 1e-24 && err2 > 0.81 * prev)) { ++it; break; }
    prev = err2;
    if (it == 47) return false;
  }
  if (iters_out) *iters_out = it;
  for (int idx = g.tid; idx < 30; idx += g.nt) Uout[idx] = U[idx];
  g.sync();
  return true;
}

// x_t = S^-1 r for the four vectors of the sweep, S = Ls Ls' (lower factor in a.S, reciprocal diagonal in a.rS); r and
// the result live in a.Z rows T0 .. T0 + n3.  Generic: one vector per thread, row-oriented substitution.
template <class G>
__host__ __device__ inline void arrow_schur_solve(const G& g, const ArrowWs& a, int n3) {
  const int T0 = 3 * RSX_S;
  for (int v = g.tid; v < ARW_K; v += g.nt) {
    double* z = a.Z + T0 * ARW_K + v;
    for (int r = 0; r < n3; ++r) {
      double s = z[r * ARW_K];
      for (int k = 0; k < r; ++k) s -= a.S[r * n3 + k] * z[k * ARW_K];
      z[r * ARW_K] = s * a.rS[r];
    }
    for (int r = n3 - 1; r >= 0; --r) {
      double s = z[r * ARW_K];
      for (int k = r + 1; k < n3; ++k) s -= a.S[k * n3 + r] * z[k * ARW_K];
      z[r * ARW_K] = s * a.rS[r];
    }
  }
  g.sync();
}
#ifdef __CUDA_ARCH__
// One warp: the row-oriented version keeps 4 lanes busy with two serial 3C x 3C substitutions — 28 % of all the
// instructions the warp fit kernel executed (ncu source page, r02).  Here all 32 lanes work column-oriented: lane
// (v = lane & 3, g8 = lane >> 2) keeps rows g8, g8 + 8, ... of vector v in registers; at step k the owner of row k scales
// it, the value travels to the 8 lanes of its vector by one shuffle and every lane updates the rows it still owns.
// 2 x 3C steps of ~14 instructions instead of 2 x (3C)^2 / 2 dependent shared-memory multiply-adds on 4 lanes.
template <int NT>
__device__ __forceinline__ void arrow_schur_solve_warp(const ArrowWs& a, int n3, int lane) {
  const int T0 = 3 * RSX_S;
  const int v = lane & 3, g8 = lane >> 2;
  double val[NT];
#pragma unroll
  for (int t = 0; t < NT; ++t) { const int r = g8 + 8 * t; val[t] = r < n3 ? a.Z[(T0 + r) * ARW_K + v] : 0.0; }
#pragma unroll 1
  for (int k = 0; k < n3; ++k) {                                   // forward: Ls y = r
    const int tk = k >> 3;
    double mine = 0.0;
#pragma unroll
    for (int t = 0; t < NT; ++t) mine = t == tk ? val[t] : mine;
    const double zk = __shfl_sync(0xffffffffu, mine * a.rS[k], ((k & 7) << 2) | v);
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int r = g8 + 8 * t;
      if (r == k) val[t] = zk;
      else if (r > k && r < n3) val[t] -= a.S[r * n3 + k] * zk;
    }
  }
#pragma unroll 1
  for (int k = n3 - 1; k >= 0; --k) {                              // backward: Ls' x = y
    const int tk = k >> 3;
    double mine = 0.0;
#pragma unroll
    for (int t = 0; t < NT; ++t) mine = t == tk ? val[t] : mine;
    const double zk = __shfl_sync(0xffffffffu, mine * a.rS[k], ((k & 7) << 2) | v);
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int r = g8 + 8 * t;
      if (r == k) val[t] = zk;
      else if (r < k) val[t] -= a.S[k * n3 + r] * zk;
    }
  }
#pragma unroll
  for (int t = 0; t < NT; ++t) { const int r = g8 + 8 * t; if (r < n3) a.Z[(T0 + r) * ARW_K + v] = val[t]; }
  __syncwarp();
}
__device__ __forceinline__ void arrow_schur_solve(const WarpCtx& g, const ArrowWs& a, int n3) {
  if (n3 <= 16) arrow_schur_solve_warp<2>(a, n3, g.tid);
  else if (n3 <= 40) arrow_schur_solve_warp<5>(a, n3, g.tid);
  else arrow_schur_solve<WarpCtx>(g, a, n3);
}
#endif

// H: [nu][ld] dense symmetric (read only), R: [C][9] rotations (for the start vectors), kb: [nu][4] basis out.
template <class G>
__host__ __device__ inline bool arrow_null4(const G& g, const double* H, int ld, int C, const double* R,
                                            const ArrowWs& a, double* kb, int* sweeps_out,
                                            int max_sweeps = ARW_MAX_SWEEPS) {
  const int n3 = 3 * C, T0 = 3 * RSX_S, nu = T0 + n3;
  // ---- shift: far below lambda_1 (>= 1e-7 of the mean diagonal here), far above the round-off of the factorisation
  double tr[1] = {0.0};
  for (int i = g.tid; i < nu; i += g.nt) tr[0] += H[i * ld + i];
  group_allsum(g, tr, a.red, a.out);
  const double sig = 1e-13 * tr[0] / nu;
  if (!(sig > 0)) return false;
  // ---- D_i + sig I = L_i L_i'
  for (int i = g.tid; i < RSX_S; i += g.nt) {
    const double* d = H + (3 * i) * ld + 3 * i;
    double* o = a.Ld + 9 * i;
    const double i0 = rsx_rsqrt_fast(d[0] + sig);
    const double l10 = d[ld] * i0, l20 = d[2 * ld] * i0;
    const double i1 = rsx_rsqrt_fast(d[ld + 1] + sig - l10 * l10);
    const double l21 = (d[2 * ld + 1] - l20 * l10) * i1;
    const double i2 = rsx_rsqrt_fast(d[2 * ld + 2] + sig - l20 * l20 - l21 * l21);
    o[0] = (d[0] + sig) * i0; o[1] = l10; o[2] = (d[ld + 1] + sig - l10 * l10) * i1; o[3] = l20; o[4] = l21;
    o[5] = (d[2 * ld + 2] + sig - l20 * l20 - l21 * l21) * i2; o[6] = i0; o[7] = i1; o[8] = i2;
  }
  g.sync();
  // ---- W_i = L_i^-1 B_i
  for (int t = g.tid; t < RSX_S * n3; t += g.nt) {
    const int i = t / n3, c = t - i * n3;
    const double* L = a.Ld + 9 * i;
    const double b0 = H[(3 * i) * ld + T0 + c], b1 = H[(3 * i + 1) * ld + T0 + c], b2 = H[(3 * i + 2) * ld + T0 + c];
    const double w0 = b0 * L[6];
    const double w1 = (b1 - L[1] * w0) * L[7];
    const double w2 = (b2 - L[3] * w0 - L[4] * w1) * L[8];
    a.Wm[(3 * i) * n3 + c] = w0; a.Wm[(3 * i + 1) * n3 + c] = w1; a.Wm[(3 * i + 2) * n3 + c] = w2;
  }
  g.sync();
  // ---- S = T + sig I - W' W (lower part), then its Cholesky factor in place
  for (int t = g.tid; t < n3 * n3; t += g.nt) {
    const int r = t / n3, c = t - r * n3;
    if (c > r) continue;
    // three independent partial sums: one running sum is a chain of 30 dependent FP64 multiply-adds, and this
    // latency-bound kernel (4 warps per scheduler, issue slots half empty) waits out every one of them
    double s0 = H[(T0 + r) * ld + T0 + c] + (r == c ? sig : 0.0), s1 = 0.0, s2 = 0.0;
#pragma unroll 2
    for (int k = 0; k < 3 * RSX_S; k += 3) {
      s0 -= a.Wm[k * n3 + r] * a.Wm[k * n3 + c];
      s1 -= a.Wm[(k + 1) * n3 + r] * a.Wm[(k + 1) * n3 + c];
      s2 -= a.Wm[(k + 2) * n3 + r] * a.Wm[(k + 2) * n3 + c];
    }
    a.S[r * n3 + c] = s0 + (s1 + s2);
  }
  g.sync();
  for (int k = 0; k < n3; ++k) {
    const double dk = a.S[k * n3 + k];
    if (!(dk > 0)) return false;                                   // same value in every thread
    const double ik = rsx_rsqrt_fast(dk);
    g.sync();
    for (int j = k + g.tid; j < n3; j += g.nt) {
      if (j == k) { a.S[k * n3 + k] = dk * ik; a.rS[k] = ik; }
      else a.S[j * n3 + k] *= ik;
    }
    g.sync();
    const int w = n3 - k - 1;
    for (int t = g.tid; t < w * w; t += g.nt) {
      const int r = k + 1 + t / w, c = k + 1 + t % w;
      if (c <= r) a.S[r * n3 + c] -= a.S[r * n3 + k] * a.S[c * n3 + k];
    }
    g.sync();
  }
  // ---- start: the three world shifts (exact null vectors of the collinearity rows: X_i += d, t_c -= R_c d) and one
  // fixed pseudo-random vector for the scale-like direction
  for (int i = g.tid; i < nu; i += g.nt) {
    double* q = a.Q + i * ARW_K;
    if (i < T0) { q[0] = (i % 3 == 0); q[1] = (i % 3 == 1); q[2] = (i % 3 == 2); }
    else { const int c = (i - T0) / 3, r = (i - T0) % 3; q[0] = -R[9 * c + 3 * r]; q[1] = -R[9 * c + 3 * r + 1]; q[2] = -R[9 * c + 3 * r + 2]; }
    const unsigned hsh = (unsigned)(i + 1) * 2654435761u;
    q[3] = (double)(hsh >> 8) * (1.0 / 16777216.0) - 0.5;
  }
  g.sync();
  if (!cholqr4(g, a.Q, nu, a.red, a.out)) return false;
  if (!cholqr4(g, a.Q, nu, a.red, a.out)) return false;
  double prev = 1e300;
  bool done = false;
  int sweep = 0;
  for (; sweep < max_sweeps && !done; ++sweep) {
    // ---- Z = (H + sig I)^-1 Q by block elimination
    for (int t = g.tid; t < RSX_S * ARW_K; t += g.nt) {            // y_i = L_i^-1 q_i
      const int i = t / ARW_K, v = t - i * ARW_K;
      const double* L = a.Ld + 9 * i;
      const double y0 = a.Q[(3 * i) * ARW_K + v] * L[6];
      const double y1 = (a.Q[(3 * i + 1) * ARW_K + v] - L[1] * y0) * L[7];
      const double y2 = (a.Q[(3 * i + 2) * ARW_K + v] - L[3] * y0 - L[4] * y1) * L[8];
      a.Z[(3 * i) * ARW_K + v] = y0; a.Z[(3 * i + 1) * ARW_K + v] = y1; a.Z[(3 * i + 2) * ARW_K + v] = y2;
    }
    g.sync();
    for (int t = g.tid; t < n3 * ARW_K; t += g.nt) {               // r = q_t - W' y
      const int r = t / ARW_K, v = t - r * ARW_K;
      double s0 = a.Q[(T0 + r) * ARW_K + v], s1 = 0.0, s2 = 0.0;      // three chains of 10, not one of 30 (see above)
#pragma unroll 2
      for (int k = 0; k < 3 * RSX_S; k += 3) {
        s0 -= a.Wm[k * n3 + r] * a.Z[k * ARW_K + v];
        s1 -= a.Wm[(k + 1) * n3 + r] * a.Z[(k + 1) * ARW_K + v];
        s2 -= a.Wm[(k + 2) * n3 + r] * a.Z[(k + 2) * ARW_K + v];
      }
      a.Z[(T0 + r) * ARW_K + v] = s0 + (s1 + s2);
    }
    g.sync();
    arrow_schur_solve(g, a, n3);                                   // x_t = S^-1 r (syncs inside)
    for (int t = g.tid; t < RSX_S * ARW_K; t += g.nt) {            // x_i = L_i^-T (y_i - W_i x_t)
      const int i = t / ARW_K, v = t - i * ARW_K;
      const double* L = a.Ld + 9 * i;
      double y0 = a.Z[(3 * i) * ARW_K + v], y1 = a.Z[(3 * i + 1) * ARW_K + v], y2 = a.Z[(3 * i + 2) * ARW_K + v];
      double u0 = 0.0, u1 = 0.0, u2 = 0.0;                          // second set of chains (n3 = 3C is a multiple of 3, odd or even)
      int c = 0;
      for (; c + 1 < n3; c += 2) {
        const double xt = a.Z[(T0 + c) * ARW_K + v], xu = a.Z[(T0 + c + 1) * ARW_K + v];
        y0 -= a.Wm[(3 * i) * n3 + c] * xt; y1 -= a.Wm[(3 * i + 1) * n3 + c] * xt; y2 -= a.Wm[(3 * i + 2) * n3 + c] * xt;
        u0 -= a.Wm[(3 * i) * n3 + c + 1] * xu; u1 -= a.Wm[(3 * i + 1) * n3 + c + 1] * xu; u2 -= a.Wm[(3 * i + 2) * n3 + c + 1] * xu;
      }
      if (c < n3) {
        const double xt = a.Z[(T0 + c) * ARW_K + v];
        y0 -= a.Wm[(3 * i) * n3 + c] * xt; y1 -= a.Wm[(3 * i + 1) * n3 + c] * xt; y2 -= a.Wm[(3 * i + 2) * n3 + c] * xt;
      }
      y0 += u0; y1 += u1; y2 += u2;
      const double x2 = y2 * L[8];
      const double x1 = (y1 - L[4] * x2) * L[7];
      const double x0 = (y0 - L[1] * x1 - L[3] * x2) * L[6];
      a.Z[(3 * i) * ARW_K + v] = x0; a.Z[(3 * i + 1) * ARW_K + v] = x1; a.Z[(3 * i + 2) * ARW_K + v] = x2;
    }
    g.sync();
    if (!cholqr4(g, a.Z, nu, a.red, a.out)) return false;
    if (!cholqr4(g, a.Z, nu, a.red, a.out)) return false;
    // ---- change of the subspace: || Z - Q (Q' Z) ||_F^2
    double gm[16];
    for (int e = 0; e < 16; ++e) gm[e] = 0.0;
    for (int i = g.tid; i < nu; i += g.nt)
      for (int p = 0; p < ARW_K; ++p)
        for (int q = 0; q < ARW_K; ++q) gm[4 * p + q] += a.Q[i * ARW_K + p] * a.Z[i * ARW_K + q];
    group_allsum(g, gm, a.red, a.out);
    double er[1] = {0.0};
    for (int i = g.tid; i < nu; i += g.nt)
      for (int q = 0; q < ARW_K; ++q) {
        double d = a.Z[i * ARW_K + q];
        for (int p = 0; p < ARW_K; ++p) d -= a.Q[i * ARW_K + p] * gm[4 * p + q];
        er[0] += d * d;
      }
    group_allsum(g, er, a.red, a.out);
    const double err2 = er[0];
    for (int i = g.tid; i < nu * ARW_K; i += g.nt) a.Q[i] = a.Z[i];
    g.sync();
    // settled: the subspace moved by less than 1e-12 (the error left is that times rho / (1 - rho)), or it sits on
    // the round-off floor of the solves (tiny and no longer contracting)
#if defined(ARW_DEBUG) && !defined(__CUDA_ARCH__)
    printf("  sweep %d err2 %.3e\n", sweep, err2);
#endif
    // what is LEFT after a movement of sqrt(err2) is that times rho / (1 - rho), rho^2 ~ err2 / prev the observed
    // contraction: a slowly converging run (rho ~ 0.8, possible now that up to 160 sweeps are allowed) must move less
    const double rho2 = err2 < prev ? err2 / prev : 1.0, omr = 1.0 - sqrt(rho2);
    const double amp2 = rho2 > 0.25 ? rho2 / (omr * omr) : 1.0;
    // A run still moving after ARW_LATE sweeps has a poor gap: its sample is ill-conditioned and amplifies what is left
    // of the subspace error by orders of magnitude in the score (1e-12 -> 2e-8 seen), so it must settle 100x tighter
    // (or sit on the round-off floor).  With that, letting slow runs go on (200 sweeps where a second pass exists)
    // keeps the scores within 1e-9 of the Jacobi result and cuts the hypotheses left to that pass from 3.5 % to 0.4 %.
    done = (err2 * amp2 < (sweep < ARW_LATE ? 1e-24 : ARW_LATE_TOL)) || (err2 < 1e-21 && err2 > 0.81 * prev);
    if (!(err2 == err2)) return false;                             // NaN
    prev = err2;
  }
  if (sweeps_out) *sweeps_out = sweep;
  if (!done) return false;
  for (int i = g.tid; i < nu * ARW_K; i += g.nt) kb[i] = a.Q[i];
  g.sync();
  return true;
}
