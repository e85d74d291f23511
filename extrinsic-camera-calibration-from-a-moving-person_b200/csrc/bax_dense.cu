// Reduced camera system: FP64 single-CTA blocked Cholesky + triangular solves (SPD because alpha > 0).
#include "bax_internal.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

#define DN_THREADS 1024
#define NBK 32
#define LDS_ 33

// In-place lower Cholesky of A (n x n, row-major, leading dimension ld, global/L2 resident).
// Dg: [32][33] shared, Pn: [(n-32)][33] shared.  Right-looking, 32-wide panels, 4x4 register tiles.
__device__ __noinline__ void chol32_warp(double* Dg, double* rinv);
__device__ void chol_blocked(double* __restrict__ A, int n, int ld, double* Dg, double* Pn) {
  __shared__ double rinv_s[NBK];
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int j0 = 0; j0 < n; j0 += NBK) {
    const int nb = min(NBK, n - j0);
    for (int idx = tid; idx < NBK * NBK; idx += nt) {      // lower part; identity beyond nb (chol32_warp works on 32 x 32)
      const int r = idx >> 5, c = idx & 31;
      Dg[r * LDS_ + c] = (r < nb && c <= r) ? A[(size_t)(j0 + r) * ld + j0 + c] : (r == c ? 1.0 : 0.0);
    }
    __syncthreads();
    chol32_warp(Dg, rinv_s);                               // ends with a CTA barrier
    for (int idx = tid; idx < nb * nb; idx += nt) {
      int r = idx / nb, c = idx % nb;
      if (c <= r) A[(size_t)(j0 + r) * ld + j0 + c] = Dg[r * LDS_ + c];
    }
    const int base = j0 + nb, m = n - base;
    if (m > 0) {
      for (int idx = tid; idx < m * nb; idx += nt) {
        int r = idx / nb, c = idx % nb;
        Pn[r * LDS_ + c] = A[(size_t)(base + r) * ld + j0 + c];
      }
      __syncthreads();
      for (int r = tid; r < m; r += nt) {        // L21 = A21 L11^-T, one row per thread (1 / L_tt from the factorisation,
        double* row = Pn + r * LDS_;             // as in the cooperative kernel's trsm32_cta: no FP64 division per entry)
        for (int t = 0; t < nb; ++t) {
          double v = row[t];
          for (int u = 0; u < t; ++u) v -= row[u] * Dg[t * LDS_ + u];
          row[t] = v * rinv_s[t];
        }
      }
      __syncthreads();
      for (int idx = tid; idx < m * nb; idx += nt) {
        int r = idx / nb, c = idx % nb;
        A[(size_t)(base + r) * ld + j0 + c] = Pn[r * LDS_ + c];
      }
      const int m4 = (m + 3) >> 2;
      for (int tile = tid; tile < m4 * m4; tile += nt) {
        int ti = tile / m4, tk = tile % m4;
        if (tk > ti) continue;
        double acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
        int ra[4], rk[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { ra[a] = min(4 * ti + a, m - 1); rk[a] = min(4 * tk + a, m - 1); }
        for (int t = 0; t < nb; ++t) {
          double pa[4], pk[4];
#pragma unroll
          for (int a = 0; a < 4; ++a) { pa[a] = Pn[ra[a] * LDS_ + t]; pk[a] = Pn[rk[a] * LDS_ + t]; }
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] += pa[a] * pk[b];
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            int i = 4 * ti + a, k = 4 * tk + b;
            if (i < m && k <= i) A[(size_t)(base + i) * ld + base + k] -= acc[a][b];
          }
      }
    }
    __syncthreads();
  }
}

// x <- L^-1 x (fwd) or L^-T x (bwd); x in shared memory, L lower in global.
__device__ void trsv_blocked(const double* __restrict__ A, int n, int ld, double* x, double* Dg, bool transpose) {
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  const int nblk = (n + NBK - 1) / NBK;
  for (int bi = 0; bi < nblk; ++bi) {
    const int j0 = transpose ? (nblk - 1 - bi) * NBK : bi * NBK;
    const int nb = min(NBK, n - j0);
    for (int idx = tid; idx < nb * nb; idx += nt) {
      int r = idx / nb, c = idx % nb;
      Dg[r * LDS_ + c] = A[(size_t)(j0 + r) * ld + j0 + c];
    }
    __syncthreads();
    if (warp == 0) {
      // lane l carries the running right-hand side of unknown l; one broadcast + one shared-memory read per column
      double acc = lane < nb ? x[j0 + lane] : 0.0;
      const double rinv = lane < nb ? 1.0 / Dg[lane * LDS_ + lane] : 0.0;
      double mine = 0.0;
      // selects instead of branches, fixed trip count (steps k >= nb see acc = rinv = 0 and change nothing)
      if (!transpose) {
#pragma unroll
        for (int k = 0; k < NBK; ++k) {
          const double xk = __shfl_sync(0xffffffffu, acc, k) * __shfl_sync(0xffffffffu, rinv, k);
          mine = lane == k ? xk : mine;
          const double l = (lane > k && lane < nb) ? Dg[lane * LDS_ + k] : 0.0;
          acc -= l * xk;
        }
      } else {
#pragma unroll
        for (int k = NBK - 1; k >= 0; --k) {
          const double xk = __shfl_sync(0xffffffffu, acc, k) * __shfl_sync(0xffffffffu, rinv, k);
          mine = lane == k ? xk : mine;
          const double l = (lane < k && k < nb) ? Dg[k * LDS_ + lane] : 0.0;
          acc -= l * xk;
        }
      }
      if (lane < nb) x[j0 + lane] = mine;
    }
    __syncthreads();
    // the NBK loads of an unknown are issued together (a run-time trip count left them one L2 round trip each:
    // 166k of the cooperative solve's 714k cycles at 384 unknowns went into this loop)
    if (!transpose) {
      for (int i = j0 + nb + tid; i < n; i += nt) {
        const double* row = A + (size_t)i * ld + j0;
        double a[NBK];
#pragma unroll
        for (int t = 0; t < NBK; ++t) a[t] = t < nb ? row[t] : 0.0;
        double v = x[i];
#pragma unroll
        for (int t = 0; t < NBK; ++t) if (t < nb) v -= a[t] * x[j0 + t];
        x[i] = v;
      }
    } else {
      for (int i = tid; i < j0; i += nt) {
        double a[NBK];
#pragma unroll
        for (int t = 0; t < NBK; ++t) a[t] = t < nb ? A[(size_t)(j0 + t) * ld + i] : 0.0;
        double v = x[i];
#pragma unroll
        for (int t = 0; t < NBK; ++t) if (t < nb) v -= a[t] * x[j0 + t];
        x[i] = v;
      }
    }
    __syncthreads();
  }
}

// workspace layout (doubles): Sf [n*n] | Z [B*n] | pc [n] | qc [n] | LT [B*B] | hz [B] | w [B]
struct DenseWs { double *Sf, *Z, *pc, *qc, *LT, *hz, *w; };
__device__ __host__ inline DenseWs dense_ws(double* ws, int n, int B) {
  DenseWs d; d.Sf = ws; d.Z = d.Sf + (size_t)n * n; d.pc = d.Z + (size_t)BAX_MAX_B * n; d.qc = d.pc + n;
  d.LT = d.qc + n; d.hz = d.LT + BAX_MAX_B * BAX_MAX_B; d.w = d.hz + BAX_MAX_B; return d;
}

// small serial helpers on B x B (B <= 16) systems held in shared memory
__device__ void small_chol(double* T, int B) {     // in place lower, row-major ld = BAX_MAX_B
  for (int k = 0; k < B; ++k) {
    double d = sqrt(T[k * BAX_MAX_B + k]); T[k * BAX_MAX_B + k] = d;
    const double rd = 1.0 / d;                       // one division per column (FP64 divide latency dominates here)
    for (int i = k + 1; i < B; ++i) T[i * BAX_MAX_B + k] *= rd;
    for (int i = k + 1; i < B; ++i)
      for (int j = k + 1; j <= i; ++j) T[i * BAX_MAX_B + j] -= T[i * BAX_MAX_B + k] * T[j * BAX_MAX_B + k];
  }
}
// The same factorisation through the warp routine (T padded with the identity into a 32 x 32 block): one thread
// working through shared memory took 45k cycles for B = 12 — a third of the whole reduced solve of a 4-camera problem.
// Leaves 1 / L_aa in rinvT[BAX_MAX_B].  Ends with a CTA barrier.
__device__ void small_chol_warp(double* T, int B, double* Dg, double* rinv32, double* rinvT) {
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int idx = tid; idx < NBK * NBK; idx += nt) {
    const int r = idx >> 5, c = idx & 31;
    Dg[r * LDS_ + c] = (r < B && c <= r) ? T[r * BAX_MAX_B + c] : (r == c ? 1.0 : 0.0);
  }
  __syncthreads();
  chol32_warp(Dg, rinv32);
  for (int idx = tid; idx < BAX_MAX_B * BAX_MAX_B; idx += nt) {
    const int r = idx / BAX_MAX_B, c = idx % BAX_MAX_B;
    if (r < B && c <= r) T[idx] = Dg[r * LDS_ + c];
  }
  if (tid < BAX_MAX_B) rinvT[tid] = tid < B ? rinv32[tid] : 1.0;
  __syncthreads();
}
__device__ void small_fwd_r(const double* L, const double* ri, int B, double* v) {       // reciprocals instead of divisions
  for (int a = 0; a < B; ++a) { double s = v[a]; for (int b = 0; b < a; ++b) s -= L[a * BAX_MAX_B + b] * v[b]; v[a] = s * ri[a]; }
}
__device__ void small_bwd_r(const double* L, const double* ri, int B, double* v) {
  for (int a = B - 1; a >= 0; --a) { double s = v[a]; for (int b = a + 1; b < B; ++b) s -= L[b * BAX_MAX_B + a] * v[b]; v[a] = s * ri[a]; }
}
__device__ void small_fwd(const double* L, int B, double* v) {
  for (int a = 0; a < B; ++a) { double s = v[a]; for (int b = 0; b < a; ++b) s -= L[a * BAX_MAX_B + b] * v[b]; v[a] = s / L[a * BAX_MAX_B + a]; }
}
__device__ void small_bwd(const double* L, int B, double* v) {
  for (int a = B - 1; a >= 0; --a) { double s = v[a]; for (int b = a + 1; b < B; ++b) s -= L[b * BAX_MAX_B + a] * v[b]; v[a] = s / L[a * BAX_MAX_B + a]; }
}

// NT = 1024 caps the kernel at 64 registers, which spills the warp factorisation's 32-double row; small systems (little
// to share out anyway) run the NT = 256 instance.
template <int NT>
__global__ void __launch_bounds__(NT)
k_dense_solve1(const double* __restrict__ schur, int R, const double* __restrict__ lin, int n, int B, double alpha,
               double* __restrict__ ws, double* __restrict__ coef, double* __restrict__ step /*camera part, theta order*/) {
  extern __shared__ double smem[];
  double* Dg = smem;                        // 32*33
  double* xs = Dg + NBK * LDS_;             // n
  double* sLT = xs + ((n + 1) & ~1);        // 16*16
  double* sh = sLT + BAX_MAX_B * BAX_MAX_B; // 16 (h0 / hz)
  double* Pn = sh + BAX_MAX_B;              // (n-32)*33
  const int tid = threadIdx.x, nt = blockDim.x;
  DenseWs d = dense_ws(ws, n, B);
  const double* Hcc = lin; const double* gc = lin + (size_t)n * n;
  // 1. T = I + Ut^T Ut, Cholesky; hz = LT^-1 h0
  for (int i = tid; i < BAX_MAX_B * BAX_MAX_B; i += nt) {
    int a = i / BAX_MAX_B, b = i % BAX_MAX_B;
    sLT[i] = (a < B && b < B) ? schur[(size_t)(n + a) * R + n + b] + (a == b ? 1.0 : 0.0) : (a == b ? 1.0 : 0.0);
  }
  if (tid < BAX_MAX_B) sh[tid] = tid < B ? schur[(size_t)(n + tid) * R + n + B] : 0.0;
  __syncthreads();
  __shared__ double s_rinv32[NBK], s_trinv[BAX_MAX_B];
  small_chol_warp(sLT, B, Dg, s_rinv32, s_trinv);
  if (tid == 0) small_fwd_r(sLT, s_trinv, B, sh);
  __syncthreads();
  for (int i = tid; i < BAX_MAX_B * BAX_MAX_B; i += nt) d.LT[i] = sLT[i];
  if (tid < BAX_MAX_B) d.hz[tid] = sh[tid];
  // 2. Z = LT^-1 G^T  (B x n), one camera parameter per thread
  for (int i = tid; i < n; i += nt) {
    double z[BAX_MAX_B];
#pragma unroll
    for (int a = 0; a < BAX_MAX_B; ++a) z[a] = a < B ? schur[(size_t)i * R + n + a] : 0.0;
#pragma unroll
    for (int a = 0; a < BAX_MAX_B; ++a) {
      double s = z[a];
#pragma unroll
      for (int b = 0; b < BAX_MAX_B; ++b) if (b < a) s -= sLT[a * BAX_MAX_B + b] * z[b];
      z[a] = s * s_trinv[a];
    }
#pragma unroll
    for (int a = 0; a < BAX_MAX_B; ++a) if (a < B) d.Z[(size_t)a * n + i] = z[a];
    double r = -gc[i] - schur[(size_t)i * R + n + B];
#pragma unroll
    for (int a = 0; a < BAX_MAX_B; ++a) if (a < B) r += z[a] * sh[a];
    xs[i] = r;
  }
  __syncthreads();
  // 3. S = Hcc + alpha I - Y Y^T + Z^T Z   (lower triangle)
  for (int idx = tid; idx < n * n; idx += nt) {
    int i = idx / n, j = idx % n;
    if (j > i) continue;
    double v = Hcc[idx] - schur[(size_t)i * R + j] + (i == j ? alpha : 0.0);
    for (int a = 0; a < B; ++a) v += d.Z[(size_t)a * n + i] * d.Z[(size_t)a * n + j];
    d.Sf[idx] = v;
  }
  __syncthreads();
  // 4. factor and solve
  chol_blocked(d.Sf, n, n, Dg, Pn);
  trsv_blocked(d.Sf, n, n, xs, Dg, false);
  trsv_blocked(d.Sf, n, n, xs, Dg, true);
  for (int i = tid; i < n; i += nt) {           // internal order 6c + {rvec, t} -> theta order (all rvecs, then all tvecs)
    d.pc[i] = xs[i]; coef[i] = xs[i];
    const int c = i / 6, a = i % 6, C = n / 6;
    step[a < 3 ? 3 * c + a : 3 * C + 3 * c + (a - 3)] = xs[i];
  }
  // 5. w = T^-1 (h0 - G^T pc)
  if (tid < B) {
    double hv = schur[(size_t)(n + tid) * R + n + B];
    for (int i = 0; i < n; ++i) hv -= schur[(size_t)i * R + n + tid] * xs[i];
    sh[tid] = hv;
  }
  __syncthreads();
  if (tid == 0) { small_fwd_r(sLT, s_trinv, B, sh); small_bwd_r(sLT, s_trinv, B, sh); }
  __syncthreads();
  if (tid < B) { d.w[tid] = sh[tid]; coef[n + tid] = sh[tid]; }
}

// ------------------------------------------------------------------ cooperative multi-CTA version of solve 1
// The single-CTA kernel above spends ~1.3 ms on a 384 x 384 system (one SM, barrier-bound).  Here G CTAs share the
// work and meet at grid barriers (cooperative launch): Z and the right-hand side, the assembly of S, then a
// right-looking blocked Cholesky in which EVERY CTA factors the 32 x 32 diagonal block redundantly in shared memory
// (no broadcast step), solves its share of the panel rows, and after one barrier updates its share of the 32 x 32
// trailing tiles.  CTA 0 finishes with the two triangular solves.  Same arithmetic per entry as the single-CTA path
// except for the summation order inside the trailing updates.
#define CG_THREADS 256
// 32 x 32 blocks are worked on by 256 threads: thread (r = tid >> 3, cg = tid & 7) keeps elements (r, cg + 8 q),
// q < 4, in registers; a column travels through a 32-entry shared buffer.
__device__ __forceinline__ double sel4(const double (&e)[4], int q) {
  return q == 0 ? e[0] : q == 1 ? e[1] : q == 2 ? e[2] : e[3];
}
// Factor the block held in Dg (row-major, stride LDS_; rows / columns >= nb padded with the identity) in place and
// leave 1 / L_kk in rinv[32].  Per column: publish it, barrier, everyone scales and updates its own elements.
__device__ void chol32_cta(double* Dg, double* rinv, double* colbuf /*[2][32]*/) {
  const int tid = threadIdx.x, r = tid >> 3, cg = tid & 7;
  double e[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) e[q] = Dg[r * LDS_ + cg + 8 * q];
  for (int k = 0; k < NBK; ++k) {
    double* cb = colbuf + (k & 1) * NBK;           // double-buffered: one barrier per column
    if (cg == (k & 7)) cb[r] = sel4(e, k >> 3);
    __syncthreads();
    const double ri = rsqrt(cb[k]);
    const double dkk = cb[k] * ri;
    const double lr = cb[r] * ri;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int c = cg + 8 * q;
      if (c == k) e[q] = r == k ? dkk : lr;        // final column k (rows < k: unused upper part)
      else if (c > k && r >= c) e[q] = fma(-lr, cb[c] * ri, e[q]);
    }
    if (tid == 0) rinv[k] = ri;
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) Dg[r * LDS_ + cg + 8 * q] = e[q];
  __syncthreads();
}
// The same factorisation by ONE warp out of registers: lane r keeps row r (32 doubles), the pivot and the column
// entries travel by shuffle, no barrier and no shared-memory round trip per column (the 256-thread version above spends
// ~720 cycles per column on them: 23k cycles per block, a third of the cooperative solve at 384 unknowns).  Same
// arithmetic per entry as chol32_cta (bitwise).  Ends with a CTA barrier.
// (noinline: one copy per kernel.  The fully unrolled body is ~50 KB of code; a second inlined copy — the B x B system of
// the bone-length rows goes through it too — slowed every call from 8.6k to 11.2k cycles.  A rolled column loop that
// shifts the row by one register per column keeps the code at 4 KB but issues twice the shuffles: 15.3k cycles.)
__device__ __noinline__ void chol32_warp(double* Dg, double* rinv) {
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    double e[NBK];
#pragma unroll
    for (int c = 0; c < NBK; ++c) e[c] = Dg[lane * LDS_ + c];
#pragma unroll
    for (int k = 0; k < NBK; ++k) {
      const double piv = __shfl_sync(0xffffffffu, e[k], k);
      const double ri = rsqrt(piv);
      const double lr = e[k] * ri;                 // L[lane][k] (lanes < k: unused upper part)
      e[k] = lr;
      if (lane == k) rinv[k] = ri;
#pragma unroll
      for (int c = k + 1; c < NBK; ++c) {
        const double lc = __shfl_sync(0xffffffffu, lr, c);       // L[c][k]
        e[c] = fma(-lr, lc, e[c]);
      }
    }
#pragma unroll
    for (int c = 0; c < NBK; ++c) Dg[lane * LDS_ + c] = e[c];
  }
  __syncthreads();
}
// X <- X L^-T for 32 rows held in Pn (row-major, stride LDS_), L = Dg with reciprocal diagonal rinv
__device__ void trsm32_cta(double* Pn, const double* Dg, const double* rinv) {
  const int tid = threadIdx.x, r = tid >> 3, cg = tid & 7, lane = tid & 31;
  double e[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) e[q] = Pn[r * LDS_ + cg + 8 * q];
  // Branch-free and fully unrolled: written with `if (t == u) .. else if (t > u) ..` and a run-time u, the eight column
  // owners of a warp took different paths through four reconvergence regions per step and a step cost ~400 cycles
  // (172k of the cooperative solve's 714k cycles at 384 unknowns); as selects it is one shuffle and four FMAs.
#pragma unroll
  for (int u = 0; u < NBK; ++u) {
    double x = e[u >> 3] * rinv[u];
    x = __shfl_sync(0xffffffffu, x, (lane & ~7) | (u & 7));      // from the owner of column u in this row
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int t = cg + 8 * q;
      const double upd = fma(-x, Dg[t * LDS_ + u], e[q]);
      e[q] = t == u ? x : (t > u ? upd : e[q]);
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) Pn[r * LDS_ + cg + 8 * q] = e[q];
}

__global__ void __launch_bounds__(CG_THREADS)
k_dense_solve1_coop(const double* __restrict__ schur, int R, const double* __restrict__ lin, int n, int B, double alpha,
                    double* __restrict__ ws, double* __restrict__ coef, int prof, int chol_warp, double* __restrict__ step) {
  long long tp[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; long long tc = clock64();
#define PROF(i) { long long t_ = clock64(); tp[i] += t_ - tc; tc = t_; }
  cg::grid_group grid = cg::this_grid();
  extern __shared__ double smem[];
  double* Dg = smem;                        // 32*33
  double* Ta = Dg + NBK * LDS_;             // 32*33
  double* Tb = Ta + NBK * LDS_;             // 32*33
  double* sLT = Tb + NBK * LDS_;            // 16*16
  double* sh = sLT + BAX_MAX_B * BAX_MAX_B; // 16
  double* rinv = sh + BAX_MAX_B;            // 32
  double* colbuf = rinv + NBK;              // 2*32
  double* Pn = colbuf + 2 * NBK;            // 32*33 panel rows in flight
  double* xs = Pn + NBK * LDS_;             // n (CTA 0 only)
  const int tid = threadIdx.x, nt = blockDim.x, G = gridDim.x, cta = blockIdx.x;
  const int gtid = cta * nt + tid, gnt = G * nt;
  DenseWs d = dense_ws(ws, n, B);
  const double* Hcc = lin; const double* gc = lin + (size_t)n * n;
  // 1. T = I + Ut^T Ut, Cholesky; hz = LT^-1 h0   (every CTA, redundantly: B <= 15)
  for (int i = tid; i < BAX_MAX_B * BAX_MAX_B; i += nt) {
    int a = i / BAX_MAX_B, b = i % BAX_MAX_B;
    sLT[i] = (a < B && b < B) ? schur[(size_t)(n + a) * R + n + b] + (a == b ? 1.0 : 0.0) : (a == b ? 1.0 : 0.0);
  }
  if (tid < BAX_MAX_B) sh[tid] = tid < B ? schur[(size_t)(n + tid) * R + n + B] : 0.0;
  __syncthreads();
  __shared__ double s_trinv[BAX_MAX_B];
  small_chol_warp(sLT, B, Dg, rinv, s_trinv);
  if (tid == 0) small_fwd_r(sLT, s_trinv, B, sh);
  __syncthreads();
  if (cta == 0) {
    for (int i = tid; i < BAX_MAX_B * BAX_MAX_B; i += nt) d.LT[i] = sLT[i];
    if (tid < BAX_MAX_B) d.hz[tid] = sh[tid];
  }
  // 2. Z = LT^-1 G^T (B x n) and the right-hand side, one camera parameter per thread of the grid
  for (int i = gtid; i < n; i += gnt) {
    double z[BAX_MAX_B];
#pragma unroll
    for (int a = 0; a < BAX_MAX_B; ++a) z[a] = a < B ? schur[(size_t)i * R + n + a] : 0.0;
#pragma unroll
    for (int a = 0; a < BAX_MAX_B; ++a) {
      double v = z[a];
#pragma unroll
      for (int b = 0; b < BAX_MAX_B; ++b) if (b < a) v -= sLT[a * BAX_MAX_B + b] * z[b];
      z[a] = v * s_trinv[a];
    }
#pragma unroll
    for (int a = 0; a < BAX_MAX_B; ++a) if (a < B) d.Z[(size_t)a * n + i] = z[a];
    double r = -gc[i] - schur[(size_t)i * R + n + B];
#pragma unroll
    for (int a = 0; a < BAX_MAX_B; ++a) if (a < B) r += z[a] * sh[a];
    d.qc[i] = r;                            // right-hand side parked in qc until CTA 0 picks it up
  }
  PROF(0)
  grid.sync();
  PROF(1)
  // 3. S = Hcc + alpha I - Y Y^T + Z^T Z   (lower triangle)
  for (int idx = gtid; idx < n * n; idx += gnt) {
    int i = idx / n, j = idx % n;
    if (j > i) continue;
    double v = Hcc[idx] - schur[(size_t)i * R + j] + (i == j ? alpha : 0.0);
    for (int a = 0; a < B; ++a) v += d.Z[(size_t)a * n + i] * d.Z[(size_t)a * n + j];
    d.Sf[idx] = v;
  }
  PROF(2)
  grid.sync();
  PROF(1)
  // 4. blocked Cholesky of Sf
  double* A = d.Sf;
  for (int j0 = 0; j0 < n; j0 += NBK) {
    const int nb = min(NBK, n - j0);
    for (int idx = tid; idx < NBK * NBK; idx += nt) {
      const int r = idx >> 5, c = idx & 31;
      Dg[r * LDS_ + c] = (r < nb && c <= r) ? A[(size_t)(j0 + r) * n + j0 + c] : (r == c ? 1.0 : 0.0);
    }
    __syncthreads();
    PROF(3)
    if (chol_warp) chol32_warp(Dg, rinv); else chol32_cta(Dg, rinv, colbuf);   // both end with a CTA barrier
    PROF(4)
    if (cta == 0)
      for (int idx = tid; idx < nb * nb; idx += nt) {
        int r = idx / nb, c = idx % nb;
        if (c <= r) A[(size_t)(j0 + r) * n + j0 + c] = Dg[r * LDS_ + c];
      }
    const int base = j0 + nb, m = n - base;
    // panel rows: L21 = A21 L11^-T, one thread per row; the right-hand side rides along as row m (so the forward
    // substitution y = L^-1 rhs is a by-product of the factorisation)
    for (int rb = cta * NBK; rb <= m; rb += G * NBK) {
      const int nr = min(NBK, m + 1 - rb);                      // rows rb .. rb+nr-1 (row m = right-hand side)
      __syncthreads();
      for (int idx = tid; idx < NBK * NBK; idx += nt) {
        const int r = idx >> 5, c = idx & 31;
        const double* src = rb + r < m ? A + (size_t)(base + rb + r) * n + j0 : d.qc + j0;
        Pn[r * LDS_ + c] = (r < nr && c < nb) ? src[c] : 0.0;
      }
      __syncthreads();
      PROF(8)
      trsm32_cta(Pn, Dg, rinv);
      __syncthreads();
      PROF(9)
      for (int idx = tid; idx < nr * NBK; idx += nt) {
        const int r = idx >> 5, c = idx & 31;
        double* dst = rb + r < m ? A + (size_t)(base + rb + r) * n + j0 : d.qc + j0;
        if (c < nb) dst[c] = Pn[r * LDS_ + c];
      }
    }
    PROF(5)
    if (m <= 0) break;
    grid.sync();
    PROF(1)
    // trailing update, 32 x 32 tiles of the lower triangle dealt round-robin
    const int mb = (m + NBK - 1) / NBK, ntile = mb * (mb + 1) / 2;
    for (int tile = cta; tile < ntile; tile += G) {
      int ti = (int)((sqrt(8.0 * tile + 1.0) - 1.0) * 0.5);
      while ((ti + 1) * (ti + 2) / 2 <= tile) ++ti;
      while (ti * (ti + 1) / 2 > tile) --ti;
      const int tk = tile - ti * (ti + 1) / 2;
      __syncthreads();
      for (int idx = tid; idx < NBK * NBK; idx += nt) {
        const int r = idx / NBK, c = idx % NBK;
        const int ri = ti * NBK + r, rk = tk * NBK + r;
        Ta[r * LDS_ + c] = (ri < m && c < nb) ? A[(size_t)(base + ri) * n + j0 + c] : 0.0;
        Tb[r * LDS_ + c] = (rk < m && c < nb) ? A[(size_t)(base + rk) * n + j0 + c] : 0.0;
      }
      // 256 threads x (2 x 2) outputs: rows {ty, ty+16}, cols {tx, tx+16}
      const int tx = tid & 15, ty = tid >> 4;
      const int i0 = ti * NBK + ty, i1 = i0 + 16, k0 = tk * NBK + tx, k1 = k0 + 16;
      const bool w00 = i0 < m && k0 <= i0, w01 = i0 < m && k1 <= i0 && k1 < m, w10 = i1 < m && k0 <= i1,
                 w11 = i1 < m && k1 <= i1 && k1 < m;
      double* o00 = A + (size_t)(base + i0) * n + base + k0; double* o01 = o00 + 16;
      double* o10 = A + (size_t)(base + i1) * n + base + k0; double* o11 = o10 + 16;
      double a00 = w00 ? *o00 : 0.0, a01 = w01 ? *o01 : 0.0, a10 = w10 ? *o10 : 0.0, a11 = w11 ? *o11 : 0.0;
      __syncthreads();
#pragma unroll 8
      for (int t = 0; t < NBK; ++t) {
        const double p0 = Ta[ty * LDS_ + t], p1 = Ta[(ty + 16) * LDS_ + t];
        const double q0 = Tb[tx * LDS_ + t], q1 = Tb[(tx + 16) * LDS_ + t];
        a00 = fma(-p0, q0, a00); a01 = fma(-p0, q1, a01); a10 = fma(-p1, q0, a10); a11 = fma(-p1, q1, a11);
      }
      if (w00) *o00 = a00;
      if (w01) *o01 = a01;
      if (w10) *o10 = a10;
      if (w11) *o11 = a11;
    }
    for (int i = gnt - 1 - gtid; i < m; i += gnt) {     // right-hand side "row": rhs_i -= L21[i,:] . y_j  (last CTAs first)
      const double* li = A + (size_t)(base + i) * n + j0;
      double v = d.qc[base + i];
      for (int t = 0; t < nb; ++t) v -= li[t] * d.qc[j0 + t];
      d.qc[base + i] = v;
    }
    PROF(6)
    grid.sync();
    PROF(1)
  }
  if (cta != 0) return;
  // 5. solves (CTA 0)
  __syncthreads();
  for (int i = tid; i < n; i += nt) xs[i] = d.qc[i];
  __syncthreads();
  trsv_blocked(d.Sf, n, n, xs, Dg, true);       // the forward substitution happened during the factorisation
  for (int i = tid; i < n; i += nt) {
    d.pc[i] = xs[i]; coef[i] = xs[i];
    const int c = i / 6, a = i % 6, C = n / 6;
    step[a < 3 ? 3 * c + a : 3 * C + 3 * c + (a - 3)] = xs[i];
  }
  // 6. w = T^-1 (h0 - G^T pc)
  __syncthreads();
  if (tid < B) {
    double hv = schur[(size_t)(n + tid) * R + n + B];
    for (int i = 0; i < n; ++i) hv -= schur[(size_t)i * R + n + tid] * xs[i];
    sh[tid] = hv;
  }
  __syncthreads();
  if (tid == 0) { small_fwd_r(sLT, s_trinv, B, sh); small_bwd_r(sLT, s_trinv, B, sh); }
  __syncthreads();
  if (tid < B) { d.w[tid] = sh[tid]; coef[n + tid] = sh[tid]; }
  PROF(7)
  if (prof && tid == 0)
    printf("dense coop cycles: init %lld gridsync %lld assemble %lld diagload %lld factor %lld panel load %lld trsm %lld store %lld trailing %lld solve %lld\n",
           tp[0], tp[1], tp[2], tp[3], tp[4], tp[8], tp[9], tp[5], tp[6], tp[7]);
#undef PROF
}

// second reduced solve for phi'(alpha) and the step scalars
//   rhs2 = [ rhs0' (n) | h0' (B) | sum|p_p|^2 | sum g_p.p_p | sum |L^-1 p_p|^2 | . ]
//   out  = { |p|^2, g.p, p^T (H + alpha I)^-1 p, 0 }
__global__ void __launch_bounds__(DN_THREADS)
k_dense_solve2(const double* __restrict__ schur, int R, const double* __restrict__ lin, int n, int B,
               const double* __restrict__ rhs2, double* __restrict__ ws, int need_phi, double* __restrict__ out) {
  extern __shared__ double smem[];
  double* Dg = smem;
  double* xs = Dg + NBK * LDS_;
  double* sLT = xs + ((n + 1) & ~1);
  double* sh = sLT + BAX_MAX_B * BAX_MAX_B;
  double* sred = sh + BAX_MAX_B;            // 4*32
  const int tid = threadIdx.x, nt = blockDim.x;
  DenseWs d = dense_ws(ws, n, B);
  const double* gc = lin + (size_t)n * n;
  double acc[4] = {0, 0, 0, 0};             // |pc|^2, gc.pc, pc.qc, rhs0'.qc
  if (need_phi) {
    for (int i = tid; i < BAX_MAX_B * BAX_MAX_B; i += nt) sLT[i] = d.LT[i];
    if (tid < BAX_MAX_B) sh[tid] = tid < B ? rhs2[n + tid] : 0.0;
    __syncthreads();
    if (tid == 0) small_fwd(sLT, B, sh);    // hz' = LT^-1 h0'
    __syncthreads();
    for (int i = tid; i < n; i += nt) {
      double r = d.pc[i] - rhs2[i];
      for (int a = 0; a < B; ++a) r += d.Z[(size_t)a * n + i] * sh[a];
      xs[i] = r;
    }
    __syncthreads();
    trsv_blocked(d.Sf, n, n, xs, Dg, false);
    trsv_blocked(d.Sf, n, n, xs, Dg, true);   // qc
    if (tid < B) {
      double hv = rhs2[n + tid];
      for (int i = 0; i < n; ++i) hv -= schur[(size_t)i * R + n + tid] * xs[i];
      sh[tid] = hv;
    }
    __syncthreads();
    if (tid == 0) { small_fwd(sLT, B, sh); small_bwd(sLT, B, sh); }   // w'
    __syncthreads();
  }
  for (int i = tid; i < n; i += nt) {
    double p = d.pc[i];
    acc[0] += p * p; acc[1] += gc[i] * p;
    if (need_phi) { acc[2] += p * xs[i]; acc[3] += rhs2[i] * xs[i]; }
  }
  block_sum<4>(acc, sred);
  if (tid == 0) {
    double hw = 0;
    if (need_phi) for (int a = 0; a < B; ++a) hw += rhs2[n + a] * sh[a];
    out[0] = acc[0] + rhs2[n + B];
    out[1] = acc[1] + rhs2[n + B + 1];
    out[2] = need_phi ? acc[2] + rhs2[n + B + 2] - acc[3] - hw : 0.0;
    out[3] = 0.0;
  }
}

// camera part of the step in theta order: internal 6c+{0..2} -> 3c+k, 6c+{3..5} -> 3C+3c+k
__global__ void k_scatter_cam_step(const double* __restrict__ coef, int C, double* __restrict__ step) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 6 * C) return;
  int c = i / 6, a = i % 6;
  step[a < 3 ? 3 * c + a : 3 * C + 3 * c + (a - 3)] = coef[i];
}
int bax_scatter_cam_step(bax_handle* h, double* step, cudaStream_t s) {
  k_scatter_cam_step<<<(6 * h->C + 127) / 128, 128, 0, s>>>(h->coef, h->C, step);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}

static size_t dense_smem(int n) {
  size_t d = NBK * LDS_ + ((n + 1) & ~1) + BAX_MAX_B * BAX_MAX_B + BAX_MAX_B + 4 * 32;
  size_t pn = n > NBK ? (size_t)(n - NBK) * LDS_ : 0;
  return (d + pn) * sizeof(double);
}

int bax_dense_solve1(bax_handle* h, double* step, cudaStream_t s) {   // also writes the camera part of the step (theta order)
  const int n = 6 * h->C;
  if (n > NBK && h->dense_coop) {
    const int mb = (n + NBK - 1) / NBK - 1;
    int G = mb * (mb + 1) / 2;
    if (G > 64) G = 64;
    if (G > h->sm_count) G = h->sm_count;
    if (G < 1) G = 1;
    size_t sm = sizeof(double) * (3 * NBK * LDS_ + BAX_MAX_B * BAX_MAX_B + BAX_MAX_B + 3 * NBK + NBK * LDS_ + ((n + 1) & ~1));
    BAX_CUDA(cudaFuncSetAttribute(k_dense_solve1_coop, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    const double* schur = h->buf[BAX_BUF_SCHUR]; const double* lin = h->buf[BAX_BUF_LIN];
    int R = h->R, nn = n, B = h->B; double alpha = h->alpha; double* ws = h->dense_ws; double* coef = h->coef;
    int prof = h->dense_coop > 1;
    static const int chol_warp = getenv("BAX_CHOL_WARP") ? atoi(getenv("BAX_CHOL_WARP")) : 1;   // 0: the 256-thread diagonal-block factorisation
    int cw = chol_warp;
    void* args[] = {(void*)&schur, (void*)&R, (void*)&lin, (void*)&nn, (void*)&B, (void*)&alpha, (void*)&ws, (void*)&coef, (void*)&prof, (void*)&cw, (void*)&step};
    BaxTimer t(h, BAX_T_DENSE, s, 1);
    BAX_CUDA(cudaLaunchCooperativeKernel((void*)k_dense_solve1_coop, dim3(G), dim3(CG_THREADS), args, sm, s));
    return BAX_OK;
  }
  size_t sm = dense_smem(n);
  if (sm > 227 * 1024) { bax_set_error("reduced camera system of %d unknowns exceeds the single-CTA solver", n); return BAX_E_ARG; }
  BaxTimer t(h, BAX_T_DENSE, s, 1);
  static const int small_nt = getenv("BAX_DENSE_SMALL_NT") ? atoi(getenv("BAX_DENSE_SMALL_NT")) : 256;   // tuning aid: 1024 = one instance for all
  if (n <= 96 && small_nt == 256) {
    BAX_CUDA(cudaFuncSetAttribute(k_dense_solve1<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    k_dense_solve1<256><<<1, 256, sm, s>>>(h->buf[BAX_BUF_SCHUR], h->R, h->buf[BAX_BUF_LIN], n, h->B, h->alpha, h->dense_ws, h->coef, step);
  } else {
    BAX_CUDA(cudaFuncSetAttribute(k_dense_solve1<DN_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    k_dense_solve1<DN_THREADS><<<1, DN_THREADS, sm, s>>>(h->buf[BAX_BUF_SCHUR], h->R, h->buf[BAX_BUF_LIN], n, h->B, h->alpha,
                                                         h->dense_ws, h->coef, step);
  }
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}
int bax_dense_solve2(bax_handle* h, cudaStream_t s) {
  const int n = 6 * h->C;
  size_t sm = dense_smem(n);
  BAX_CUDA(cudaFuncSetAttribute(k_dense_solve2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  BaxTimer t(h, BAX_T_DENSE, s, 1);
  k_dense_solve2<<<1, DN_THREADS, sm, s>>>(h->buf[BAX_BUF_SCHUR], h->R, h->buf[BAX_BUF_LIN], n, h->B,
                                           h->buf[BAX_BUF_RHS2], h->dense_ws, h->need_phi, h->buf[BAX_BUF_STEP]);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}
