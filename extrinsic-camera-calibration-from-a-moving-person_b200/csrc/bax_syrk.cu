// out = X X^T for a short-and-wide FP64 matrix X [rows][K] (row-major, K contiguous, ldx % 32 == 0, columns
// K..ldx zero).  Lower 64x64 tiles x split-K CTAs, cp.async double-buffered shared-memory staging, 4x4
// register tiles with row-strided ownership (conflict-free 128-bit shared loads), deterministic two-stage sum.
#include "bax_internal.cuh"
#include <stdlib.h>
#include <stdio.h>

#define ST 64            // tile edge
#define KC 32            // K columns per stage
#define LDT (KC + 2)     // padded row stride (doubles): 272 B, keeps 16-B alignment, spreads banks

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, bool valid) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void load_tile(double* dst, const double* __restrict__ X, int row0, int rows, int64_t ldx, int64_t k0) {
  // 64 rows x 32 columns = 1024 16-byte words; 256 threads x 4
  const int t = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int w = t + 256 * i;
    int r = w >> 4, c2 = (w & 15) * 2;
    bool ok = (row0 + r) < rows;
    const double* src = X + (size_t)(ok ? row0 + r : 0) * ldx + k0 + c2;
    cp_async16(dst + r * LDT + c2, src, ok);
  }
}

__global__ void __launch_bounds__(256)
k_syrk(const double* __restrict__ X, int rows, int64_t ldx, int64_t kper, int64_t Kpad, double* __restrict__ part) {
  extern __shared__ __align__(16) double sm[];   // [2 stages][A|B][64][LDT]
  const int t = blockIdx.x;
  int bi = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
  while ((bi + 1) * (bi + 2) / 2 <= t) ++bi;
  while (bi * (bi + 1) / 2 > t) --bi;
  const int bj = t - bi * (bi + 1) / 2;
  const bool diag = (bi == bj);
  const int64_t kbeg = (int64_t)blockIdx.y * kper;
  int64_t kend = kbeg + kper; if (kend > Kpad) kend = Kpad;
  const int nchunk = (int)((kend - kbeg + KC - 1) / KC);
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  double acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
  const int stage_sz = 2 * ST * LDT;
  if (nchunk > 0) {
    load_tile(sm, X, bi * ST, rows, ldx, kbeg);
    if (!diag) load_tile(sm + ST * LDT, X, bj * ST, rows, ldx, kbeg);
    cp_async_commit();
  }
  for (int ch = 0; ch < nchunk; ++ch) {
    double* cur = sm + (ch & 1) * stage_sz;
    if (ch + 1 < nchunk) {
      double* nxt = sm + ((ch + 1) & 1) * stage_sz;
      load_tile(nxt, X, bi * ST, rows, ldx, kbeg + (int64_t)(ch + 1) * KC);
      if (!diag) load_tile(nxt + ST * LDT, X, bj * ST, rows, ldx, kbeg + (int64_t)(ch + 1) * KC);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const double* As = cur;
    const double* Bs = diag ? cur : cur + ST * LDT;
#pragma unroll 4
    for (int kk = 0; kk < KC; kk += 2) {
      double2 a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const double2*>(As + (ty + 16 * i) * LDT + kk);
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const double2*>(Bs + (tx + 16 * j) * LDT + kk);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] += a[i].x * b[j].x + a[i].y * b[j].y;
    }
    __syncthreads();
  }
  double* o = part + ((size_t)blockIdx.y * gridDim.x + t) * (ST * ST);
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) o[(ty + 16 * i) * ST + tx + 16 * j] = acc[i][j];
}

// Second stage of the split-K sum: 32 outputs x 8 slices of the split range per block (a small problem is cut into up
// to ~200 pieces to fill the GPU; one thread walking all of them per output took 21 us for a 37 x 37 result).  The order
// of the additions is fixed: every slice sums its pieces z = s, s + 8, ... in turn, then the slices are added 0..7.
__global__ void __launch_bounds__(256)
k_syrk_reduce(const double* __restrict__ part, int rows, int ntiles, int nsplit, double* __restrict__ out, int ldo) {
  __shared__ double red[8][33];
  const int o = threadIdx.x & 31, sl = threadIdx.x >> 5;
  const int idx = blockIdx.x * 32 + o;
  const bool ok = idx < rows * rows;
  int i = 0, j = 0;
  double acc = 0.0;
  if (ok) {
    i = idx / rows; j = idx % rows;
    const int r = i >= j ? i : j, c = i >= j ? j : i;
    const int bi = r / ST, bj = c / ST;
    const int t = bi * (bi + 1) / 2 + bj;
    const size_t off = (size_t)t * (ST * ST) + (size_t)(r % ST) * ST + (c % ST);
    for (int z = sl; z < nsplit; z += 8) acc += part[(size_t)z * ntiles * (ST * ST) + off];
  }
  red[sl][o] = acc;
  __syncthreads();
  if (sl == 0 && ok) {
    double v = red[0][o];
#pragma unroll
    for (int s = 1; s < 8; ++s) v += red[s][o];
    out[(size_t)i * ldo + j] = v;
  }
}

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// ------------------------------------------------------------------ up to 64 rows on the FP64 tensor pipe
// The reference's own scenes have 3-5 cameras: 6C + B + 1 = 31..43 rows (and 3C + 1 = 10..16 for the orientation rows).
// k_syrk above computes a full 64 x 64 tile for them (4096 products per column where 37 rows need 703) on the DFMA pipe
// and cuts K into ~200 CTAs of 256 threads that fill the whole GPU for a 20-30 us kernel — which is also what keeps the
// streams of a sweep over many small problems (bax_batch_solve) from overlapping.  Here one WARP is a split-K unit: it
// keeps the NT8 (NT8 + 1) / 2 lower m8n8 accumulator tiles of the whole product in registers, reads its operand
// fragments X[8 i + g][k + tig] straight from global memory (four lanes share a 32-byte sector, four consecutive steps a
// 128-byte line held in L1; Yhat of such a problem lives in L2) one 16-column step ahead, and issues NT8 (NT8 + 1) / 2
// DMMA per 4 columns.  Partials go out in the 64 x 64 layout k_syrk_reduce already sums in a fixed order.
template <int NT8>
__device__ __forceinline__ void ssm_load(double (&f)[4][NT8], const double* const (&rp)[NT8], const bool (&ok)[NT8], int64_t k) {
#pragma unroll
  for (int s4 = 0; s4 < 4; ++s4)
#pragma unroll
    for (int i = 0; i < NT8; ++i) f[s4][i] = ok[i] ? rp[i][k + 4 * s4] : 0.0;
}
template <int NT8>
__device__ __forceinline__ void ssm_mma(double (&acc)[NT8 * (NT8 + 1) / 2][2], const double (&f)[4][NT8]) {
#pragma unroll
  for (int s4 = 0; s4 < 4; ++s4) {
    int t = 0;
#pragma unroll
    for (int i = 0; i < NT8; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) { dmma884(acc[t], f[s4][i], f[s4][j]); ++t; }
  }
}
template <int NT8>
__global__ void __launch_bounds__(128)
k_syrk_small(const double* __restrict__ X, int rows, int64_t ldx, int64_t kw /*columns per warp, multiple of 32*/, int64_t Kpad,
             int nunits, double* __restrict__ part /*[nunits][64 x 64]*/) {
  const int lane = threadIdx.x & 31, g = lane >> 2, tig = lane & 3;
  const int unit = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (unit >= nunits) return;
  const int64_t kbeg = (int64_t)unit * kw;
  int64_t kend = kbeg + kw; if (kend > Kpad) kend = Kpad;
  const double* rp[NT8]; bool ok[NT8];
#pragma unroll
  for (int i = 0; i < NT8; ++i) { ok[i] = 8 * i + g < rows; rp[i] = X + (size_t)(ok[i] ? 8 * i + g : 0) * ldx + tig; }
  double acc[NT8 * (NT8 + 1) / 2][2];
#pragma unroll
  for (int t = 0; t < NT8 * (NT8 + 1) / 2; ++t) { acc[t][0] = 0.0; acc[t][1] = 0.0; }
  double fa[4][NT8], fb[4][NT8];
  ssm_load<NT8>(fa, rp, ok, kbeg);
  for (int64_t k = kbeg; k < kend; k += 32) {    // kw and Kpad are multiples of 32: two 16-column steps per round
    ssm_load<NT8>(fb, rp, ok, k + 16);
    ssm_mma<NT8>(acc, fa);
    if (k + 32 < kend) ssm_load<NT8>(fa, rp, ok, k + 32);
    ssm_mma<NT8>(acc, fb);
  }
  double* o = part + (size_t)unit * (ST * ST);
  int t = 0;
#pragma unroll
  for (int i = 0; i < NT8; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      *reinterpret_cast<double2*>(o + (8 * i + g) * ST + 8 * j + 2 * tig) = make_double2(acc[t][0], acc[t][1]);
      ++t;
    }
}

// ------------------------------------------------------------------ 96 x 96 tiles on the FP64 tensor pipe (DMMA)
// Measured on this B200 (tools/micro/fp64_peak.cu): mma.sync.m8n8k4.f64 sustains 37.1 TFLOP/s, the DFMA vector pipe
// 34 TFLOP/s and only with >= 16 resident warps; a 6 x 6 DFMA register tile reached 61 % pipe utilisation.  Here a
// CTA of 4 warps owns a 96 x 96 tile; in an off-diagonal tile each warp owns a 48 x 48 quadrant = 6 x 6 m8n8 accumulator
// tiles (72 doubles per thread) and per 4 K columns issues 12 64-bit shared loads and 36 DMMA (diagonal tiles: below).  Both operands are rows of X, read as
// X[row0 + (lane >> 2)][k0 + (lane & 3)]: with a row stride of 20 doubles the 32 lanes hit 32 distinct 8-byte banks
// slots twice (two wavefronts, the minimum for 256 bytes).  3-stage cp.async pipeline of 16 K columns, 2 CTAs / SM.
#define T9 96
#ifndef KC9
#define KC9 16
#define ST9 3
#endif
#define LD9 (KC9 + 4)
#define STAGE9 (2 * T9 * LD9)

__device__ __forceinline__ void load_tile96(double* dst, const double* __restrict__ X, int row0, int rows, int64_t ldx, int64_t k0) {
  const int t = threadIdx.x;           // 96 rows x KC9 columns = 48 KC9 16-byte words; 128 threads x (3 KC9 / 8)
#pragma unroll
  for (int i = 0; i < 3 * KC9 / 8; ++i) {
    int w = t + 128 * i;
    int r = w / (KC9 / 2), c2 = (w % (KC9 / 2)) * 2;
    bool ok = (row0 + r) < rows;
    const double* src = X + (size_t)(ok ? row0 + r : 0) * ldx + k0 + c2;
    cp_async16(dst + r * LD9 + c2, src, ok);
  }
}


// Trailing rows ("extras": the bone-length and right-hand-side rows of Yhat, the residual row of the orientation
// Jacobian) ride along in the diagonal tiles: every tile row of a diagonal tile also forms its 8 x 16 piece of the panel
// X[tile rows] . X[extras]^T, and tile 0 the 16 x 16 corner  X[extras] . X[extras]^T.  The extras cost one more 16-row
// slab per stage and no extra pass over X.
#define NE9 16
#define STAGE9E STAGE9                           // the extras slab of a diagonal tile lies where an off-diagonal tile keeps B
#define EOFF9 (T9 * LD9)
#define EP9 ((T9 + NE9) * NE9)                   // doubles of one panel partial: [96 panel rows + 16 corner rows][16]

// Work of a DIAGONAL tile.  Only its lower triangle is read back, so computing 48 x 48 quadrants there wastes a third of
// the tensor-pipe time of the tile (3 full quadrants + the 96 x 16 extras panel: the slowest warp issues 36 DMMA per 4
// columns, as in an off-diagonal tile).  Instead the 12 x 12 grid of m8n8 accumulator tiles is cut along its lower
// triangle: 78 tiles (r, c <= r) dealt to the four warps by whole tile rows {0, 11, 5}, {1, 10, 6}, {2, 9, 4}, {3, 8, 7}
// (19, 20, 18, 21 tiles), and the 12 x 2 tiles of the extras panel by tile rows {0, 11, 5}, {1, 10, 6}, {2, 9, 4, 3},
// {8, 7}: 25, 26, 26, 25 tiles per warp (the 2 x 2 extras corner of tile 0 adds 2 to the first and the last warp).
// Row and column fragments of a diagonal tile are the same shared-memory words, so a warp loads at most 12 + 2 fragments
// per 4 columns.  A diagonal-tile CTA then costs about 0.8 of an off-diagonal one per K column (measured; 26 / 36 by
// the DMMA count), and K is cut per tile kind (bax_syrk_aug) so that both kinds of CTA finish together.
#define DIAG_COST 0.8

__host__ __device__ constexpr int dg_row(int w, int q) {
  return w == 0 ? (q == 0 ? 0 : q == 1 ? 11 : 5) : w == 1 ? (q == 0 ? 1 : q == 1 ? 10 : 6)
       : w == 2 ? (q == 0 ? 2 : q == 1 ? 9 : 4) : (q == 0 ? 3 : q == 1 ? 8 : 7);
}
__host__ __device__ constexpr int dg_base(int w, int q) {      // first accumulator of tile row q of warp w
  return q == 0 ? 0 : q == 1 ? dg_row(w, 0) + 1 : dg_row(w, 0) + dg_row(w, 1) + 2;
}
__host__ __device__ constexpr int dg_ebase(int w) { return dg_row(w, 0) + dg_row(w, 1) + dg_row(w, 2) + 3; }
__host__ __device__ constexpr int dg_maxrow(int w) { return dg_row(w, 1); }
__host__ __device__ constexpr int dg_nerow(int w) { return w == 2 ? 4 : w == 3 ? 2 : 3; }   // extras-panel tile rows
__host__ __device__ constexpr int dg_erow(int w, int q) {
  return w == 3 ? (q == 0 ? 8 : 7) : q < 3 ? dg_row(w, q) : 3;
}
__host__ __device__ constexpr int dg_cbase(int w) { return dg_ebase(w) + 2 * dg_nerow(w); }
static_assert(dg_cbase(0) + 2 <= 36 && dg_cbase(1) <= 36 && dg_cbase(2) <= 36 && dg_cbase(3) + 2 <= 36, "accumulators");

template <int R, int BASE>
__device__ __forceinline__ void dg_row_mma(double (&acc)[36][2], const double (&f)[12]) {
#pragma unroll
  for (int c = 0; c <= R; ++c) dmma884(acc[BASE + c], f[R], f[c]);
}
template <int R, int BASE>
__device__ __forceinline__ void dg_row_store(const double (&acc)[36][2], double* o, int g, int tig) {
#pragma unroll
  for (int c = 0; c <= R; ++c)
    *reinterpret_cast<double2*>(o + (8 * R + g) * T9 + 8 * c + 2 * tig) = make_double2(acc[BASE + c][0], acc[BASE + c][1]);
}

// one 16-column chunk of warp W of a diagonal tile; WITH_E / CORNER are CTA-uniform and kept out of the unrolled loop so
// that the fragment loads of the next 4 columns can be hoisted over the DMMAs of the current ones
template <int W, bool WITH_E, bool CORNER>
__device__ __forceinline__ void dg_chunk(double (&acc)[36][2], const double* As, const double* Es, int g, int tig) {
  constexpr int EB = dg_ebase(W), CB = dg_cbase(W);
  const double* fp = As + g * LD9 + tig;
  const double* ep = Es + g * LD9 + tig;
#pragma unroll 2                                 // not 4: see the note on code size at k_syrk96
  for (int k4 = 0; k4 < KC9; k4 += 4) {
    double f[12];
#pragma unroll
    for (int c = 0; c <= dg_maxrow(W); ++c) f[c] = fp[8 * c * LD9 + k4];
    dg_row_mma<dg_row(W, 0), dg_base(W, 0)>(acc, f);
    dg_row_mma<dg_row(W, 1), dg_base(W, 1)>(acc, f);
    dg_row_mma<dg_row(W, 2), dg_base(W, 2)>(acc, f);
    if (WITH_E) {
      const double e0 = ep[k4], e1 = ep[8 * LD9 + k4];
#pragma unroll
      for (int q = 0; q < dg_nerow(W); ++q) {
        dmma884(acc[EB + 2 * q], f[dg_erow(W, q)], e0); dmma884(acc[EB + 2 * q + 1], f[dg_erow(W, q)], e1);
      }
      if (CORNER && W == 0) { dmma884(acc[CB], e0, e0); dmma884(acc[CB + 1], e0, e1); }
      if (CORNER && W == 3) { dmma884(acc[CB], e1, e0); dmma884(acc[CB + 1], e1, e1); }
    }
  }
}

template <int W>
__device__ __forceinline__ void dg_chunk_any(double (&acc)[36][2], const double* As, const double* Es, bool with_e, bool corner,
                                             int g, int tig) {
  if (!with_e) dg_chunk<W, false, false>(acc, As, Es, g, tig);
  else if (corner && (W == 0 || W == 3)) dg_chunk<W, true, true>(acc, As, Es, g, tig);
  else dg_chunk<W, true, false>(acc, As, Es, g, tig);
}

template <int W>
__device__ __forceinline__ void dg_store(const double (&acc)[36][2], double* o, double* oe, bool with_e, bool corner, int g, int tig) {
  constexpr int EB = dg_ebase(W), CB = dg_cbase(W);
  dg_row_store<dg_row(W, 0), dg_base(W, 0)>(acc, o, g, tig);
  dg_row_store<dg_row(W, 1), dg_base(W, 1)>(acc, o, g, tig);
  dg_row_store<dg_row(W, 2), dg_base(W, 2)>(acc, o, g, tig);
  if (with_e) {
#pragma unroll
    for (int q = 0; q < dg_nerow(W); ++q)
#pragma unroll
      for (int j = 0; j < 2; ++j)
        *reinterpret_cast<double2*>(oe + (8 * dg_erow(W, q) + g) * NE9 + 8 * j + 2 * tig) = make_double2(acc[EB + 2 * q + j][0], acc[EB + 2 * q + j][1]);
    if (corner && (W == 0 || W == 3)) {          // corner rows 0..7 from the first warp, 8..15 from the last
#pragma unroll
      for (int j = 0; j < 2; ++j)
        *reinterpret_cast<double2*>(oe + (T9 + (W == 3 ? 8 : 0) + g) * NE9 + 8 * j + 2 * tig) = make_double2(acc[CB + j][0], acc[CB + j][1]);
    }
  }
}

// Code size matters here.  An SM that hosts one CTA of each kind runs five different inner loops (the off-diagonal one
// and the four warps' of a diagonal tile); with the 16 columns of a chunk fully unrolled they are 4.8 + 4 x 3.6 KB, and
// builds with IDENTICAL instruction streams and control codes but different code placement ran the 64-camera product in
// 3.19 or 3.42 ms (pure diagonal-tile grids were unaffected: one loop kind per SM).  Unrolled by two (8 columns) the loops
// are half that, stay inside the instruction caches whatever their placement, and every build measures 3.20 ms.
// CTAs [0, noff * ns_off): off-diagonal tiles, K cut into ns_off pieces of kper_off columns;
// CTAs [noff * ns_off, + nt * ns_diag): diagonal tiles, K cut into ns_diag pieces of kper_diag columns.
// part is indexed [z][tile][96 x 96] with tile = bi (bi + 1) / 2 + bj and z < the split count of that tile's kind.
__global__ void __launch_bounds__(128, 2)
k_syrk96(const double* __restrict__ X, int rows, int ne, int64_t ldx, int nt, int ns_off, int ns_diag, int64_t kper_off,
         int64_t kper_diag, int64_t Kpad, double* __restrict__ part, double* __restrict__ epart) {
  extern __shared__ __align__(16) double sm[];   // [ST9 stages][A|B|E][..][LD9]
  const int noff = nt * (nt - 1) / 2, ntiles = noff + nt;
  int b = blockIdx.x, bi, bj, z;
  const bool diag = b >= noff * ns_off;
  if (!diag) {
    z = b / noff; b -= z * noff;                 // b-th off-diagonal tile: b = bi (bi - 1) / 2 + bj, bj < bi
    bi = (int)((sqrt(8.0 * b + 1.0) + 1.0) * 0.5);
    while (bi * (bi - 1) / 2 > b) --bi;
    while ((bi + 1) * bi / 2 <= b) ++bi;
    bj = b - bi * (bi - 1) / 2;
  } else {
    b -= noff * ns_off;
    z = b / nt; bi = bj = b - z * nt;
  }
  const int t = bi * (bi + 1) / 2 + bj;
  const int64_t kper = diag ? kper_diag : kper_off;
  const int64_t kbeg = (int64_t)z * kper;
  int64_t kend = kbeg + kper; if (kend > Kpad) kend = Kpad;
  const int nchunk = (int)((kend - kbeg + KC9 - 1) / KC9);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wy = warp >> 1, wx = warp & 1, g = lane >> 2, tig = lane & 3;
  const bool with_e = diag && ne > 0;
  const bool corner = bi == 0;
  double acc[36][2];
#pragma unroll
  for (int i = 0; i < 36; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
  auto load_stage = [&](int st, int64_t k0) {
    double* d = sm + st * STAGE9E;
    load_tile96(d, X, bi * T9, rows, ldx, k0);
    if (!diag) load_tile96(d + T9 * LD9, X, bj * T9, rows, ldx, k0);
    else if (with_e) {                           // 16 rows x KC9 columns = 8 KC9 16-byte words
#pragma unroll
      for (int i = 0; i < KC9 / 16; ++i) {
        const int w = threadIdx.x + 128 * i;
        const int r = w / (KC9 / 2), c2 = (w % (KC9 / 2)) * 2;
        const bool ok = r < ne;
        cp_async16(d + EOFF9 + r * LD9 + c2, X + (size_t)(ok ? rows + r : 0) * ldx + k0 + c2, ok);
      }
    }
  };
#pragma unroll
  for (int st = 0; st < ST9 - 1; ++st) {
    if (st < nchunk) load_stage(st, kbeg + (int64_t)st * KC9);
    cp_async_commit();
  }
  for (int ch = 0; ch < nchunk; ++ch) {
    cp_async_wait<ST9 - 2>();                    // chunk ch has landed (this thread's copies)
    __syncthreads();                             // ... everyone's; and everyone is done reading chunk ch-1
    {
      const int nx = ch + ST9 - 1;
      if (nx < nchunk) load_stage(nx % ST9, kbeg + (int64_t)nx * KC9);
      cp_async_commit();
    }
    const double* As = sm + (ch % ST9) * STAGE9E;
    if (!diag) {
      const double* ap = As + (48 * wy + g) * LD9 + tig;
      const double* bp = As + T9 * LD9 + (48 * wx + g) * LD9 + tig;
#pragma unroll 2
      for (int k4 = 0; k4 < KC9; k4 += 4) {
        double a[6], bb[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) a[i] = ap[8 * i * LD9 + k4];
#pragma unroll
        for (int j = 0; j < 6; ++j) bb[j] = bp[8 * j * LD9 + k4];
#pragma unroll
        for (int i = 0; i < 6; ++i)
#pragma unroll
          for (int j = 0; j < 6; ++j) dmma884(acc[6 * i + j], a[i], bb[j]);
      }
    } else {
      const double* Es = As + EOFF9;
      switch (warp) {                            // uniform per warp
        case 0: dg_chunk_any<0>(acc, As, Es, with_e, corner, g, tig); break;
        case 1: dg_chunk_any<1>(acc, As, Es, with_e, corner, g, tig); break;
        case 2: dg_chunk_any<2>(acc, As, Es, with_e, corner, g, tig); break;
        default: dg_chunk_any<3>(acc, As, Es, with_e, corner, g, tig); break;
      }
    }
  }
  cp_async_wait<0>();
  double* o = part + ((size_t)z * ntiles + t) * (T9 * T9);
  if (!diag) {
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 6; ++j)
        *reinterpret_cast<double2*>(o + (48 * wy + 8 * i + g) * T9 + 48 * wx + 8 * j + 2 * tig) = make_double2(acc[6 * i + j][0], acc[6 * i + j][1]);
  } else {
    double* oe = epart + ((size_t)z * ntiles + t) * EP9;      // indexed like `part`; only diagonal tiles are written
    switch (warp) {
      case 0: dg_store<0>(acc, o, oe, with_e, corner, g, tig); break;
      case 1: dg_store<1>(acc, o, oe, with_e, corner, g, tig); break;
      case 2: dg_store<2>(acc, o, oe, with_e, corner, g, tig); break;
      default: dg_store<3>(acc, o, oe, with_e, corner, g, tig); break;
    }
  }
}

// extras of the output: out[r][rows + e] (and its mirror) for r < rows, and the ne x ne corner
__global__ void k_syrk96_reduce_e(const double* __restrict__ epart, int rows, int ne, int ntiles, int nsplit /*of the diagonal tiles*/,
                                  double* __restrict__ out, int ldo) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (rows + ne) * ne) return;
  const int r = idx / ne, e = idx % ne;
  int tile, row;
  if (r < rows) { const int b = r / T9; tile = b * (b + 1) / 2 + b; row = r % T9; }
  else { tile = 0; row = T9 + (r - rows); }
  double acc = 0.0;
  for (int z = 0; z < nsplit; ++z) acc += epart[((size_t)z * ntiles + tile) * EP9 + row * NE9 + e];
  out[(size_t)r * ldo + rows + e] = acc;
  if (r < rows) out[(size_t)(rows + e) * ldo + r] = acc;
}

__global__ void k_syrk96_reduce(const double* __restrict__ part, int rows, int ntiles, int ns_off, int ns_diag, double* __restrict__ out, int ldo) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * rows) return;
  int i = idx / rows, j = idx % rows;
  int r = i >= j ? i : j, c = i >= j ? j : i;
  int bi = r / T9, bj = c / T9;
  int t = bi * (bi + 1) / 2 + bj;
  size_t off = (size_t)t * (T9 * T9) + (size_t)(r % T9) * T9 + (c % T9);
  const int nsplit = bi == bj ? ns_diag : ns_off;
  double acc = 0.0;
  for (int z = 0; z < nsplit; ++z) acc += part[(size_t)z * ntiles * (T9 * T9) + off];
  out[(size_t)i * ldo + j] = acc;
}

static int ensure_part(bax_handle* h, size_t need) {
  if (need > h->syrk_part_bytes) {
    bax_dev_free(h, h->syrk_part);
    h->syrk_part = nullptr; h->syrk_part_bytes = 0;
    BAX_ALLOC(h, h->syrk_part, need);
    BAX_CUDA(cudaStreamSynchronize(h->mem_stream));
    h->syrk_part_bytes = need;
  }
  return BAX_OK;
}

// out ((rows + ne) x (rows + ne), leading dimension ldo, full symmetric) = X X^T for the first rows + ne rows of X;
// the ne <= 16 trailing rows ride along in the diagonal tiles (see k_syrk96).
int bax_syrk_aug(bax_handle* h, const double* X, int rows, int ne, int64_t K, int64_t ldx, double* out, int ldo, cudaStream_t s) {
  if (ldx % KC != 0) { bax_set_error("bax_syrk: ldx must be a multiple of %d", KC); return BAX_E_ARG; }
  if (ne < 0 || ne > NE9) { bax_set_error("bax_syrk: at most %d trailing rows", NE9); return BAX_E_ARG; }
  const int64_t Kpad = ((K + KC - 1) / KC) * KC;
  const int64_t nchunks = Kpad / KC;
  if (rows + ne > ST) {
    const int nt = (rows + T9 - 1) / T9;
    const int ntiles = nt * (nt + 1) / 2;
    const int noff = ntiles - nt;
    // Two CTAs per SM, a single wave.  A diagonal-tile CTA costs DIAG_COST (measured ~0.8) of an off-diagonal one per K
    // column, so the diagonal tiles are cut into fewer, longer pieces.  Of all (ns_off, ns_diag) whose CTAs fit the wave,
    // take the pair with the shortest slowest CTA (64 cameras on 148 SMs: 32 and 26 pieces, 296 CTAs; measured against
    // 33/24, 32/25, 32/24, 31/27, 34/23: the rule's pick is the fastest).  Neither kind is cut below 8 chunks (256
    // columns) per CTA.  BAX_SYRK_NS=a,d and BAX_SYRK_DIAG_COST override the rule for such measurements.
    const int64_t slots = 2 * (int64_t)h->sm_count;
    const int64_t maxsplit = (nchunks + 7) / 8 > 1 ? (nchunks + 7) / 8 : 1;
    int64_t ns_off = 0, ns_diag = 1;
    static const double rho = getenv("BAX_SYRK_DIAG_COST") ? atof(getenv("BAX_SYRK_DIAG_COST")) : DIAG_COST;   // tuning aids
    static const char* force = getenv("BAX_SYRK_NS");
    if (noff > 0) {
      double best = 1e300;
      ns_off = 1;
      for (int64_t a = 1; a * noff + nt <= slots; ++a) {
        const int64_t d = (slots - a * noff) / nt;                // the most diagonal pieces that still fit
        const double tmax = 1.0 / a > rho / d ? 1.0 / a : rho / d;
        if (tmax < best * (1.0 - 1e-12)) { best = tmax; ns_off = a; ns_diag = d; }
      }
      // no finer diagonal cut than needed to stay under the off-diagonal CTAs' time
      while (ns_diag > 1 && rho / (ns_diag - 1) <= 1.0 / ns_off) --ns_diag;
      if (force) { int a = 0, d = 0; if (sscanf(force, "%d,%d", &a, &d) == 2 && a > 0 && d > 0) { ns_off = a; ns_diag = d; } }
      if (ns_off > maxsplit) ns_off = maxsplit;
    } else {
      ns_diag = slots / nt;
    }
    if (ns_diag > maxsplit) ns_diag = maxsplit;
    if (ns_diag < 1) ns_diag = 1;
    int64_t kper_off = 0, kper_diag = ((nchunks + ns_diag - 1) / ns_diag) * KC;   // multiples of KC (32), hence of KC9
    ns_diag = (Kpad + kper_diag - 1) / kper_diag;
    if (noff > 0) { kper_off = ((nchunks + ns_off - 1) / ns_off) * KC; ns_off = (Kpad + kper_off - 1) / kper_off; }
    const int64_t nsmax = ns_off > ns_diag ? ns_off : ns_diag;
    const size_t main_doubles = (size_t)nsmax * ntiles * T9 * T9;
    const size_t e_doubles = ne > 0 ? (size_t)ns_diag * ntiles * EP9 : 0;
    int rc = ensure_part(h, sizeof(double) * (main_doubles + e_doubles)); if (rc) return rc;
    double* epart = h->syrk_part + main_doubles;
    const size_t smem = sizeof(double) * ST9 * STAGE9E;
    BAX_CUDA(cudaFuncSetAttribute(k_syrk96, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    BaxTimer t(h, BAX_T_SYRK, s, ne > 0 ? 3 : 2);
    k_syrk96<<<(unsigned)(noff * ns_off + nt * ns_diag), 128, smem, s>>>(X, rows, ne, ldx, nt, (int)ns_off, (int)ns_diag, kper_off,
                                                                          kper_diag, Kpad, h->syrk_part, epart);
    BAX_LAUNCH_CHECK();
    k_syrk96_reduce<<<(rows * rows + 255) / 256, 256, 0, s>>>(h->syrk_part, rows, ntiles, (int)ns_off, (int)ns_diag, out, ldo);
    BAX_LAUNCH_CHECK();
    if (ne > 0) {
      k_syrk96_reduce_e<<<((rows + ne) * ne + 255) / 256, 256, 0, s>>>(epart, rows, ne, ntiles, (int)ns_diag, out, ldo);
      BAX_LAUNCH_CHECK();
    }
    return BAX_OK;
  }
  rows += ne;                                    // small systems: one 64 x 64 tile covers everything
  static const int small_mma = getenv("BAX_SYRK_SMALL") ? atoi(getenv("BAX_SYRK_SMALL")) : 1;   // 0: the DFMA tile kernel below
  if (small_mma && rows <= 48) {                 // 7 and 8 tile rows do not fit the register file with a double-buffered operand set
    const int nt8 = (rows + 7) / 8;
    // one warp per split-K unit, at least 128 columns each, at most 8 warps per SM
    int64_t kw = (Kpad + 8 * (int64_t)h->sm_count - 1) / (8 * (int64_t)h->sm_count);
    kw = ((kw + KC - 1) / KC) * KC;
    if (kw < 128) kw = 128;
    const int nunits = (int)((Kpad + kw - 1) / kw);
    int rc = ensure_part(h, sizeof(double) * (size_t)nunits * ST * ST); if (rc) return rc;
    BaxTimer t(h, BAX_T_SYRK, s, 2);
    const int nb = (nunits + 3) / 4;
    switch (nt8) {
      case 1: case 2: k_syrk_small<2><<<nb, 128, 0, s>>>(X, rows, ldx, kw, Kpad, nunits, h->syrk_part); break;
      case 3: case 4: k_syrk_small<4><<<nb, 128, 0, s>>>(X, rows, ldx, kw, Kpad, nunits, h->syrk_part); break;
      case 5: k_syrk_small<5><<<nb, 128, 0, s>>>(X, rows, ldx, kw, Kpad, nunits, h->syrk_part); break;
      default: k_syrk_small<6><<<nb, 128, 0, s>>>(X, rows, ldx, kw, Kpad, nunits, h->syrk_part); break;
    }
    BAX_LAUNCH_CHECK();
    k_syrk_reduce<<<(rows * rows + 31) / 32, 256, 0, s>>>(h->syrk_part, rows, 1, nunits, out, ldo);
    BAX_LAUNCH_CHECK();
    return BAX_OK;
  }
  const int nt = (rows + ST - 1) / ST;
  const int ntiles = nt * (nt + 1) / 2;
  int64_t nsplit = (4 * (int64_t)h->sm_count + ntiles - 1) / ntiles;
  int64_t maxsplit = (nchunks + 7) / 8;          // at least 8 chunks (256 columns) per CTA
  if (nsplit > maxsplit) nsplit = maxsplit;
  if (nsplit < 1) nsplit = 1;
  int64_t kper = ((nchunks + nsplit - 1) / nsplit) * KC;
  nsplit = (Kpad + kper - 1) / kper;
  int rc = ensure_part(h, sizeof(double) * (size_t)nsplit * ntiles * ST * ST); if (rc) return rc;
  const size_t smem = sizeof(double) * 2 * 2 * ST * LDT;
  BAX_CUDA(cudaFuncSetAttribute(k_syrk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  {
    BaxTimer t(h, BAX_T_SYRK, s, 2);
    k_syrk<<<dim3(ntiles, (unsigned)nsplit), 256, smem, s>>>(X, rows, ldx, kper, Kpad, h->syrk_part);
    BAX_LAUNCH_CHECK();
    k_syrk_reduce<<<(rows * rows + 31) / 32, 256, 0, s>>>(h->syrk_part, rows, ntiles, (int)nsplit, out, ldo);
    BAX_LAUNCH_CHECK();
  }
  return BAX_OK;
}

int bax_syrk(bax_handle* h, const double* X, int rows, int64_t K, int64_t ldx, double* out, cudaStream_t s) {
  return bax_syrk_aug(h, X, rows, 0, K, ldx, out, rows, s);
}
