// out = X X^T for a short-and-wide FP64 matrix X [rows][K] (row-major, K contiguous, ldx % 32 == 0, columns
// K..ldx zero).  Lower 64x64 tiles x split-K CTAs, cp.async double-buffered shared-memory staging, 4x4
// register tiles with row-strided ownership (conflict-free 128-bit shared loads), deterministic two-stage sum.
#include "bax_internal.cuh"

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

// ------------------------------------------------------------------ 96 x 96 tiles on the FP64 tensor pipe (DMMA)
// Measured on this B200 (tools/micro/fp64_peak.cu): mma.sync.m8n8k4.f64 sustains 37.1 TFLOP/s, the DFMA vector pipe
// 34 TFLOP/s and only with >= 16 resident warps; a 6 x 6 DFMA register tile reached 61 % pipe utilisation.  Here a
// CTA of 4 warps owns a 96 x 96 tile, each warp a 48 x 48 quadrant = 6 x 6 m8n8 accumulator tiles (72 doubles per
// thread); per 4 K columns a warp issues 12 64-bit shared loads and 36 DMMA.  Both operands are rows of X, read as
// X[row0 + (lane >> 2)][k0 + (lane & 3)]: with a row stride of 20 doubles the 32 lanes hit 32 distinct 8-byte banks
// slots twice (two wavefronts, the minimum for 256 bytes).  3-stage cp.async pipeline of 16 K columns, 2 CTAs / SM.
#define T9 96
#define KC9 16
#define LD9 20
#define ST9 3
#define STAGE9 (2 * T9 * LD9)

__device__ __forceinline__ void load_tile96(double* dst, const double* __restrict__ X, int row0, int rows, int64_t ldx, int64_t k0) {
  const int t = threadIdx.x;           // 96 rows x 16 columns = 768 16-byte words; 128 threads x 6
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    int w = t + 128 * i;
    int r = w >> 3, c2 = (w & 7) * 2;
    bool ok = (row0 + r) < rows;
    const double* src = X + (size_t)(ok ? row0 + r : 0) * ldx + k0 + c2;
    cp_async16(dst + r * LD9 + c2, src, ok);
  }
}

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// Trailing rows ("extras": the bone-length and right-hand-side rows of Yhat, the residual row of the orientation
// Jacobian): in every diagonal tile the quadrant above the diagonal is never read back, so the warp that owns it
// computes instead the 96 x 16 panel  X[tile rows] . X[extras]^T  (12 x 2 accumulator tiles, 24 DMMA per 4 columns,
// fewer than the 36 of a full quadrant) and, in tile 0, the 16 x 16 corner  X[extras] . X[extras]^T.  The extras cost
// one more 16-row slab per stage and no extra pass over X.
#define NE9 16
#define STAGE9E (STAGE9 + NE9 * LD9)
#define EP9 ((T9 + NE9) * NE9)                   // doubles of one panel partial: [96 panel rows + 16 corner rows][16]

__global__ void __launch_bounds__(128, 2)
k_syrk96(const double* __restrict__ X, int rows, int ne, int64_t ldx, int64_t kper, int64_t Kpad,
         double* __restrict__ part, double* __restrict__ epart) {
  extern __shared__ __align__(16) double sm[];   // [ST9 stages][A|B|E][..][LD9]
  const int t = blockIdx.x;
  int bi = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
  while ((bi + 1) * (bi + 2) / 2 <= t) ++bi;
  while (bi * (bi + 1) / 2 > t) --bi;
  const int bj = t - bi * (bi + 1) / 2;
  const bool diag = (bi == bj);
  const int64_t kbeg = (int64_t)blockIdx.y * kper;
  int64_t kend = kbeg + kper; if (kend > Kpad) kend = Kpad;
  const int nchunk = (int)((kend - kbeg + KC9 - 1) / KC9);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wy = warp >> 1, wx = warp & 1, g = lane >> 2, tig = lane & 3;
  const bool upper = diag && wx > wy;            // quadrant strictly above the diagonal: never read back
  const bool with_e = diag && ne > 0;
  double acc[6][6][2];
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
  auto load_stage = [&](int st, int64_t k0) {
    double* d = sm + st * STAGE9E;
    load_tile96(d, X, bi * T9, rows, ldx, k0);
    if (!diag) load_tile96(d + T9 * LD9, X, bj * T9, rows, ldx, k0);
    else if (with_e) {                           // 16 rows x 16 columns = 128 16-byte words, one per thread
      const int r = threadIdx.x >> 3, c2 = (threadIdx.x & 7) * 2;
      const bool ok = r < ne;
      cp_async16(d + 2 * T9 * LD9 + r * LD9 + c2, X + (size_t)(ok ? rows + r : 0) * ldx + k0 + c2, ok);
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
    if (!upper) {
      const double* Bs = diag ? As : As + T9 * LD9;
      const double* ap = As + (48 * wy + g) * LD9 + tig;
      const double* bp = Bs + (48 * wx + g) * LD9 + tig;
#pragma unroll
      for (int k4 = 0; k4 < KC9; k4 += 4) {
        double a[6], b[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) a[i] = ap[8 * i * LD9 + k4];
#pragma unroll
        for (int j = 0; j < 6; ++j) b[j] = bp[8 * j * LD9 + k4];
#pragma unroll
        for (int i = 0; i < 6; ++i)
#pragma unroll
          for (int j = 0; j < 6; ++j) dmma884(acc[i][j], a[i], b[j]);
      }
    } else if (with_e) {
      // panel rows 8 i + g (i < 6) -> acc[i][0..1], rows 48 + 8 i + g -> acc[i][2..3]; corner -> acc[0..1][4..5]
      const double* ap = As + g * LD9 + tig;
      const double* ep = As + 2 * T9 * LD9 + g * LD9 + tig;
#pragma unroll
      for (int k4 = 0; k4 < KC9; k4 += 4) {
        double a[12], e[2];
#pragma unroll
        for (int i = 0; i < 12; ++i) a[i] = ap[8 * i * LD9 + k4];
        e[0] = ep[k4]; e[1] = ep[8 * LD9 + k4];
#pragma unroll
        for (int i = 0; i < 6; ++i) {
          dmma884(acc[i][0], a[i], e[0]); dmma884(acc[i][1], a[i], e[1]);
          dmma884(acc[i][2], a[6 + i], e[0]); dmma884(acc[i][3], a[6 + i], e[1]);
        }
        if (bi == 0) {
          dmma884(acc[0][4], e[0], e[0]); dmma884(acc[0][5], e[0], e[1]);
          dmma884(acc[1][4], e[1], e[0]); dmma884(acc[1][5], e[1], e[1]);
        }
      }
    }
  }
  cp_async_wait<0>();
  if (!upper) {
    double* o = part + ((size_t)blockIdx.y * gridDim.x + t) * (T9 * T9);
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 6; ++j)
        *reinterpret_cast<double2*>(o + (48 * wy + 8 * i + g) * T9 + 48 * wx + 8 * j + 2 * tig) = make_double2(acc[i][j][0], acc[i][j][1]);
  } else if (with_e) {
    double* o = epart + ((size_t)blockIdx.y * gridDim.x + t) * EP9;   // indexed like `part`; only diagonal tiles are written
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        *reinterpret_cast<double2*>(o + (8 * i + g) * NE9 + 8 * j + 2 * tig) = make_double2(acc[i][j][0], acc[i][j][1]);
        *reinterpret_cast<double2*>(o + (48 + 8 * i + g) * NE9 + 8 * j + 2 * tig) = make_double2(acc[i][2 + j][0], acc[i][2 + j][1]);
      }
    if (bi == 0) {
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j)
          *reinterpret_cast<double2*>(o + (T9 + 8 * i + g) * NE9 + 8 * j + 2 * tig) = make_double2(acc[i][4 + j][0], acc[i][4 + j][1]);
    }
  }
}

// extras of the output: out[r][rows + e] (and its mirror) for r < rows, and the ne x ne corner
__global__ void k_syrk96_reduce_e(const double* __restrict__ epart, int rows, int ne, int ntiles, int nsplit,
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

__global__ void k_syrk96_reduce(const double* __restrict__ part, int rows, int ntiles, int nsplit, double* __restrict__ out, int ldo) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * rows) return;
  int i = idx / rows, j = idx % rows;
  int r = i >= j ? i : j, c = i >= j ? j : i;
  int bi = r / T9, bj = c / T9;
  int t = bi * (bi + 1) / 2 + bj;
  size_t off = (size_t)t * (T9 * T9) + (size_t)(r % T9) * T9 + (c % T9);
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
    int64_t nsplit = (2 * h->sm_count) / ntiles; // two CTAs per SM: a single wave
    if (nsplit < 1) nsplit = 1;
    int64_t maxsplit = (nchunks + 7) / 8;
    if (nsplit > maxsplit) nsplit = maxsplit;
    if (nsplit < 1) nsplit = 1;
    int64_t kper = ((nchunks + nsplit - 1) / nsplit) * KC;      // multiple of KC (32), hence of KC9
    nsplit = (Kpad + kper - 1) / kper;
    const size_t main_doubles = (size_t)nsplit * ntiles * T9 * T9;
    const size_t e_doubles = ne > 0 ? (size_t)nsplit * ntiles * EP9 : 0;
    int rc = ensure_part(h, sizeof(double) * (main_doubles + e_doubles)); if (rc) return rc;
    double* epart = h->syrk_part + main_doubles;
    const size_t smem = sizeof(double) * ST9 * STAGE9E;
    BAX_CUDA(cudaFuncSetAttribute(k_syrk96, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    BaxTimer t(h, BAX_T_SYRK, s, ne > 0 ? 3 : 2);
    k_syrk96<<<dim3(ntiles, (unsigned)nsplit), 128, smem, s>>>(X, rows, ne, ldx, kper, Kpad, h->syrk_part, epart);
    BAX_LAUNCH_CHECK();
    k_syrk96_reduce<<<(rows * rows + 255) / 256, 256, 0, s>>>(h->syrk_part, rows, ntiles, (int)nsplit, out, ldo);
    BAX_LAUNCH_CHECK();
    if (ne > 0) {
      k_syrk96_reduce_e<<<((rows + ne) * ne + 255) / 256, 256, 0, s>>>(epart, rows, ne, ntiles, (int)nsplit, out, ldo);
      BAX_LAUNCH_CHECK();
    }
    return BAX_OK;
  }
  rows += ne;                                    // small systems: one 64 x 64 tile covers everything
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
