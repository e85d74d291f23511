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

__global__ void k_syrk_reduce(const double* __restrict__ part, int rows, int ntiles, int nsplit, double* __restrict__ out) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * rows) return;
  int i = idx / rows, j = idx % rows;
  int r = i >= j ? i : j, c = i >= j ? j : i;
  int bi = r / ST, bj = c / ST;
  int t = bi * (bi + 1) / 2 + bj;
  size_t off = (size_t)t * (ST * ST) + (size_t)(r % ST) * ST + (c % ST);
  double acc = 0.0;
  for (int z = 0; z < nsplit; ++z) acc += part[(size_t)z * ntiles * (ST * ST) + off];
  out[idx] = acc;
}

int bax_syrk(bax_handle* h, const double* X, int rows, int64_t K, int64_t ldx, double* out, cudaStream_t s) {
  if (ldx % KC != 0) { bax_set_error("bax_syrk: ldx must be a multiple of %d", KC); return BAX_E_ARG; }
  const int nt = (rows + ST - 1) / ST;
  const int ntiles = nt * (nt + 1) / 2;
  const int64_t Kpad = ((K + KC - 1) / KC) * KC;
  const int64_t nchunks = Kpad / KC;
  int64_t nsplit = (4 * (int64_t)h->sm_count + ntiles - 1) / ntiles;
  int64_t maxsplit = (nchunks + 7) / 8;          // at least 8 chunks (256 columns) per CTA
  if (nsplit > maxsplit) nsplit = maxsplit;
  if (nsplit < 1) nsplit = 1;
  int64_t kper = ((nchunks + nsplit - 1) / nsplit) * KC;
  nsplit = (Kpad + kper - 1) / kper;
  size_t need = sizeof(double) * (size_t)nsplit * ntiles * ST * ST;
  if (need > h->syrk_part_bytes) {
    if (h->syrk_part) cudaFree(h->syrk_part);
    h->syrk_part = nullptr; h->syrk_part_bytes = 0;
    BAX_CUDA(cudaMalloc(&h->syrk_part, need));
    h->syrk_part_bytes = need;
  }
  const size_t smem = sizeof(double) * 2 * 2 * ST * LDT;
  BAX_CUDA(cudaFuncSetAttribute(k_syrk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  {
    BaxTimer t(h, BAX_T_SYRK, s, 2);
    k_syrk<<<dim3(ntiles, (unsigned)nsplit), 256, smem, s>>>(X, rows, ldx, kper, Kpad, h->syrk_part);
    BAX_LAUNCH_CHECK();
    k_syrk_reduce<<<(rows * rows + 255) / 256, 256, 0, s>>>(h->syrk_part, rows, ntiles, (int)nsplit, out);
    BAX_LAUNCH_CHECK();
  }
  return BAX_OK;
}
