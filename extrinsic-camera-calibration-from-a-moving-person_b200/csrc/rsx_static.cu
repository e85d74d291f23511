// rsx_argmin_static: the RANSAC winner search (first minimum of the scores under strict '<', calib_ransac.py:139,182)
// for callers that may not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  It has no counterpart in the reference.
//
// rsx_argmin (rsx.cu) needs an rsx_handle, whose creation allocates device and pinned memory, and it searches with one
// CTA.  This variant takes no handle and spreads the search over the GPU: one CTA per RSX_STATIC_CHUNK scores writes
// its partial winner into a file-scope device array, and the CTA that finishes last folds the partials.  The array is
// static, so every GPU that loads libbax.so holds its RSX_STATIC_SCRATCH_BYTES (65 792 B) until the process exits.
//
// What to know before calling it:
//  * nothing here calls cudaMalloc / cudaFree / cudaHostAlloc or any allocating library routine; the only runtime
//    calls are cudaSetDevice, cudaGetSymbolAddress (first call per device), cudaMemsetAsync and the kernel launch.
//    The first call on a device resolves the scratch address (which may load the module): make that call once before
//    entering graph capture.
//  * H is limited to RSX_STATIC_MAX_H; larger inputs are rejected with BAX_E_ARG.
//  * every call shares the one array per device, and the array keeps the previous call's contents.  The arrival
//    counter is cleared on the caller's stream at the start of every call; the partials need no clearing, since a call
//    reads only the entries its own CTAs wrote.  Two calls that overlap on the same GPU (different streams, or
//    different threads) RACE on the array: order them on one stream or with events.
#include "bax_internal.cuh"
#include "../../include/rsx.h"
#include <math.h>
#include <mutex>

#define RSX_STATIC_THREADS 256
#define RSX_STATIC_PER_THREAD (RSX_STATIC_CHUNK / RSX_STATIC_THREADS)
#define RSX_STATIC_MAX_CTAS (RSX_STATIC_MAX_H / RSX_STATIC_CHUNK)
#define RSX_STATIC_MAX_DEVICES 64

static_assert(RSX_STATIC_CHUNK % RSX_STATIC_THREADS == 0, "a CTA's chunk is a whole number of scores per thread");
static_assert(RSX_STATIC_SCRATCH_BYTES == 256 + 16 * (long long)RSX_STATIC_MAX_CTAS, "header and kernel disagree");

// [0, 256): arrival counter (one unsigned int, padded); then RSX_STATIC_MAX_CTAS partial scores (double) and
// RSX_STATIC_MAX_CTAS partial indices (long long).  In HOST code this name is a host-side stand-in, not a device
// address: the address a kernel can use comes from cudaGetSymbolAddress (scratch_for_device below).
static __device__ __align__(256) unsigned char rsx_static_scratch[RSX_STATIC_SCRATCH_BYTES];

namespace {

// (value, index) ordering of k_rsx_argmin: lower score wins, equal scores go to the lower index, index -1 = "none yet"
__device__ __forceinline__ void take_better(double& bv, long long& bi, double ov, long long oi) {
  if (oi >= 0 && (ov < bv || (ov == bv && (bi < 0 || oi < bi)))) { bv = ov; bi = oi; }
}

__device__ __forceinline__ void cta_fold(double& bv, long long& bi, double* sv, long long* si) {
  for (int o = 16; o > 0; o >>= 1) {
    const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const long long oi = __shfl_xor_sync(0xffffffffu, bi, o);
    take_better(bv, bi, ov, oi);
  }
  __syncthreads();                                        // sv / si may still be read from an earlier fold
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0)
    for (int w = 1; w < RSX_STATIC_THREADS / 32; ++w) take_better(bv, bi, sv[w], si[w]);
}

__global__ void __launch_bounds__(RSX_STATIC_THREADS)
k_rsx_argmin_static(const double* __restrict__ score, int64_t H, unsigned char* __restrict__ scratch,
                    double* __restrict__ out /*[2]: index, score*/) {
  __shared__ double sv[RSX_STATIC_THREADS / 32];
  __shared__ long long si[RSX_STATIC_THREADS / 32];
  __shared__ bool last;
  unsigned int* arrived = (unsigned int*)scratch;
  double* pv = (double*)(scratch + 256);
  long long* pi = (long long*)(scratch + 256 + 8 * (size_t)RSX_STATIC_MAX_CTAS);

  // NaN and +inf never pass 's < bv', as in the reference's 'if score < best' starting from np.inf
  double bv = INFINITY; long long bi = -1;
  const int64_t base = (int64_t)blockIdx.x * RSX_STATIC_CHUNK;
#pragma unroll
  for (int k = 0; k < RSX_STATIC_PER_THREAD; ++k) {
    const int64_t i = base + (int64_t)k * RSX_STATIC_THREADS + threadIdx.x;    // ascending per thread: first minimum kept
    if (i < H) { const double s = __ldg(score + i); if (s < bv) { bv = s; bi = i; } }
  }
  cta_fold(bv, bi, sv, si);
  if (threadIdx.x == 0) {
    pv[blockIdx.x] = bv; pi[blockIdx.x] = bi;
    __threadfence();                                      // the partial is visible before the arrival is counted
    last = atomicAdd(arrived, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  bv = INFINITY; bi = -1;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += RSX_STATIC_THREADS)       // ascending CTAs = ascending indices
    take_better(bv, bi, __ldcg(pv + b), __ldcg(pi + b));
  cta_fold(bv, bi, sv, si);
  if (threadIdx.x == 0) { out[0] = (double)bi; out[1] = bv; *arrived = 0u; }
}

std::mutex g_mu;
unsigned char* g_scratch[RSX_STATIC_MAX_DEVICES];         // device address of rsx_static_scratch, per device

// once per device, with that device current
int scratch_for_device(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_scratch[device]) {
    void* p = nullptr;
    BAX_CUDA(cudaGetSymbolAddress(&p, rsx_static_scratch));
    if (!p) { bax_set_error("rsx_argmin_static: cudaGetSymbolAddress returned a null address"); return BAX_E_CUDA; }
    g_scratch[device] = (unsigned char*)p;
  }
  *out = g_scratch[device];
  return BAX_OK;
}

}  // namespace

extern "C" int rsx_argmin_static(int device, const double* score, int64_t H, double* out, void* stream) {
  if (!score || !out || H < 1) { bax_set_error("rsx_argmin_static: bad argument"); return BAX_E_ARG; }
  if (H > (int64_t)RSX_STATIC_MAX_H) {
    bax_set_error("rsx_argmin_static: H = %lld exceeds RSX_STATIC_MAX_H = %lld (the static scratch holds one partial "
                  "per %d scores); use rsx_argmin", (long long)H, (long long)RSX_STATIC_MAX_H, (int)RSX_STATIC_CHUNK);
    return BAX_E_ARG;
  }
  if (device < 0 || device >= RSX_STATIC_MAX_DEVICES) { bax_set_error("rsx_argmin_static: bad device %d", device); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(device));
  unsigned char* scratch = nullptr;
  const int rc = scratch_for_device(device, &scratch);
  if (rc) return rc;
  // the array keeps what the last call left: the arrival counter must start at zero (a call that was aborted, or a
  // fault in an overlapping call, may have left it anywhere)
  BAX_CUDA(cudaMemsetAsync(scratch, 0, 256, s));
  const int grid = (int)((H + RSX_STATIC_CHUNK - 1) / RSX_STATIC_CHUNK);
  k_rsx_argmin_static<<<grid, RSX_STATIC_THREADS, 0, s>>>(score, H, scratch, out);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}
