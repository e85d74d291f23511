// HBM ceiling for the traffic mix of the residual/Jacobian kernel (24 B read : 160 B written per observation),
// beside a 1:1 copy and a pure fill, with plain 16-byte LSU accesses.  Prints GB/s (read + written bytes).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hbm_mix hbm_mix.cu && ./hbm_mix
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_copy(const double2* __restrict__ a, double2* __restrict__ b, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = a[i];
}
__global__ void k_fill(double2* __restrict__ b, size_t n, double v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = make_double2(v, v);
}
// W 16-byte words written per 16-byte word read (W = 7 is the k_rj mix: 24+ B in, 160 B out)
template <int W>
__global__ void k_mix(const double2* __restrict__ a, double2* __restrict__ b, size_t n /*words read*/) {
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x; i0 < n; i0 += (size_t)gridDim.x * blockDim.x) {
    size_t i = i0 + threadIdx.x;
    if (i < n) {
      double2 v = a[i];
#pragma unroll
      for (int w = 0; w < W; ++w) b[(i0 * W) + (size_t)w * blockDim.x + threadIdx.x] = make_double2(v.x + w, v.y);
    }
  }
}

template <class F>
static float time_ms(F f, int reps) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) f();
  cudaEventRecord(e0);
  for (int i = 0; i < reps; ++i) f();
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms / reps;
}

int main() {
  const size_t nw = (size_t)1 << 27;          // 2 GiB of 16-byte words written
  double2 *a, *b;
  cudaMalloc(&a, nw * 16); cudaMalloc(&b, nw * 16);
  cudaMemset(a, 0, nw * 16); cudaMemset(b, 0, nw * 16);
  int sm = 148; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
  for (int mult : {4, 8, 16, 32}) {
    int g = sm * mult;
    float t1 = time_ms([&] { k_copy<<<g, 256>>>(a, b, nw); }, 10);
    float t2 = time_ms([&] { k_fill<<<g, 256>>>(b, nw, 1.0); }, 10);
    float t3 = time_ms([&] { k_mix<7><<<g, 256>>>(a, b, nw / 7); }, 10);
    float t4 = time_ms([&] { cudaMemsetAsync(b, 0, nw * 16); }, 10);
    float t5 = time_ms([&] { cudaMemcpyAsync(b, a, nw * 16, cudaMemcpyDeviceToDevice); }, 10);
    printf("{\"grid\": %d, \"copy_GBps\": %.0f, \"fill_GBps\": %.0f, \"mix_1r7w_GBps\": %.0f, \"memset_GBps\": %.0f, \"memcpy_GBps\": %.0f}\n",
           g, 2.0 * nw * 16 / t1 / 1e6, nw * 16 / t2 / 1e6, (nw / 7) * 8.0 * 16 / t3 / 1e6, nw * 16 / t4 / 1e6, 2.0 * nw * 16 / t5 / 1e6);
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
