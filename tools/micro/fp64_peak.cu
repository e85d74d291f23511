// FP64 peak micro-benchmark: DFMA (vector pipe) and DMMA (mma.sync.m8n8k4.f64) throughput on this GPU.
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void k_dfma(double* out, int iters, double a, double b) {
  double acc[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) acc[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void k_dmma(double* out, int iters, double a, double b) {
  double c[ILP][2];
#pragma unroll
  for (int i = 0; i < ILP; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int sms = p.multiProcessorCount;
  double* out; cudaMalloc(&out, sizeof(double) * sms * 8 * 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int threads : {128, 256, 512, 1024}) {
    for (int bps : {1, 2}) {
      if (threads * bps > 2048) continue;
      k_dfma<16><<<sms * bps, threads>>>(out, 100, 1.0000001, 1e-9);
      cudaDeviceSynchronize();
      cudaEventRecord(e0);
      k_dfma<16><<<sms * bps, threads>>>(out, iters, 1.0000001, 1e-9);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double fl = 2.0 * 16 * iters * (double)threads * bps * sms;
      printf("DFMA threads=%4d blocks/SM=%d : %.2f TFLOP/s  (%.1f FMA/clk/SM at %d MHz nominal)\n", threads, bps, fl / ms * 1e-9,
             fl / 2 / (ms * 1e-3) / sms / (p.clockRate * 1e3), p.clockRate / 1000);
    }
  }
  for (int threads : {128, 256, 512, 1024}) {
    k_dmma<8><<<sms, threads>>>(out, 100, 1.0000001, 1e-9);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k_dmma<8><<<sms, threads>>>(out, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double fl = 2.0 * 8 * 8 * 4 * 8 * iters * (double)(threads / 32) * sms;   // m8n8k4 = 256 FMA per warp instr
    printf("DMMA threads=%4d : %.2f TFLOP/s\n", threads, fl / ms * 1e-9);
  }
  printf("%s, %d SMs\n", p.name, sms);
  return 0;
}
