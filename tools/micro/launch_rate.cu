// How many small kernels per second can one process feed to a B200 from T host threads, one stream each —
// as plain launches and as replays of a captured graph of NODES dependent kernels?  (Decides whether the cfg5 sweep,
// ~190 small launches per problem from 12 worker threads, is bound by the launch path.)
//   nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a launch_rate.cu -o launch_rate && ./launch_rate
#include <cuda_runtime.h>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <thread>
#include <vector>

__global__ void k_tiny(double* p, double a) { if (threadIdx.x == 0 && blockIdx.x == 0) p[0] = p[0] * a + 1.0; }

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main() {
  const int NODES = 20, REPS = 2000;
  for (int T : {1, 2, 4, 8, 12, 16}) {
    for (int mode = 0; mode < 2; ++mode) {
      std::vector<std::thread> th;
      std::vector<double> secs(T);
      std::atomic<int> ready(0);
      for (int t = 0; t < T; ++t) th.emplace_back([&, t]() {
        cudaSetDevice(0);
        cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
        double* d; cudaMalloc(&d, 8); cudaMemsetAsync(d, 0, 8, s);
        cudaGraphExec_t ge = nullptr;
        if (mode == 1) {
          cudaGraph_t g;
          cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
          for (int i = 0; i < NODES; ++i) k_tiny<<<4, 128, 0, s>>>(d, 0.5);
          cudaStreamEndCapture(s, &g);
          cudaGraphInstantiate(&ge, g, 0);
        }
        for (int i = 0; i < 50; ++i) k_tiny<<<4, 128, 0, s>>>(d, 0.5);
        cudaStreamSynchronize(s);
        ready.fetch_add(1); while (ready.load() < T) {}
        const double t0 = now();
        for (int r = 0; r < REPS; ++r) {
          if (mode == 0) for (int i = 0; i < NODES; ++i) k_tiny<<<4, 128, 0, s>>>(d, 0.5);
          else cudaGraphLaunch(ge, s);
          if ((r & 7) == 7) cudaStreamSynchronize(s);          // a read-back every 8 sequences, as the solver has them
        }
        cudaStreamSynchronize(s);
        secs[t] = now() - t0;
      });
      for (auto& x : th) x.join();
      double mx = 0; for (double x : secs) mx = x > mx ? x : mx;
      printf("threads %2d  %s  %.3e kernels/s  (%.2f us per kernel per thread)\n", T, mode ? "graph x20 " : "launches  ",
             (double)T * REPS * NODES / mx, mx / (REPS * NODES) * 1e6);
    }
  }
  return 0;
}
