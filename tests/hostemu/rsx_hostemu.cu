// TEST INFRASTRUCTURE ONLY: runs the __host__ __device__ numerics of csrc/rsx_*.cuh on the CPU (one "thread",
// syncs are no-ops) so the `-m "not gpu"` tests can check the device math against the oracle without a GPU.
// Never linked into libbax.so, never imported by the package.
#define RSX_HOST_STATS 1
#include "../../extrinsic-camera-calibration-from-a-moving-person_b200/csrc/rsx_fit.cuh"
#include <stdlib.h>
#include <vector>

extern "C" int emu_jacobi(int n, double* A /*[n][n] in, diag out*/, double* V /*[n][n]*/) {
  const int ld = n | 1;
  std::vector<double> a((size_t)n * ld), v((size_t)n * ld), cs(2 * (n / 2 + 1));
  int flag[2];
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) a[i * ld + j] = A[i * n + j];
  CtaCtx g{0, 1};
  int sweeps = cta_jacobi(g, a.data(), v.data(), n, ld, cs.data(), flag, 60);
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) { A[i * n + j] = a[i * ld + j]; V[i * n + j] = v[i * ld + j]; }
  return sweeps;
}
extern "C" void emu_eig4(const double* A16, double* x4) {
  double a[4][4], x[4];
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) a[i][j] = A16[4 * i + j];
  eig4_smallest(a, x);
  for (int i = 0; i < 4; ++i) x4[i] = x[i];
}
extern "C" void emu_eig4_fast(const double* A16, double* x4) {
  double m[10], x[4];
  int e = 0;
  for (int i = 0; i < 4; ++i) for (int j = i; j < 4; ++j) m[e++] = A16[4 * i + j];
  eig4_smallest_fast(m, x);
  for (int i = 0; i < 4; ++i) x4[i] = x[i];
}
extern "C" void emu_eig4_mode(const double* A16, double* x4, int mode, int norm) {
  double m[10], x[4];
  int e = 0;
  for (int i = 0; i < 4; ++i) for (int j = i; j < 4; ++j) m[e++] = A16[4 * i + j];
  if (mode == 0) eig4_smallest_invit(m, x);
  else if (norm) eig4_smallest_rqi<true>(m, x);
  else eig4_smallest_rqi<false>(m, x);
  for (int i = 0; i < 4; ++i) x4[i] = x[i];
}
extern "C" void emu_rqi_stats(long long* out4, int reset) {
  for (int i = 0; i < 4; ++i) { out4[i] = rsx_host_stats[i]; if (reset) rsx_host_stats[i] = 0; }
}
extern "C" void emu_fit(int C, long long N, long long M, const double* v, const double* n, const int* sv, const int* sm,
                        double* Rt_out) {
  const int nu = 3 * (RSX_S + C), ld = nu | 1;
  std::vector<double> buf(fit_ws_small_doubles(C) + 2 * (size_t)nu * ld);
  FitWs ws; fit_ws_carve(buf.data(), nullptr, C, ws);
  CtaCtx g{0, 1};
  fit_hypothesis(g, C, N, M, v, n, sv, sm, ws, Rt_out);
}
// mode 0: Jacobi only; 1: arrow subspace iteration first.  Returns the sweep count of the arrow path (-1: fell back).
extern "C" int emu_fit_mode(int C, long long N, long long M, const double* v, const double* n, const int* sv, const int* sm,
                            double* Rt_out, int mode) {
  const int nu = 3 * (RSX_S + C), ld = nu | 1;
  std::vector<double> buf(fit_ws_small_doubles(C) + 2 * (size_t)nu * ld);
  FitWs ws; fit_ws_carve(buf.data(), nullptr, C, ws);
  ws.iflag[8] = -2;
  CtaCtx g{0, 1};
  ws.iflag[9] = -1;
  fit_hypothesis(g, C, N, M, v, n, sv, sm, ws, Rt_out, mode != 0);
  if (mode == 2) return ws.iflag[9];          // iterations of the top-3 orthogonal iteration (-1: not run / failed)
  return ws.iflag[8];
}
extern "C" void emu_score_sum(int C, long long N, long long M, const double* v, const double* n, const double* Rt,
                              double* parts2, int mode) {
  CtaCtx g{0, 1};
  parts2[0] = mode ? score_points_sum<1>(g, C, M, n, Rt) : score_points_sum<0>(g, C, M, n, Rt);
  std::vector<double> Gc(9 * C);
  score_dirs_prepare(g, C, Rt, Gc.data());
  parts2[1] = mode ? score_dirs_sum(g, C, N, v, Gc.data()) : score_dirs(g, C, N, v, Rt, nullptr, 0.0, nullptr);
}
extern "C" void emu_score(int C, long long N, long long M, const double* v, const double* n, const double* Rt,
                          double* parts2, double* e_out, double* e3d_out, const double* K00, double th2d, double th3d,
                          unsigned char* mask2d, unsigned char* mask3d) {
  CtaCtx g{0, 1};
  parts2[0] = score_points(g, C, M, n, Rt, e_out, K00, th2d, mask2d);
  parts2[1] = score_dirs(g, C, N, v, Rt, e3d_out, th3d, mask3d);
}
