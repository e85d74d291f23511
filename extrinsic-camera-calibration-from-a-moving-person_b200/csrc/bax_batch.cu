// Sweep over many independent, identically shaped BA problems (BASELINE configs[4]: 4096 four-camera scenes;
// reference: run_all.sh calling ba.py once per sequence, ba.py:157-210 each) behind ONE call across the C ABI.
//
// Round 1 drove such a sweep from 12 Python worker threads: the device work of a problem (~40 launches and 2-3 scalar
// read-backs per trust-region iteration) already ran inside single C calls with the GIL released, but everything
// around it — tensor creation, the uploads, to_theta / from_theta, result dictionaries — is interpreter time under the
// GIL, ~1 ms per problem, which capped the sweep at ~800 problems/s whatever the number of streams.  Here the workers
// are native threads: each owns a CUDA stream, a libbax handle, pinned staging and device buffers sized for one
// problem, pulls problem indices from an atomic counter and runs, per problem, exactly the sequence ba_main runs
// (bax_problem_set -> bax_triangulate -> theta0 = [rvec | t | X] -> bax_trf_begin -> bax_trf_solve), so the results
// are bit-identical to single ba_main calls.  Replicas only: problems share nothing, there is no collective.
#include "bax_internal.cuh"
#include <atomic>
#include <mutex>
#include <chrono>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Worker {
  cudaStream_t s = nullptr;
  bax_handle* h = nullptr;
  char* pin = nullptr;       // pinned staging: inputs of one problem, then its theta
  char* dev = nullptr;       // device mirror of the inputs
  double *x = nullptr, *x_new = nullptr, *step = nullptr, *X0 = nullptr;
  int device = -1;
  size_t pin_bytes = 0, dev_bytes = 0, n = 0, p3 = 0;   // capacities
  int rc = 0;
  std::string err;
};

// Workers outlive a call: creating twelve of them (pinned staging, device buffers, handle, stream) costs ~60 ms, a
// tenth of a 512-problem sweep.  They are parked here and handed to the next call whose problems fit.
std::mutex g_pool_mu;
std::vector<Worker> g_pool;

void worker_free(Worker& w) {
  if (w.device >= 0) cudaSetDevice(w.device);
  if (w.s) cudaStreamSynchronize(w.s);
  if (w.h) bax_destroy(w.h);
  if (w.X0) cudaFree(w.X0);
  if (w.step) cudaFree(w.step);
  if (w.x_new) cudaFree(w.x_new);
  if (w.x) cudaFree(w.x);
  if (w.dev) cudaFree(w.dev);
  if (w.pin) cudaFreeHost(w.pin);
  if (w.s) cudaStreamDestroy(w.s);
  w = Worker();
}

struct Layout {              // byte offsets of one problem's inputs inside the staging block
  size_t K, y, s, p3, s3, R, t, rv, total;
};

}  // namespace

extern "C" int bax_batch_solve(const bax_batch_args* a) {
  if (!a || a->n_problems < 0 || a->C < 1 || a->N < 1 || a->J < 1 || !a->K || !a->y2d || !a->s2d || !a->R0 || !a->t0 ||
      !a->rvec0 || !a->theta_out || !a->info_out || (a->B > 0 && !a->bones)) {
    bax_set_error("bax_batch_solve: bad argument"); return BAX_E_ARG;
  }
  if (a->n_problems == 0) return BAX_OK;
  const int C = a->C, N = a->N, J = a->J, J3 = a->J3, B = a->B;
  const int64_t P = (int64_t)N * J, n = 6 * (int64_t)C + 3 * P;
  const bool has3 = a->p3d && a->s3d;
  Layout L;
  size_t o = 0;
  auto take = [&](size_t doubles) { size_t at = o; o += ((doubles * sizeof(double) + 255) / 256) * 256; return at; };
  L.K = take(9 * (size_t)C); L.y = take(2 * (size_t)C * P); L.s = take((size_t)C * P);
  L.p3 = take(has3 ? 3 * (size_t)C * N * J3 : 0); L.s3 = take(has3 ? (size_t)C * N * J3 : 0);
  L.R = take(9 * (size_t)C); L.t = take(3 * (size_t)C); L.rv = take(3 * (size_t)C);
  L.total = o;
  const size_t pin_bytes = L.total > (size_t)n * sizeof(double) ? L.total : (size_t)n * sizeof(double);
  int nw = a->n_workers < 1 ? 1 : a->n_workers;
  if (nw > a->n_problems) nw = a->n_problems;
  std::vector<Worker> W(nw);
  {  // parked workers of earlier calls that fit this problem shape
    std::lock_guard<std::mutex> lk(g_pool_mu);
    int got = 0;
    for (size_t i = 0; i < g_pool.size() && got < nw;) {
      Worker& c = g_pool[i];
      if (c.device == a->device && c.pin_bytes >= pin_bytes && c.dev_bytes >= L.total && c.n >= (size_t)n && c.p3 >= (size_t)(3 * P)) {
        W[got++] = c; g_pool.erase(g_pool.begin() + i);
      } else ++i;
    }
  }
  std::atomic<int> next(0);
  std::atomic<int> failed(0);
  // spinning host waits are the faster ones as long as every worker thread has a core; the caller asks for sleeping
  // waits when it has to oversubscribe the host
  const bool block = a->blocking_sync != 0;
  const bool prof = getenv("BAX_BATCH_PROFILE") != nullptr;       // per-phase host wall time, summed over workers
  std::atomic<long long> t_ns[6];
  for (auto& t : t_ns) t.store(0);
  auto now = []() { return std::chrono::steady_clock::now(); };
  auto add = [&](int i, std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
    t_ns[i].fetch_add(std::chrono::duration_cast<std::chrono::nanoseconds>(b - a).count());
  };

  auto run = [&](int wi) {
    Worker& w = W[wi];
    auto fail = [&](int rc, const char* what) {
      w.rc = rc; w.err = std::string(what) + ": " + bax_last_error(); failed.store(1);
    };
    if (cudaSetDevice(a->device) != cudaSuccess) { w.rc = BAX_E_CUDA; w.err = "cudaSetDevice"; failed.store(1); return; }
    int rc = 0;
    w.rc = 0; w.err.clear();
    if (!w.h) {                                   // a fresh worker (none parked that fits)
      bool ok = cudaStreamCreateWithFlags(&w.s, cudaStreamNonBlocking) == cudaSuccess &&
                cudaHostAlloc((void**)&w.pin, pin_bytes, cudaHostAllocDefault) == cudaSuccess &&
                cudaMalloc((void**)&w.dev, L.total) == cudaSuccess &&
                cudaMalloc((void**)&w.x, sizeof(double) * n) == cudaSuccess &&
                cudaMalloc((void**)&w.x_new, sizeof(double) * n) == cudaSuccess &&
                cudaMalloc((void**)&w.step, sizeof(double) * n) == cudaSuccess &&
                cudaMalloc((void**)&w.X0, sizeof(double) * 3 * P) == cudaSuccess;
      w.device = a->device; w.pin_bytes = pin_bytes; w.dev_bytes = L.total; w.n = (size_t)n; w.p3 = (size_t)(3 * P);
      if (!ok) { w.rc = BAX_E_CUDA; w.err = "bax_batch_solve: worker allocation failed"; failed.store(1); return; }
      rc = bax_create(a->device, &w.h);
      if (rc) { fail(rc, "bax_create"); return; }
    }
    bax_set_blocking_sync(w.h, block ? 1 : 0);
    for (;;) {
      if (failed.load()) return;
      const int k = next.fetch_add(1);
      if (k >= a->n_problems) return;
      double* info = a->info_out + (size_t)k * 8;
      for (int e = 0; e < 8; ++e) info[e] = 0.0;
      // ---- stage and upload this problem's inputs
      auto T0 = now();
      memcpy(w.pin + L.K, a->K[k], sizeof(double) * 9 * C);
      memcpy(w.pin + L.y, a->y2d[k], sizeof(double) * 2 * C * P);
      memcpy(w.pin + L.s, a->s2d[k], sizeof(double) * C * P);
      if (has3) {
        memcpy(w.pin + L.p3, a->p3d[k], sizeof(double) * 3 * C * N * J3);
        memcpy(w.pin + L.s3, a->s3d[k], sizeof(double) * C * N * J3);
      }
      memcpy(w.pin + L.R, a->R0[k], sizeof(double) * 9 * C);
      memcpy(w.pin + L.t, a->t0[k], sizeof(double) * 3 * C);
      memcpy(w.pin + L.rv, a->rvec0[k], sizeof(double) * 3 * C);
      if (cudaMemcpyAsync(w.dev, w.pin, L.total, cudaMemcpyHostToDevice, w.s) != cudaSuccess) { w.rc = BAX_E_CUDA; w.err = "upload"; failed.store(1); return; }
      const double* dK = (const double*)(w.dev + L.K); const double* dy = (const double*)(w.dev + L.y);
      const double* ds = (const double*)(w.dev + L.s);
      const double* dp3 = has3 ? (const double*)(w.dev + L.p3) : nullptr;
      const double* ds3 = has3 ? (const double*)(w.dev + L.s3) : nullptr;
      const double* dR = (const double*)(w.dev + L.R); const double* dt = (const double*)(w.dev + L.t);
      const double* drv = (const double*)(w.dev + L.rv);
      auto T1 = now();
      rc = bax_problem_set(w.h, C, N, J, has3 ? J3 : J, B, dK, dy, ds, dp3, ds3, a->bones, a->lambda1, a->lambda2, w.s);
      if (rc) { fail(rc, "bax_problem_set"); return; }
      auto T2 = now();
      int32_t tri = 0;
      rc = bax_triangulate(w.h, dR, dt, w.X0, &tri, w.s);
      if (rc) { fail(rc, "bax_triangulate"); return; }
      info[7] = (double)tri;
      auto T3 = now();
      if (tri == 1) { info[0] = -10.0; continue; }                   // "not enough constraint" (util.py:320-321)
      // theta0 = [rvec | t | X]  (ba.py:18-20)
      cudaMemcpyAsync(w.x, drv, sizeof(double) * 3 * C, cudaMemcpyDeviceToDevice, w.s);
      cudaMemcpyAsync(w.x + 3 * C, dt, sizeof(double) * 3 * C, cudaMemcpyDeviceToDevice, w.s);
      cudaMemcpyAsync(w.x + 6 * C, w.X0, sizeof(double) * 3 * P, cudaMemcpyDeviceToDevice, w.s);
      bax_trf_state st;
      memset(&st, 0, sizeof(st));
      st.ftol = a->ftol; st.xtol = a->xtol; st.gtol = a->gtol; st.max_nfev = a->max_nfev;
      st.x = w.x; st.x_new = w.x_new; st.step = w.step;
      rc = bax_trf_begin(w.h, &st, w.s);
      if (rc == BAX_E_STATE && !(st.cost == st.cost && st.cost - st.cost == 0.0)) { info[0] = -11.0; continue; }   // non-finite start
      if (rc) { fail(rc, "bax_trf_begin"); return; }
      auto T4 = now();
      rc = bax_trf_solve(w.h, &st, w.s);
      if (rc) { fail(rc, "bax_trf_solve"); return; }
      auto T5 = now();
      if (cudaMemcpyAsync(w.pin, st.x, sizeof(double) * n, cudaMemcpyDeviceToHost, w.s) != cudaSuccess ||
          bax_stream_wait(w.h, w.s) != cudaSuccess) { w.rc = BAX_E_CUDA; w.err = "download"; failed.store(1); return; }
      memcpy(a->theta_out + (size_t)k * n, w.pin, sizeof(double) * n);
      info[0] = (double)st.status; info[1] = st.cost; info[2] = st.initial_cost; info[3] = (double)st.nfev;
      info[4] = (double)st.njev; info[5] = (double)st.iteration; info[6] = st.g_inf;
      if (prof) { auto T6 = now(); add(0, T0, T1); add(1, T1, T2); add(2, T2, T3); add(3, T3, T4); add(4, T4, T5); add(5, T5, T6); }
    }
  };

  std::vector<std::thread> th;
  for (int i = 1; i < nw; ++i) th.emplace_back(run, i);
  run(0);
  for (auto& t : th) t.join();
  if (prof)
    fprintf(stderr, "bax_batch_solve: %d problems, %d workers; host ms per problem: stage+upload %.3f, problem_set %.3f, "
            "triangulate %.3f, theta0+begin %.3f, solve %.3f, download %.3f\n", a->n_problems, nw,
            t_ns[0] * 1e-6 / a->n_problems, t_ns[1] * 1e-6 / a->n_problems, t_ns[2] * 1e-6 / a->n_problems,
            t_ns[3] * 1e-6 / a->n_problems, t_ns[4] * 1e-6 / a->n_problems, t_ns[5] * 1e-6 / a->n_problems);
  int rc_all = BAX_OK;
  std::string msg;
  for (auto& w : W) {
    if (w.rc && !rc_all) { rc_all = w.rc; msg = w.err; }
    if (w.s) cudaStreamSynchronize(w.s);
  }
  {
    std::lock_guard<std::mutex> lk(g_pool_mu);
    for (auto& w : W) {
      if (w.rc || !w.h || g_pool.size() >= 64) worker_free(w);     // a failed worker is not trusted again
      else g_pool.push_back(w);
    }
  }
  if (rc_all) { bax_set_error("bax_batch_solve: %s", msg.c_str()); return rc_all; }
  return BAX_OK;
}

// Free the parked workers (streams, handles, pinned and device buffers) of earlier bax_batch_solve calls.
extern "C" int bax_batch_release(void) {
  std::lock_guard<std::mutex> lk(g_pool_mu);
  for (auto& w : g_pool) worker_free(w);
  g_pool.clear();
  return BAX_OK;
}
