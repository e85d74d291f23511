// RANSAC linear calibration on the GPU: one CTA fits one minimal-sample hypothesis, one CTA scores one model
// against every ray and bone direction (reference calib_ransac.py:134-209).
#include "rsx_handle.cuh"
#include "rsx_fit.cuh"
#include <stdlib.h>
#include <string.h>

#define FIT_THREADS 128
#define SCORE_THREADS 256

// ------------------------------------------------------------------ kernels
__global__ void __launch_bounds__(FIT_THREADS)
k_rsx_fit(int C, int64_t N, int64_t M, const double* __restrict__ v, const double* __restrict__ n,
          const int32_t* __restrict__ samp_v, const int32_t* __restrict__ samp_m, int64_t H,
          double* __restrict__ Rt, double* __restrict__ hv_global /*null: H, V in shared memory*/) {
  extern __shared__ __align__(16) double sm[];
  const int nu = 3 * (RSX_S + C), ld = nu | 1;
  FitWs ws;
  fit_ws_carve(sm, hv_global ? hv_global + (size_t)blockIdx.x * 2 * nu * ld : nullptr, C, ws);
  CtaCtx g{(int)threadIdx.x, (int)blockDim.x};
  for (int64_t h = blockIdx.x; h < H; h += gridDim.x)
    fit_hypothesis(g, C, N, M, v, n, samp_v + h * RSX_S, samp_m + h * RSX_S, ws, Rt + (size_t)h * C * 12);
}

// One WARP per hypothesis (few cameras): the arrow-structured subspace iteration of rsx_arrow.cuh needs only warp-wide
// steps, so FITW_WARPS hypotheses share a CTA, each with its own shared-memory workspace, and __syncwarp replaces the
// CTA barrier.  A hypothesis whose iteration does not settle is finished by the same warp with the Jacobi solver.
#define FITW_WARPS 2
// WARPS hypotheses per CTA.  LOCKSTEP: the CTA takes WARPS consecutive hypotheses at a time and meets at a barrier after
// each batch, so that the warps of an SM walk through the same phase of the (large) code together: fewer
// instruction-fetch stalls than the free-running version, at the price of waiting for the batch's slowest fit
// (default where the compact layout fits 16 warps; env RSX_FIT_LOCKSTEP=0 gives the free-running kernel).
template <bool DEFER_ONLY, int WARPS, bool LOCKSTEP>
__global__ void __launch_bounds__(32 * WARPS, LOCKSTEP ? 1 : 16 / WARPS)
k_rsx_fit_warp(int C, int64_t N, int64_t M, const double* __restrict__ v, const double* __restrict__ n,
               const int32_t* __restrict__ samp_v, const int32_t* __restrict__ samp_m, int64_t H,
               double* __restrict__ Rt, unsigned long long* __restrict__ stats /*[4]: fall-backs, sweeps, work counter, deferred*/,
               int32_t* __restrict__ deferred /*null: the warp runs the Jacobi fall-back itself*/,
               double* h_global /*[warps of the grid][nu][ld] or null: H in shared memory too*/,
               int compact /*compact shared-memory layout (needs deferred and h_global)*/,
               uint8_t* __restrict__ deferred_flag /*[H] or null: set to 1 for the hypotheses left to the second pass*/) {
  extern __shared__ __align__(16) double sm[];
  const int nu = 3 * (RSX_S + C), ld = nu | 1;
  // The dense normal matrix H is written once and read once (the arrow solver copies its blocks out), so it lives in
  // an L2-resident global slab per warp; the solver's workspace (V) stays in shared memory.  That takes a warp's
  // shared memory from 36 to 22 KB at 4 cameras: 10 resident warps per SM instead of 6 for a latency-bound kernel.
  const size_t per_warp = compact ? ((fit_ws_compact_small_doubles(C) + fit_ws_compact_v_doubles(C) + 1) & ~(size_t)1)
                                  : ((fit_ws_small_doubles(C) + (h_global ? 1 : 2) * (size_t)nu * ld + 1) & ~(size_t)1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  FitWs ws;
  fit_ws_carve(sm + warp * per_warp, nullptr, C, ws,
               h_global ? h_global + ((size_t)blockIdx.x * WARPS + warp) * (size_t)nu * ld : nullptr, compact != 0);
  WarpCtx g{lane, 32};
  __shared__ long long s_base;
  // hypotheses are handed out one at a time from a global counter: a fit's cost varies (sweep count, the rare Jacobi
  // fall-back is ~50x a normal fit), so a fixed stride leaves most SMs idle while the slowest warp finishes
  for (;;) {
    long long h = 0;
    if (LOCKSTEP) {
      if (threadIdx.x == 0) s_base = (long long)atomicAdd(stats + 2, (unsigned long long)WARPS);
      __syncthreads();
      const long long base = s_base;
      if (base >= H) break;                                   // uniform over the CTA
      h = base + warp;
    } else {
      if (lane == 0) h = (long long)atomicAdd(stats + 2, 1ULL);
      h = __shfl_sync(0xffffffffu, h, 0);
      if (h >= H) break;
    }
    if (h < H) {
      const bool done = fit_hypothesis_t<DEFER_ONLY>(g, C, N, M, v, n, samp_v + h * RSX_S, samp_m + h * RSX_S, ws,
                                                     Rt + (size_t)h * C * 12, true, deferred != nullptr);
      if (lane == 0) {
        const int sw = ws.iflag[8];
        if (sw < 0) atomicAdd(stats, 1ULL); else atomicAdd(stats + 1, (unsigned long long)sw);
        if (!done) {
          deferred[atomicAdd(stats + 3, 1ULL)] = (int32_t)h;             // second pass: k_rsx_fit_deferred
          if (deferred_flag) deferred_flag[h] = 1;
        }
      }
    }
    __syncwarp();
    if (LOCKSTEP) __syncthreads();
  }
}

// Second pass over the hypotheses whose arrow iteration did not settle: one CTA each, the parallel cyclic Jacobi
// diagonalisation.  A warp of k_rsx_fit_warp running it inline takes ~50 normal fits, and the few warps that met one
// near the end of a chunk kept the whole GPU waiting (ncu: SMs active 62 % of the kernel's time).
__global__ void __launch_bounds__(FIT_THREADS, 4)
k_rsx_fit_deferred(int C, int64_t N, int64_t M, const double* __restrict__ v, const double* __restrict__ n,
                   const int32_t* __restrict__ samp_v, const int32_t* __restrict__ samp_m,
                   const int32_t* __restrict__ deferred, const unsigned long long* __restrict__ stats,
                   unsigned long long* __restrict__ stats_total, double* __restrict__ Rt) {
  extern __shared__ __align__(16) double sm[];
  FitWs ws;
  fit_ws_carve(sm, nullptr, C, ws);
  CtaCtx g{(int)threadIdx.x, (int)blockDim.x};
  const unsigned long long nd = stats[3];
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(stats_total, nd);      // cumulative count behind rsx_fit_stats
  for (unsigned long long i = blockIdx.x; i < nd; i += gridDim.x) {
    const int64_t h = deferred[i];
    fit_hypothesis(g, C, N, M, v, n, samp_v + h * RSX_S, samp_m + h * RSX_S, ws, Rt + (size_t)h * C * 12, false);
  }
}

// Launch shape: 256 threads x 3 CTAs/SM (80 registers).  Measured alternatives at cfg2 (score pass of 100 k
// hypotheses, 83.4 ms): 256x2 and 128x4 at 118 registers 92-96 ms, 128x7 at 72 registers 84.2, 512x1 97.8; taking a
// thread's rays two at a time to interleave their chains (128 registers, 2 CTAs/SM) 92.2 -- no gain, dropped.
template <int MODE, int MINB>
__global__ void __launch_bounds__(SCORE_THREADS, MINB)
k_rsx_score(int C, int64_t N, int64_t M, const double* __restrict__ v, const double* __restrict__ n,
            const double* __restrict__ Rt, int64_t H, double* __restrict__ score, double* __restrict__ parts,
            const uint8_t* __restrict__ skip /*[H] or null: hypotheses whose model is not ready yet (second fit pass)*/,
            const int32_t* __restrict__ list /*null, or: score list[0 .. *n_list) instead of 0 .. H*/,
            const unsigned long long* __restrict__ n_list) {
  extern __shared__ __align__(16) double sm[];       // [C][12] model, 2*32 reduction scratch, [C][9] R_0 R_c^T
  double* sRt = sm;
  double* red = sm + ((12 * C + 1) & ~1);
  double* sG = red + 64;
  CtaCtx g{(int)threadIdx.x, (int)blockDim.x};
  const int64_t count = list ? (int64_t)*n_list : H;
  for (int64_t it = blockIdx.x; it < count; it += gridDim.x) {
    const int64_t h = list ? (int64_t)list[it] : it;
    if (skip && skip[h]) continue;                       // uniform over the CTA
    __syncthreads();
    for (int i = threadIdx.x; i < 12 * C; i += blockDim.x) sRt[i] = Rt[(size_t)h * C * 12 + i];
    __syncthreads();
    double acc[2];
    if (MODE == 1) { score_dirs_prepare(g, C, sRt, sG); __syncthreads(); }
    acc[0] = score_points_sum<MODE>(g, C, M, n, sRt);
    acc[1] = MODE == 1 ? score_dirs_sum(g, C, N, v, sG) : score_dirs(g, C, N, v, sRt, nullptr, 0.0, nullptr);
    block_sum<2>(acc, red);
    if (threadIdx.x == 0) {
      score[h] = sqrt(acc[0]) / (double)M + sqrt(acc[1]) / (double)N;      // calib_ransac.py:179-181
      if (parts) { parts[2 * h] = acc[0]; parts[2 * h + 1] = acc[1]; }
    }
  }
}

// Scoring with an exact bail-out (rsx_prune).  RANSAC only keeps the hypothesis with the smallest score, and the score
// sqrt(sum_rays e^2) / M + sqrt(sum_dirs ang^2) / N only grows while rays are added.  So a CTA first forms the (cheap)
// direction term, then walks the rays in pieces of PRUNE_RAYS per thread and after each piece compares the lower bound
// sqrt(partial) / M + direction term with the best complete score any CTA has published so far (atomicMin on the bits of
// a positive double, fit_stats[5]); a hypothesis whose bound is already LARGER than that cannot be the first minimum and
// is dropped (score = +inf, counted in fit_stats[6]).  Ties are never dropped (strict '>'), NaN sums never compare
// true, and a hypothesis that survives accumulates exactly the sequence k_rsx_score<1> accumulates: the winner, its
// score and everything derived from it are bit-identical to the unpruned pass; which losers are cut short depends on
// timing.  (The reference scores every hypothesis in full, calib_ransac.py:154-181; its result depends on the minimum only.)
// Measured and not kept: walking the directions piecewise alongside the rays instead of first — the bound is weaker while
// the direction term is incomplete: 33.5 against 29.7 ms per 100 k hypotheses.
#define PRUNE_RAYS 2                             // measured per 100 k hypotheses: 1 -> 29.0 ms, 2 -> 29.7, 4 -> 31.4, 8 -> 34.8 (the whole call; 78 unpruned)
template <int MINB>
__global__ void __launch_bounds__(SCORE_THREADS, MINB)
k_rsx_score_prune(int C, int64_t N, int64_t M, const double* __restrict__ v, const double* __restrict__ n,
                  const double* __restrict__ Rt, int64_t H, double* __restrict__ score,
                  const uint8_t* __restrict__ skip, const int32_t* __restrict__ list, const unsigned long long* __restrict__ n_list,
                  unsigned long long* __restrict__ best_bits, unsigned long long* __restrict__ pruned) {
  extern __shared__ __align__(16) double sm[];       // [C][12] model, 2*32 reduction scratch, [C][9] R_0 R_c^T
  double* sRt = sm;
  double* red = sm + ((12 * C + 1) & ~1);
  double* sG = red + 64;
  __shared__ double s_dirs; __shared__ int s_bail;
  CtaCtx g{(int)threadIdx.x, (int)blockDim.x};
  const int64_t count = list ? (int64_t)*n_list : H;
  const int64_t piece = (int64_t)PRUNE_RAYS * blockDim.x;
  for (int64_t it = blockIdx.x; it < count; it += gridDim.x) {
    const int64_t h = list ? (int64_t)list[it] : it;
    if (skip && skip[h]) continue;                       // uniform over the CTA
    __syncthreads();
    for (int i = threadIdx.x; i < 12 * C; i += blockDim.x) sRt[i] = Rt[(size_t)h * C * 12 + i];
    __syncthreads();
    score_dirs_prepare(g, C, sRt, sG);
    __syncthreads();
    double a1[1] = {score_dirs_sum(g, C, N, v, sG)};
    block_sum<1>(a1, red);
    if (threadIdx.x == 0) { s_dirs = a1[0]; s_bail = 0; }
    __syncthreads();
    const double dirs_total = s_dirs, dirs_term = sqrt(dirs_total) / (double)N;
    double acc0 = 0.0;
    bool bail = false;
    for (int64_t i0 = 0; i0 < M; i0 += piece) {
      const int64_t i1 = i0 + piece < M ? i0 + piece : M;
      score_points_accum<1>(g, C, M, n, sRt, i0, i1, acc0);
      if (i1 < M) {
        double p[1] = {acc0};
        block_sum<1>(p, red);                            // ends with the value in thread 0; barriers inside
        if (threadIdx.x == 0) {
          const double lb = sqrt(p[0]) / (double)M + dirs_term;
          const double best = __longlong_as_double((long long)*(volatile unsigned long long*)best_bits);
          s_bail = lb > best ? 1 : 0;
        }
        __syncthreads();
        bail = s_bail != 0;
        if (bail) break;
      }
    }
    if (bail) {
      if (threadIdx.x == 0) { score[h] = INFINITY; atomicAdd(pruned, 1ULL); }
      continue;
    }
    double a0[1] = {acc0};
    block_sum<1>(a0, red);
    if (threadIdx.x == 0) {
      const double sc = sqrt(a0[0]) / (double)M + sqrt(dirs_total) / (double)N;      // calib_ransac.py:179-181
      score[h] = sc;
      if (sc >= 0.0) atomicMin(best_bits, (unsigned long long)__double_as_longlong(sc));   // NaN: not published
    }
  }
}

// first minimum under strict '<' starting from +inf (calib_ransac.py:139,182): NaN / +inf never win
__global__ void __launch_bounds__(1024)
k_rsx_argmin(const double* __restrict__ score, int64_t H, double* __restrict__ out /*[2]: index, score*/) {
  __shared__ double sv[32]; __shared__ long long si[32];
  double bv = INFINITY; long long bi = -1;
  for (int64_t i = threadIdx.x; i < H; i += blockDim.x) {
    const double s = score[i];
    if (s < bv) { bv = s; bi = i; }                 // strided scan keeps the lowest index per thread
  }
  for (int o = 16; o > 0; o >>= 1) {
    const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const long long oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (oi >= 0 && (ov < bv || (ov == bv && (bi < 0 || oi < bi)))) { bv = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
      if (si[w] >= 0 && (sv[w] < bv || (sv[w] == bv && (bi < 0 || si[w] < bi)))) { bv = sv[w]; bi = si[w]; }
    out[0] = (double)bi; out[1] = bv;
  }
}

__global__ void __launch_bounds__(256)
k_rsx_masks(int C, int64_t N, int64_t M, const double* __restrict__ v, const double* __restrict__ n,
            const double* __restrict__ Rt, const double* __restrict__ K00, double th2d, double th3d,
            uint8_t* __restrict__ mask2d, uint8_t* __restrict__ mask3d, double* __restrict__ e_out,
            double* __restrict__ e3d_out) {
  extern __shared__ __align__(16) double sm[];       // [C][12] model, [C] focal lengths
  for (int i = threadIdx.x; i < 12 * C; i += blockDim.x) sm[i] = Rt[i];
  for (int i = threadIdx.x; i < C; i += blockDim.x) sm[12 * C + i] = K00[i];
  __syncthreads();
  CtaCtx g{(int)(blockIdx.x * blockDim.x + threadIdx.x), (int)(gridDim.x * blockDim.x)};
  score_points(g, C, M, n, sm, e_out, sm + 12 * C, th2d, mask2d);
  score_dirs(g, C, N, v, sm, e3d_out, th3d, mask3d);
}

// DLT re-triangulation of every ray set under one model (calib_linear.py:127-133 -> pycalib.calib.triangulate ==
// util.py:240-258): the eigenvector of the smallest eigenvalue of sum_c a^T a, dehomogenised unless
// np.isclose(x[3], 0)  (|x3| <= 1e-8), in which case the unit eigenvector itself is returned.  One thread per point.
__global__ void __launch_bounds__(256)
k_rsx_triangulate(int C, int64_t M, const double* __restrict__ n, const double* __restrict__ Rt, double* __restrict__ X4) {
  extern __shared__ __align__(16) double sm[];       // [C][12]
  for (int i = threadIdx.x; i < 12 * C; i += blockDim.x) sm[i] = Rt[i];
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (int64_t)gridDim.x * blockDim.x) {
    double m[10];
#pragma unroll
    for (int e = 0; e < 10; ++e) m[e] = 0.0;
    for (int c = 0; c < C; ++c) {
      const double* nn = n + ((size_t)c * M + i) * 3;
      dlt_accumulate(sm + 12 * c, nn[0], nn[1], m);
    }
    double X[4];
    eig4_smallest_fast(m, X);
    const bool at_infinity = fabs(X[3]) <= 1e-8;     // np.isclose(v[3, 0], 0): atol 1e-8 (+ rtol * 0)
    const double w = at_infinity ? 1.0 : X[3];
    double* o = X4 + (size_t)i * 4;
    o[0] = X[0] / w; o[1] = X[1] / w; o[2] = X[2] / w; o[3] = at_infinity ? X[3] : 1.0;
  }
}

__global__ void k_rsx_count(const uint8_t* __restrict__ m, int64_t n, unsigned long long* __restrict__ out) {
  unsigned long long c = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) c += m[i];
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// ------------------------------------------------------------------ handle
extern "C" int rsx_create(int device, rsx_handle** out) {
  if (!out) { bax_set_error("rsx_create: null out"); return BAX_E_ARG; }
  BAX_CUDA(cudaSetDevice(device));
  rsx_handle* h = (rsx_handle*)calloc(1, sizeof(rsx_handle));
  h->device = device;
  cudaDeviceProp prop;
  BAX_CUDA(cudaGetDeviceProperties(&prop, device));
  h->sm_count = prop.multiProcessorCount;
  BAX_CUDA(cudaEventCreate(&h->ev0));
  BAX_CUDA(cudaEventCreate(&h->ev1));
  BAX_CUDA(cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking));
  BAX_CUDA(cudaEventCreateWithFlags(&h->ev_main, cudaEventDisableTiming));
  BAX_CUDA(cudaEventCreateWithFlags(&h->ev_side, cudaEventDisableTiming));
  // default off: both the first-pass and the score kernel fill the register files, so the second pass run "under" the
  // score kernel takes SM slots from it for its whole (long) latency — 83.9 against 81.9 ms per 100 k measured
  { const char* e = getenv("RSX_OVERLAP_DEFERRED"); h->overlap_deferred = e ? atoi(e) : 0; }
  BAX_CUDA(cudaMallocHost((void**)&h->host_pinned, 64 * sizeof(double)));
  h->red_bytes = sizeof(double) * (1 << 20);
  BAX_CUDA(cudaMalloc(&h->red, h->red_bytes));
  {  // keep freed scratch of the final fit in the device pool instead of returning it to the driver at every sync
    cudaMemPool_t pool; unsigned long long keep = ~0ULL;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  { const char* e = getenv("RSX_FIT_MODE"); h->fit_mode = e ? atoi(e) : 1; }
  { const char* e = getenv("RSX_SCORE_MODE"); h->score_mode = e ? atoi(e) : 1; }   // 0: round-1 inverse iteration + per-camera divisions   // 0: CTA Jacobi kernel only
  BAX_CUDA(cudaMalloc(&h->fit_stats, 8 * sizeof(unsigned long long)));
  BAX_CUDA(cudaMemset(h->fit_stats, 0, 8 * sizeof(unsigned long long)));
  h->bx = (bax_handle*)calloc(1, sizeof(bax_handle));
  h->bx->device = device; h->bx->sm_count = h->sm_count;
  *out = h;
  return BAX_OK;
}

static void rsx_free_problem(rsx_handle* h) {
  void* ptrs[] = {h->v, h->n, h->Rt, h->fit_scratch};
  for (void* p : ptrs) if (p) cudaFree(p);
  h->v = h->n = h->Rt = h->fit_scratch = nullptr; h->cap_h = 0; h->fit_scratch_bytes = 0;
}

extern "C" int rsx_destroy(rsx_handle* h) {
  if (!h) return BAX_OK;
  cudaSetDevice(h->device);
  rsx_free_problem(h);
  if (h->bx) { if (h->bx->syrk_part) cudaFree(h->bx->syrk_part); if (h->bx->partials) cudaFree(h->bx->partials); free(h->bx); }
  cudaFree(h->red); cudaFree(h->fit_stats); if (h->deferred) cudaFree(h->deferred); cudaFreeHost(h->host_pinned);
  cudaEventDestroy(h->ev0); cudaEventDestroy(h->ev1);
  if (h->side) { cudaStreamSynchronize(h->side); cudaStreamDestroy(h->side); }
  cudaEventDestroy(h->ev_main); cudaEventDestroy(h->ev_side);
  if (h->deferred_flag) cudaFree(h->deferred_flag);
  free(h);
  return BAX_OK;
}

extern "C" int rsx_problem_set(rsx_handle* h, int C, int64_t N, int64_t M, const double* v, const double* n, void* stream) {
  if (!h || C < 2 || N < RSX_S || M < RSX_S || !v || !n) {
    bax_set_error("rsx_problem_set: needs C >= 2 cameras and at least %d directions and rays", RSX_S); return BAX_E_ARG;
  }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  rsx_free_problem(h);
  h->C = C; h->N = N; h->M = M;
  BAX_CUDA(cudaMalloc(&h->v, sizeof(double) * 3 * C * N));
  BAX_CUDA(cudaMalloc(&h->n, sizeof(double) * 3 * C * M));
  BAX_CUDA(cudaMemcpyAsync(h->v, v, sizeof(double) * 3 * C * N, cudaMemcpyDeviceToDevice, s));
  BAX_CUDA(cudaMemcpyAsync(h->n, n, sizeof(double) * 3 * C * M, cudaMemcpyDeviceToDevice, s));
  BAX_CUDA(cudaStreamSynchronize(s));
  return BAX_OK;
}

static int ensure_models(rsx_handle* h, int64_t H) {
  if (H > h->cap_h) {
    if (h->Rt) cudaFree(h->Rt);
    h->Rt = nullptr; h->cap_h = 0;
    BAX_CUDA(cudaMalloc(&h->Rt, sizeof(double) * 12 * h->C * H));
    h->cap_h = H;
  }
  return BAX_OK;
}

// overlap != null: the deferred second pass is launched on the handle's side stream (ordered after the first pass) and
// *overlap is set to 1 — the caller scores the finished hypotheses meanwhile (k_rsx_score skips h->deferred_flag) and the
// deferred ones after waiting for h->ev_side.  The second pass is one wave of CTAs at the latency of a 42 x 42 Jacobi
// diagonalisation with the SMs almost idle (ncu: 6 % warps active): run back to back it was 1.3-1.6 ms of every 20 k-
// hypothesis chunk, a third of the fit time; under the score kernel it costs nothing measurable.
int rsx_launch_fit(rsx_handle* h, const int32_t* samp_v, const int32_t* samp_m, int64_t H, double* Rt, cudaStream_t s,
                   int* overlap = nullptr) {
  if (overlap) *overlap = 0;
  h->fits_total += H;
  const int C = h->C, nu = 3 * (RSX_S + C), ld = nu | 1;
  const size_t small = fit_ws_small_doubles(C) * sizeof(double);
  const size_t hv = sizeof(double) * 2 * (size_t)nu * ld;
  {  // few cameras: one warp per hypothesis
    const bool hglob = h->fit_mode != 3;                            // env RSX_FIT_MODE=3: H in shared memory, as before
    const bool defer_ok = h->fit_mode != 2 && small + hv <= 200 * 1024;
    const bool compact = hglob && defer_ok && h->fit_mode != 5;   // env RSX_FIT_MODE=5: round-1 shared-memory layout
    const size_t per_warp = (compact ? ((fit_ws_compact_small_doubles(C) + fit_ws_compact_v_doubles(C) + 1) & ~(size_t)1)
                                     : ((fit_ws_small_doubles(C) + (hglob ? 1 : 2) * (size_t)nu * ld + 1) & ~(size_t)1)) * sizeof(double);
    static const int env_warps = getenv("RSX_FIT_WARPS") ? atoi(getenv("RSX_FIT_WARPS")) : FITW_WARPS;
    static const int env_lock = getenv("RSX_FIT_LOCKSTEP") ? atoi(getenv("RSX_FIT_LOCKSTEP")) : 0;   // lock-step batches: 17.1 against 18.2 ms per 100 k while the kernel was 15 k instructions; free-running wins again (15.4 : 15.9) now that it is 9.6 k
    const bool lockstep = env_lock != 0 && compact && per_warp * 16 <= 226 * 1024;
    const int fw = lockstep ? 16 : (env_warps == 4 ? 4 : FITW_WARPS);
    const size_t smem = per_warp * fw;
    if (h->fit_mode != 0 && (smem <= 110 * 1024 || lockstep) && arrow_ws_doubles(C, 32) <= (size_t)nu * ld) {
      const int per_sm = (int)((226 * 1024) / (smem + 1024));       // 1 KB per resident CTA is reserved by the system
      int64_t grid = (H + fw - 1) / fw;
      const int64_t cap = (int64_t)h->sm_count * per_sm;          // resident CTAs only: work is handed out dynamically
      if (grid > cap) grid = cap;
      double* hg = nullptr;
      if (hglob) {
        const size_t need = sizeof(double) * (size_t)nu * ld * fw * (size_t)cap;
        if (need > h->fit_scratch_bytes) {
          if (h->fit_scratch) cudaFree(h->fit_scratch);
          h->fit_scratch = nullptr; h->fit_scratch_bytes = 0;
          BAX_CUDA(cudaMalloc(&h->fit_scratch, need));
          h->fit_scratch_bytes = need;
        }
        hg = h->fit_scratch;
      }
      BAX_CUDA(cudaMemsetAsync((unsigned long long*)h->fit_stats + 2, 0, 2 * sizeof(unsigned long long), s));
      BAX_CUDA(cudaFuncSetAttribute(k_rsx_fit_warp<true, 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      BAX_CUDA(cudaFuncSetAttribute(k_rsx_fit_warp<true, 4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      BAX_CUDA(cudaFuncSetAttribute(k_rsx_fit_warp<true, 16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      BAX_CUDA(cudaFuncSetAttribute(k_rsx_fit_warp<false, 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int32_t* deferred = nullptr;
      if (h->fit_mode != 2 && small + hv <= 200 * 1024) {          // env RSX_FIT_MODE=2: fall-backs inline, as before
        if (H > h->cap_deferred) {
          if (h->deferred) cudaFree(h->deferred);
          h->deferred = nullptr; h->cap_deferred = 0;
          BAX_CUDA(cudaMalloc(&h->deferred, sizeof(int32_t) * H));
          h->cap_deferred = H;
        }
        deferred = h->deferred;
        if (overlap && h->side) {
          if (H > h->cap_flag) {
            if (h->deferred_flag) cudaFree(h->deferred_flag);
            h->deferred_flag = nullptr; h->cap_flag = 0;
            BAX_CUDA(cudaMalloc(&h->deferred_flag, (size_t)H));
            h->cap_flag = H;
          }
          BAX_CUDA(cudaMemsetAsync(h->deferred_flag, 0, (size_t)H, s));
        }
      }
      const bool side = deferred && overlap && h->side && h->deferred_flag;
      uint8_t* dflag = side ? h->deferred_flag : nullptr;
      RsxTimer t(h, RSX_T_FIT, s, deferred ? 2 : 1);
      unsigned long long* fst = (unsigned long long*)h->fit_stats;
      if (deferred && lockstep)
        k_rsx_fit_warp<true, 16, true><<<(unsigned)grid, 32 * 16, smem, s>>>(C, h->N, h->M, h->v, h->n, samp_v, samp_m, H, Rt, fst, deferred, hg, 1, dflag);
      else if (deferred && fw == 4)
        k_rsx_fit_warp<true, 4, false><<<(unsigned)grid, 32 * 4, smem, s>>>(C, h->N, h->M, h->v, h->n, samp_v, samp_m, H, Rt, fst, deferred, hg, compact ? 1 : 0, dflag);
      else if (deferred)
        k_rsx_fit_warp<true, 2, false><<<(unsigned)grid, 32 * 2, smem, s>>>(C, h->N, h->M, h->v, h->n, samp_v, samp_m, H, Rt, fst, deferred, hg, compact ? 1 : 0, dflag);
      else
        k_rsx_fit_warp<false, 2, false><<<(unsigned)grid, 32 * 2, smem, s>>>(C, h->N, h->M, h->v, h->n, samp_v, samp_m, H, Rt, fst, deferred, hg, 0, nullptr);
      BAX_LAUNCH_CHECK();
      if (deferred) {
        BAX_CUDA(cudaFuncSetAttribute(k_rsx_fit_deferred, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(small + hv)));
        cudaStream_t sd = s;
        if (side) {
          BAX_CUDA(cudaEventRecord(h->ev_main, s));
          BAX_CUDA(cudaStreamWaitEvent(h->side, h->ev_main, 0));
          sd = h->side;
        }
        k_rsx_fit_deferred<<<4 * h->sm_count, FIT_THREADS, small + hv, sd>>>(C, h->N, h->M, h->v, h->n, samp_v, samp_m, deferred,
                                                                          (const unsigned long long*)h->fit_stats,
                                                                          (unsigned long long*)h->fit_stats + 4, Rt);
        BAX_LAUNCH_CHECK();
        if (side) { BAX_CUDA(cudaEventRecord(h->ev_side, h->side)); *overlap = 1; }
      }
      return BAX_OK;
    }
  }
  int64_t grid = H;
  double* hvg = nullptr;
  size_t smem = small + hv;
  if (smem > 200 * 1024) {                       // many cameras: eigen-solver workspace in global memory (L2)
    smem = small;
    grid = H < 2 * h->sm_count ? H : 2 * h->sm_count;
    if (hv * grid > h->fit_scratch_bytes) {
      if (h->fit_scratch) cudaFree(h->fit_scratch);
      h->fit_scratch = nullptr; h->fit_scratch_bytes = 0;
      BAX_CUDA(cudaMalloc(&h->fit_scratch, hv * grid));
      h->fit_scratch_bytes = hv * grid;
    }
    hvg = h->fit_scratch;
  } else if (grid > 64 * (int64_t)h->sm_count) {
    grid = 64 * (int64_t)h->sm_count;
  }
  BAX_CUDA(cudaFuncSetAttribute(k_rsx_fit, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  RsxTimer t(h, RSX_T_FIT, s, 1);
  k_rsx_fit<<<(unsigned)grid, FIT_THREADS, smem, s>>>(C, h->N, h->M, h->v, h->n, samp_v, samp_m, H, Rt, hvg);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}

extern "C" int rsx_fit_score(rsx_handle* h, const int32_t* samp_v, const int32_t* samp_m, int64_t H, double* score,
                             double* Rt, double* parts, void* stream) {
  if (!h || !h->v || !samp_v || !samp_m || H < 1 || !score) { bax_set_error("rsx_fit_score: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  double* models = Rt;
  if (!models) { int rc = ensure_models(h, H); if (rc) return rc; models = h->Rt; }
  int pending = 0;
  int rc = rsx_launch_fit(h, samp_v, samp_m, H, models, s, h->overlap_deferred ? &pending : nullptr); if (rc) return rc;
  {
    const size_t smem = sizeof(double) * (((12 * h->C + 1) & ~1) + 2 * 32 + 9 * h->C);
    int64_t grid = H < 64 * (int64_t)h->sm_count ? H : 64 * (int64_t)h->sm_count;
    RsxTimer t(h, RSX_T_SCORE, s, pending ? 2 : 1);
    const uint8_t* skip = pending ? h->deferred_flag : nullptr;
    auto launch = [&](const uint8_t* sk, const int32_t* list, const unsigned long long* nl, int64_t g) {
      if (h->prune && h->score_mode == 1 && !parts)
        k_rsx_score_prune<3><<<(unsigned)g, SCORE_THREADS, smem, s>>>(h->C, h->N, h->M, h->v, h->n, models, H, score, sk, list, nl,
                                                                       (unsigned long long*)h->fit_stats + 5, (unsigned long long*)h->fit_stats + 6);
      else if (h->score_mode == 0)
        k_rsx_score<0, 3><<<(unsigned)g, SCORE_THREADS, smem, s>>>(h->C, h->N, h->M, h->v, h->n, models, H, score, parts, sk, list, nl);
      else if (h->score_mode == 2)
        k_rsx_score<1, 2><<<(unsigned)g, SCORE_THREADS, smem, s>>>(h->C, h->N, h->M, h->v, h->n, models, H, score, parts, sk, list, nl);
      else
        k_rsx_score<1, 3><<<(unsigned)g, SCORE_THREADS, smem, s>>>(h->C, h->N, h->M, h->v, h->n, models, H, score, parts, sk, list, nl);
    };
    launch(skip, nullptr, nullptr, grid);
    BAX_LAUNCH_CHECK();
    if (pending) {                                   // the deferred fits ran on the side stream meanwhile
      BAX_CUDA(cudaStreamWaitEvent(s, h->ev_side, 0));
      launch(nullptr, h->deferred, (const unsigned long long*)h->fit_stats + 3, grid < 4 * (int64_t)h->sm_count ? grid : 4 * (int64_t)h->sm_count);
      BAX_LAUNCH_CHECK();
    }
  }
  return BAX_OK;
}

// Bail-out scoring on / off (see k_rsx_score_prune); every call resets the running bound and the count.
extern "C" int rsx_prune(rsx_handle* h, int on, void* stream) {
  if (!h) { bax_set_error("rsx_prune: null handle"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  h->prune = on ? 1 : 0;
  static const unsigned long long init[2] = {0x7ff0000000000000ULL /* +inf */, 0ULL};
  BAX_CUDA(cudaMemcpyAsync((unsigned long long*)h->fit_stats + 5, init, sizeof(init), cudaMemcpyHostToDevice, s));
  return BAX_OK;
}
// out_host[0..1] = { hypotheses dropped by the bail-out since the last rsx_prune, the best complete score published so
// far (+inf if none) }.  Synchronises.
extern "C" int rsx_prune_stats(rsx_handle* h, int64_t* pruned_host, double* best_host, void* stream) {
  if (!h || !pruned_host) { bax_set_error("rsx_prune_stats: bad argument"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  unsigned long long v[2];
  BAX_CUDA(cudaMemcpyAsync(v, (unsigned long long*)h->fit_stats + 5, sizeof(v), cudaMemcpyDeviceToHost, s));
  BAX_CUDA(cudaStreamSynchronize(s));
  pruned_host[0] = (int64_t)v[1];
  if (best_host) memcpy(best_host, &v[0], sizeof(double));
  return BAX_OK;
}

extern "C" int rsx_argmin(rsx_handle* h, const double* score, int64_t H, int64_t* best_host, double* score_host, void* stream) {
  if (!h || !score || H < 1 || !best_host) { bax_set_error("rsx_argmin: bad argument"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  {
    RsxTimer t(h, RSX_T_OTHER, s, 1);
    k_rsx_argmin<<<1, 1024, 0, s>>>(score, H, h->red);
    BAX_LAUNCH_CHECK();
  }
  BAX_CUDA(cudaMemcpyAsync(h->host_pinned, h->red, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
  BAX_CUDA(cudaStreamSynchronize(s));
  best_host[0] = (int64_t)h->host_pinned[0];
  if (score_host) score_host[0] = h->host_pinned[1];
  return BAX_OK;
}

extern "C" int rsx_winner_masks(rsx_handle* h, const double* Rt, const double* K00, double th2d, double th3d,
                                uint8_t* mask2d, uint8_t* mask3d, double* e_out, double* e3d_out,
                                int64_t* counts_host, void* stream) {
  if (!h || !h->v || !Rt || !K00 || !mask2d || !mask3d) { bax_set_error("rsx_winner_masks: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  const int64_t mx = h->M > h->N ? h->M : h->N;
  int grid = (int)((mx + 255) / 256); if (grid > 8 * h->sm_count) grid = 8 * h->sm_count;
  unsigned long long* cnt = (unsigned long long*)h->red;
  BAX_CUDA(cudaMemsetAsync(cnt, 0, 2 * sizeof(unsigned long long), s));
  {
    RsxTimer t(h, RSX_T_OTHER, s, 3);
    k_rsx_masks<<<grid, 256, sizeof(double) * 13 * h->C, s>>>(h->C, h->N, h->M, h->v, h->n, Rt, K00, th2d, th3d,
                                                             mask2d, mask3d, e_out, e3d_out);
    BAX_LAUNCH_CHECK();
    k_rsx_count<<<64, 256, 0, s>>>(mask2d, h->M, cnt);
    k_rsx_count<<<64, 256, 0, s>>>(mask3d, h->N, cnt + 1);
    BAX_LAUNCH_CHECK();
  }
  unsigned long long hc[2] = {0, 0};
  BAX_CUDA(cudaMemcpyAsync(hc, cnt, sizeof(hc), cudaMemcpyDeviceToHost, s));
  BAX_CUDA(cudaStreamSynchronize(s));
  if (counts_host) { counts_host[0] = (int64_t)hc[0]; counts_host[1] = (int64_t)hc[1]; }
  return BAX_OK;
}

extern "C" int rsx_triangulate(rsx_handle* h, const double* Rt, double* X4_out, void* stream) {
  if (!h || !h->n || !Rt || !X4_out) { bax_set_error("rsx_triangulate: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  int grid = (int)((h->M + 255) / 256); if (grid > 8 * h->sm_count) grid = 8 * h->sm_count;
  RsxTimer t(h, RSX_T_OTHER, s, 1);
  k_rsx_triangulate<<<grid, 256, sizeof(double) * 12 * h->C, s>>>(h->C, h->M, h->n, Rt, X4_out);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}

extern "C" int rsx_fit_stats(rsx_handle* h, int64_t* out_host, void* stream) {
  if (!h || !out_host) { bax_set_error("rsx_fit_stats: bad argument"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  unsigned long long st[8];
  BAX_CUDA(cudaMemcpyAsync(st, h->fit_stats, sizeof(st), cudaMemcpyDeviceToHost, s));
  BAX_CUDA(cudaStreamSynchronize(s));
  out_host[0] = h->fits_total; out_host[1] = (int64_t)st[0]; out_host[2] = (int64_t)st[4]; out_host[3] = (int64_t)st[1];
  return BAX_OK;
}

extern "C" int rsx_timing_enable(rsx_handle* h, int on) {
  h->timing = on;
  for (int i = 0; i < RSX_T_COUNT; ++i) { h->t_ms[i] = 0; h->t_launch[i] = 0; }
  return BAX_OK;
}
extern "C" int rsx_timing_read(rsx_handle* h, double* ms, int64_t* nl) {
  for (int i = 0; i < RSX_T_COUNT; ++i) { if (ms) ms[i] = h->t_ms[i]; if (nl) nl[i] = h->t_launch[i]; }
  return BAX_OK;
}

// ------------------------------------------------------------------ host: NumPy-compatible sample stream
// PCG64 = pcg_setseq_128_xsl_rr_64 (NumPy's default BitGenerator), 32-bit draws served low half first.
namespace {
typedef unsigned __int128 u128;
struct Pcg64 {
  u128 state, inc; int has32; uint32_t buf;
  inline uint64_t next64() {
    const u128 mult = ((u128)2549297995355413924ULL << 64) | (u128)4865540595714422341ULL;
    state = state * mult + inc;
    const uint64_t hi = (uint64_t)(state >> 64), lo = (uint64_t)state, x = hi ^ lo;
    const unsigned r = (unsigned)(state >> 122);
    return (x >> r) | (x << ((64 - r) & 63));
  }
  inline uint32_t next32() {
    if (has32) { has32 = 0; return buf; }
    const uint64_t v = next64();
    has32 = 1; buf = (uint32_t)(v >> 32);
    return (uint32_t)v;
  }
  // random_bounded_uint64(off = 0, rng, mask = 0, use_masked = false) for rng < 2^32 - 1: Lemire's method
  inline uint32_t bounded(uint32_t rng) {
    if (rng == 0) return 0;
    const uint32_t rng_excl = rng + 1;
    uint64_t m = (uint64_t)next32() * rng_excl;
    uint32_t left = (uint32_t)m;
    if (left < rng_excl) {
      const uint32_t th = (0xffffffffu - rng) % rng_excl;
      while (left < th) { m = (uint64_t)next32() * rng_excl; left = (uint32_t)m; }
    }
    return (uint32_t)(m >> 32);
  }
  // Generator.choice(pop, size = 10, replace = False): Floyd's sampling, then _shuffle_int
  inline void choice10(int64_t pop, int32_t* out) {
    for (int k = 0; k < RSX_SAMPLE; ++k) {
      const int64_t j = pop - RSX_SAMPLE + k;
      int64_t val = bounded((uint32_t)j);
      for (int q = 0; q < k; ++q) if (out[q] == val) { val = j; break; }
      out[k] = (int32_t)val;
    }
    for (int i = RSX_SAMPLE - 1; i >= 1; --i) {
      const uint32_t j = bounded((uint32_t)i);
      const int32_t tmp = out[i]; out[i] = out[j]; out[j] = tmp;
    }
  }
};
}  // namespace

extern "C" int rsx_sample_stream(uint64_t* pcg, int64_t N, int64_t M, int64_t H, int32_t* samp_v, int32_t* samp_m) {
  if (!pcg || !samp_v || !samp_m || N < RSX_SAMPLE || M < RSX_SAMPLE || H < 0 || N >= 0x7fffffffLL || M >= 0x7fffffffLL) {
    bax_set_error("rsx_sample_stream: bad argument (populations must hold %d..2^31-2 items)", RSX_SAMPLE); return BAX_E_ARG;
  }
  Pcg64 g;
  g.state = ((u128)pcg[0] << 64) | pcg[1]; g.inc = ((u128)pcg[2] << 64) | pcg[3];
  g.has32 = (int)pcg[4]; g.buf = (uint32_t)pcg[5];
  for (int64_t h = 0; h < H; ++h) {
    g.choice10(N, samp_v + h * RSX_SAMPLE);
    g.choice10(M, samp_m + h * RSX_SAMPLE);
  }
  pcg[0] = (uint64_t)(g.state >> 64); pcg[1] = (uint64_t)g.state; pcg[4] = (uint64_t)g.has32; pcg[5] = g.buf;
  return BAX_OK;
}
