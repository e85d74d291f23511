// rsx_handle: state of the RANSAC path (include/rsx.h), shared by rsx.cu and rsx_final.cu.
#pragma once
#include "bax_internal.cuh"
#include "../../include/rsx.h"

enum { RSX_T_FIT = 0, RSX_T_SCORE = 1, RSX_T_FINAL = 2, RSX_T_OTHER = 3, RSX_T_COUNT = 4 };

struct rsx_handle {
  int device, sm_count;
  int C; int64_t N, M;
  double* v;            // [C][N][3]  (copy)
  double* n;            // [C][M][3]  (copy)
  double* Rt;           // [cap_h][C][12] models of the current chunk
  int64_t cap_h;
  double* fit_scratch;  // global H/V workspace when 3(10+C) is too large for shared memory
  size_t fit_scratch_bytes;
  int fit_mode;         // 1: warp-per-hypothesis arrow solver where it fits, Jacobi fall-backs in a second pass (env RSX_FIT_MODE=0: CTA Jacobi kernel; 2: fall-backs inline)
  int prune;            // rsx_prune: exact bail-out in the score pass (fit_stats[5] = bits of the best complete score, [6] = dropped)
  int score_mode;       // 1: Rayleigh-quotient DLT eigen-solver + common-denominator error sum (env RSX_SCORE_MODE=0: round-1 arithmetic)
  void* fit_stats;      // device [8] uint64: Jacobi fall-backs, total sweeps of the arrow solver, work counter of k_rsx_fit_warp, deferred fits of this launch, deferred fits in total
  double last_eigs[6];  // Ritz values of the 6 smallest eigenpairs of the last rsx_calib_linear (ascending)
  int64_t fits_total;   // hypotheses fitted since rsx_create
  int32_t* deferred;    // hypotheses of the current launch left to k_rsx_fit_deferred
  int64_t cap_deferred;
  uint8_t* deferred_flag;  // [cap_flag] 1 = left to the second pass (the score kernel skips it until that pass is done)
  int64_t cap_flag;
  cudaStream_t side;       // the second fit pass runs here, under the score kernel of the finished hypotheses
  cudaEvent_t ev_main, ev_side;
  int overlap_deferred;    // env RSX_OVERLAP_DEFERRED=0: second pass on the caller's stream, as in round 1
  double* red;          // small reduction scratch
  size_t red_bytes;
  double* host_pinned;  // [64]
  bax_handle* bx;       // borrowed SYRK / partial-sum machinery (only sm_count, syrk_part, partials, timing used)
  int timing;
  cudaEvent_t ev0, ev1;
  double t_ms[RSX_T_COUNT];
  int64_t t_launch[RSX_T_COUNT];
};

struct RsxTimer {
  rsx_handle* h; int which; cudaStream_t s; int nl;
  RsxTimer(rsx_handle* h_, int w, cudaStream_t s_, int launches) : h(h_), which(w), s(s_), nl(launches) {
    if (h->timing) cudaEventRecord(h->ev0, s);
  }
  ~RsxTimer() {
    h->t_launch[which] += nl;
    if (h->timing) {
      cudaEventRecord(h->ev1, s); cudaEventSynchronize(h->ev1);
      float ms = 0; cudaEventElapsedTime(&ms, h->ev0, h->ev1); h->t_ms[which] += ms;
    }
  }
};
