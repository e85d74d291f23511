/* bax.h — C ABI of the B200-native bundle-adjustment hot path (libbax.so).
 *
 * The reference (pure Python) has no FFI; its boundary for this path is the set of Python
 * functions in ba.py and util.py.  Each entry point below names the reference lines it
 * replaces.  All array arguments are CUDA DEVICE pointers unless marked "host"; every call
 * is ordered on the given cudaStream_t (passed as void*; 0 = legacy default stream), returns
 * 0 on success and a negative code on failure (bax_last_error() gives the thread-local text).
 * No exception crosses the ABI.  One handle per device; calls on a handle are not re-entrant.
 *
 * Parameter vector theta (reference ba.py:18-31):  [ rvec_0..rvec_{C-1} | t_0..t_{C-1} | x_0..x_{P-1} ],
 * P = N*J points, n = 6C + 3P.  Residual rows (reference ba.py:113-140): camera-major visible
 * reprojection rows (ex,ey interleaved), then lambda1 * orientation rows, then lambda2 * B
 * bone-length-variance rows.
 */
#ifndef BAX_H
#define BAX_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct bax_handle bax_handle;

enum { BAX_OK = 0, BAX_E_ARG = -1, BAX_E_CUDA = -2, BAX_E_STATE = -3, BAX_E_NUMERIC = -4 };

const char* bax_last_error(void);
int bax_version(void);

int bax_create(int device, bax_handle** out);
int bax_destroy(bax_handle* h);
/* Host waits of this handle's calls (scalar read-backs) sleep on a blocking-sync event instead of spinning: for callers
 * that run many handles from many threads (bax_batch_solve does it for its workers). */
int bax_set_blocking_sync(bax_handle* h, int on);

/* Bind one BA problem (replaces the argument tuple of ba.objfun, reference ba.py:113 and
 * ba_main ba.py:157-206).  y2d [C][N*J][2], s2d [C][N*J] are referenced, not copied (caller keeps
 * them alive); p3d [C][N][J3][3], s3d [C][N][J3] are consumed here: the constant unit bone
 * directions of the orientation term and their valid-row compaction (ba.py:73-91) are derived
 * once.  bones: host int32 [B][2] (util.OP_BONE, util.py:129-145); it indexes the J axis for the
 * bone-length term and the J3 axis for the orientation term, as the reference does. */
int bax_problem_set(bax_handle* h, int C, int N, int J, int J3, int B,
                    const double* K, const double* y2d, const double* s2d,
                    const double* p3d, const double* s3d, const int32_t* bones_host,
                    double lambda1, double lambda2, void* stream);

/* Point the bound problem at another copy of the observations (same shapes, same visibility pattern s2d > 0): lets a
 * caller upload the next batch into a second device buffer while kernels still read the first (double buffering of
 * the host -> device stream).  Takes effect for calls issued afterwards. */
int bax_observations_rebind(bax_handle* h, const double* y2d, const double* s2d);

/* dims_host[0..5] = { n_vis (visible observations), n_v3d (valid orientation rows),
 *                     m (residual rows), n (parameters), aug_rows (6C+13), n_points } */
int bax_problem_dims(bax_handle* h, int64_t* dims_host);

/* r_out[m] = ba.objfun(theta, ...)  (reference ba.py:113-140: objfun_nll 34-63,
 * objfun_var3d 67-96, objfun_varbone 99-110). */
int bax_residual(bax_handle* h, const double* theta, double* r_out, void* stream);

/* Fused residual + analytic Jacobian blocks of the reprojection rows, one thread per
 * (camera, frame, joint) observation, compacted to visible observations in residual-row order.
 *   r2   [n_vis][2]      Jcam [n_vis][2][6] (d/d rvec_c | d/d t_c)     Jpt [n_vis][2][3] (d/d x_p)
 *   obs_cam, obs_pt [n_vis] int32 (may be NULL): camera / point index of each compacted row pair.
 * The reference has no Jacobian: SciPy finite-differences ba.objfun (ba.py:187-206). */
int bax_residual_jacobian(bax_handle* h, const double* theta, double* r2, double* Jcam, double* Jpt,
                          int32_t* obs_cam, int32_t* obs_pt, void* stream);

/* Jacobian of the two prior families:  Jv3d [n_v3d][3C] (d/d rvecs, includes lambda1);
 * Jvb [N][B][3] = lambda2 * d r_b / d x[n, bone_b[0]]  (the entry for bone_b[1] is its negative). */
int bax_prior_jacobian(bax_handle* h, const double* theta, double* Jv3d, double* Jvb, void* stream);

/* ---- trust-region / LM step on the Schur complement (replaces what SciPy's trf_no_bounds does
 * with the dense FD Jacobian and its SVD: scipy/optimize/_lsq/trf.py:415-587, common.py:57-168).
 * Every stage is split at the points where a frame-sharded multi-GPU run must sum a small packed
 * buffer over ranks; bax_reduce_buffer() exposes that buffer (device pointer + length in doubles)
 * so the host can all-reduce it in place (NCCL via torch.distributed) between the two calls.
 * On one GPU the *_finish call simply follows. */
enum { BAX_BUF_BONE = 0, BAX_BUF_COST = 1, BAX_BUF_LIN = 2, BAX_BUF_SCHUR = 3, BAX_BUF_RHS2 = 4,
       BAX_BUF_STEP = 5 /* not reduced: trial scalars */, BAX_BUF_NORM = 6, BAX_BUF_GMAX = 7 /* MAX-reduced */,
       BAX_BUF_RHS2_HEAD = 8 /* the first 6C+B doubles of BAX_BUF_RHS2: what bax_trial_phi_prime fills */ };
int bax_reduce_buffer(bax_handle* h, int which, double** dev_ptr, int64_t* count);
/* global frame count used by the bone-length variance (sum of N over ranks); default = local N */
int bax_set_global_frames(bax_handle* h, int64_t n_frames_global);

/* cost = 0.5 * ||objfun(theta)||^2 without materialising the residual vector. */
int bax_cost_begin(bax_handle* h, const double* theta, void* stream);   /* fills BAX_BUF_BONE  */
int bax_cost_mid(bax_handle* h, void* stream);                          /* fills BAX_BUF_COST  */
int bax_cost_finish(bax_handle* h, double* cost_host, void* stream);    /* synchronises        */

/* Linearise at theta: gradient, camera blocks H_cc (reprojection + orientation rows), per-point
 * 3x3 blocks V_p and g_p.  out_host[0..4] = { cost, ||g||_inf, ||g||_2, ||theta||_2, max diag V_p } */
int bax_linearize_begin(bax_handle* h, const double* theta, void* stream); /* fills BAX_BUF_BONE */
int bax_linearize_mid(bax_handle* h, void* stream);                        /* fills BAX_BUF_LIN  */
int bax_linearize_post(bax_handle* h, void* stream);   /* fills BAX_BUF_NORM (sum) and BAX_BUF_GMAX (max) */
int bax_linearize_finish(bax_handle* h, double* out_host, void* stream);   /* synchronises       */

/* Speculative linearisation: bax_linearize_push() saves the current linearisation (pointer swap, no copy) so that
 * the next bax_linearize_* sequence can run at a trial point; bax_linearize_pop() restores the saved one if that
 * trial step is rejected.  After a push the BAX_BUF_LIN reduce buffer has a different address (query it again). */
int bax_linearize_push(bax_handle* h);
int bax_linearize_pop(bax_handle* h);

/* One damped solve (J^T J + alpha I) p = -g through the per-point Schur complement with the
 * rank-B Woodbury correction for the bone-length rows; alpha > 0.
 *   begin : per-point D_p = V_p + alpha I factored in registers, whitened blocks, SYRK  -> BAX_BUF_SCHUR
 *   mid   : FP64 single-CTA Cholesky of the reduced camera system, back-substitution    -> BAX_BUF_RHS2 / BAX_BUF_STEP
 *   finish: second reduced solve for phi'(alpha); out_host[0..3] = { ||p||^2, g.p, p^T (H+alpha I)^-1 p, 0 }
 * step [n] receives p (device). */
int bax_trial_begin(bax_handle* h, double alpha, void* stream);
int bax_trial_mid(bax_handle* h, double* step, int need_phi_prime, void* stream);
int bax_trial_finish(bax_handle* h, double* out_host, void* stream);
/* phi'(alpha) on demand, after a mid(need_phi_prime = 0) / finish pair: forms the second right-hand side
 * (-> BAX_BUF_RHS2_HEAD, to be all-reduced by a frame-sharded run); call bax_trial_finish again for out_host[2]. */
int bax_trial_phi_prime(bax_handle* h, void* stream);

/* ---- whole trust-region iterations behind one call (one GPU; a frame-sharded run all-reduces between the phases
 * above and is driven from the host language instead).  Same scalar logic as SciPy's trf_no_bounds
 * (scipy/optimize/_lsq/trf.py:415-587 — what reference ba.py:187-206 runs) with x_scale = 1: Delta_0 = ||x_0||,
 * More'-Sorensen alpha iteration, boundary-step rescaling, update_tr_radius, check_termination; every trial point
 * is linearised straight away (bax_linearize_push / _pop).  The caller fills the options and the three device
 * vectors (n doubles each; x holds the start) and keeps the struct alive between calls; x and x_new swap roles on
 * accepted steps, so the solution is wherever st->x points afterwards. */
#define BAX_TRF_MAX_HIST 64
typedef struct bax_trf_state {
  double ftol, xtol, gtol;        /* SciPy defaults of the reference call: 1e-4 (ba.py:190), 1e-8, 1e-8 */
  int64_t max_nfev;               /* <= 0: 100 * n (SciPy's default for trf) */
  double *x, *x_new, *step;       /* device, n doubles each */
  double cost, g_inf, g_2, x_norm, h_scale, Delta, alpha, initial_cost;
  int64_t nfev, njev, iteration, n_trials;
  int32_t status;                 /* -1 while running, else SciPy's status 0..4 */
  int32_t n_hist;                 /* inner passes recorded by the last bax_trf_iterate */
  double hist[BAX_TRF_MAX_HIST][8]; /* alpha, Delta, step norm, cost at the trial point, predicted, actual, More' its, accepted */
} bax_trf_state;
int bax_trf_begin(bax_handle* h, bax_trf_state* st, void* stream);    /* linearises at st->x, sets Delta_0 */
int bax_trf_iterate(bax_handle* h, bax_trf_state* st, void* stream);  /* one pass of the outer loop body */
int bax_trf_solve(bax_handle* h, bax_trf_state* st, void* stream);    /* after begin: iterate until termination; hist = last pass */

/* ---- collectives inside the library (frame-sharded runs).  With a communicator attached, bax_trf_begin /
 * _iterate / _solve all-reduce the packed buffers themselves between their phases, on the call's stream: BONE -> LIN ->
 * {NORM (sum), GMAX (max)} in one group call per linearisation, SCHUR -> RHS2 per damped solve (5 collectives per
 * trust-region iteration, none of them through the host language).  NCCL is bound at run time (dlopen of libnccl.so.2);
 * no reference counterpart (the reference is single-process).
 *   bax_comm_unique_id : HOST, rank 0: 128 bytes to be handed to every rank (ncclGetUniqueId)
 *   bax_comm_create    : every rank, collectively (ncclCommInitRank); the communicator can serve many handles
 *   bax_allreduce_attach(h, comm) : comm is an ncclComm_t (this library's or the caller's own); NULL detaches
 *   bax_comm_stats     : out_host[0..1] = { attached, collectives issued since bax_create } */
int bax_comm_unique_id(void* id128_host);
int bax_comm_create(int device, const void* id128_host, int rank, int world, void** comm_out);
int bax_comm_destroy(void* comm);
int bax_allreduce_attach(bax_handle* h, void* nccl_comm);
int bax_comm_stats(bax_handle* h, int64_t* out_host);

/* ---- sweep over independent, identically shaped problems: BASELINE configs[4] (4096 four-camera scenes); the reference
 * runs ba.py once per sequence from run_all.sh (ba.py:157-210 each).  HOST pointers throughout: per-problem arrays are
 * given as arrays of n_problems pointers (a problem may appear more than once), shapes as for bax_problem_set.
 * n_workers native threads, each with its own stream and handle, run per problem exactly what ba_main runs:
 * bax_problem_set -> bax_triangulate(R0, t0) -> theta0 = [rvec0 | t0 | X] -> bax_trf_begin -> bax_trf_solve.
 * theta_out [n_problems][6C + 3NJ]; info_out [n_problems][8] = { status (SciPy's 0..4; -10: triangulation met a
 * rank-deficient point, -11: non-finite start), cost, initial cost, nfev, njev, iterations, optimality, tri status }.
 * Replicas only, no collective.  Blocks until every problem is done. */
typedef struct bax_batch_args {
  int32_t n_problems, n_workers, device;
  int32_t C, N, J, J3, B;
  const double* const* K;        /* [C][9]        */
  const double* const* y2d;      /* [C][N*J][2]   */
  const double* const* s2d;      /* [C][N*J]      */
  const double* const* p3d;      /* [C][N][J3][3] or NULL */
  const double* const* s3d;      /* [C][N][J3]    or NULL */
  const int32_t* bones;          /* [B][2]        */
  double lambda1, lambda2;
  const double* const* R0;       /* [C][9]  start cameras (as given to ba_main) */
  const double* const* t0;       /* [C][3]  */
  const double* const* rvec0;    /* [C][3]  cv2.Rodrigues(R0) (ba.py:19) */
  double ftol, xtol, gtol;
  int64_t max_nfev;              /* <= 0: 100 n */
  double* theta_out;
  double* info_out;
  int32_t blocking_sync;         /* 1: the workers' host waits sleep (bax_set_blocking_sync) instead of spinning: for
                                    hosts with fewer cores than worker threads */
} bax_batch_args;
int bax_batch_solve(const bax_batch_args* a);
/* The workers of a call (stream, handle, pinned staging, device buffers: ~60 ms to create twelve) are parked for the next
 * call whose problems fit; this frees them. */
int bax_batch_release(void);

/* Weighted linear triangulation of every (frame, joint) — util.triangulate_with_conf,
 * reference util.py:325-354 (+ triangulate_point 308-322).  R [C][9], t [C][3]; X [N*J][3] (NaN where
 * the reference returns NaN); status_host: 0 ok, 1 = some point has rank < 3 ("not enough constraint"). */
int bax_triangulate(bax_handle* h, const double* R, const double* t, double* X, int32_t* status_host,
                    void* stream);

/* Reprojection error in pixels of every observation under cameras (R [C][9], t [C][3]) and points X [N*J][3]:
 * eval.calc_repro_error (reference eval.py:76-108) and the distance test of ba.save_mask (ba.py:248-256).
 * e_out [C][N*J]; NaN where the observation is invisible or the point is NaN. */
int bax_reproject_error(bax_handle* h, const double* R, const double* t, const double* X, double* e_out, void* stream);

/* timing hooks for bench.py: CUDA-event time (ms) accumulated on the call's stream per kernel family */
enum { BAX_T_RJ = 0, BAX_T_BUILD = 1, BAX_T_SYRK = 2, BAX_T_DENSE = 3, BAX_T_BACKSUB = 4, BAX_T_LIN = 5,
       BAX_T_COST = 6, BAX_T_COUNT = 8 };
int bax_timing_enable(bax_handle* h, int on);
int bax_timing_read(bax_handle* h, double* ms_host /*[BAX_T_COUNT]*/, int64_t* launches_host /*[BAX_T_COUNT]*/);
int bax_timing_reset(bax_handle* h);

#ifdef __cplusplus
}
#endif
#endif
