/* rsx.h — C ABI of the B200-native RANSAC linear-calibration hot path (part of libbax.so).
 *
 * The reference (pure Python) has no FFI; its boundary for this path is the set of Python functions
 * in calib_ransac.py.  Each entry point names the reference lines it replaces.  Unless marked "host",
 * array arguments are CUDA DEVICE pointers; calls are ordered on the given cudaStream_t (void*; 0 =
 * legacy default stream) and return 0 or a negative BAX_E_* code (bax_last_error() has the text).
 *
 *   v [C][N][3]  unit bone directions per camera        (calib_ransac.py:243, util.py:496-544)
 *   n [C][M][3]  K^-1-normalised rays [u, v, 1]         (calib_ransac.py:256-270)
 *   a hypothesis = 10 direction indices + 10 ray indices (calib_ransac.py:150-151)
 */
#ifndef RSX_H
#define RSX_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define RSX_SAMPLE 10   /* minimal-sample size hard-coded at calib_ransac.py:150-151 */

typedef struct rsx_handle rsx_handle;

int rsx_create(int device, rsx_handle** out);
int rsx_destroy(rsx_handle* h);

/* HOST ONLY, no GPU needed.  The reference's sequential sample stream (calib_ransac.py:148-151):
 * rng = np.random.default_rng(seed); per hypothesis rng.choice(N, 10, replace=False) then
 * rng.choice(M, 10, replace=False).  NumPy's algorithm restated: PCG64 (XSL-RR 128/64) with its 32-bit
 * half-word buffer, Lemire bounded integers, Floyd's sampling and the trailing Fisher-Yates shuffle.
 * pcg[0..5] = { state_hi, state_lo, inc_hi, inc_lo, has_uint32, uinteger } as read from
 * Generator.bit_generator.state; updated in place so a stream can be continued in chunks.
 * samp_v, samp_m: host int32 [H][10]. */
int rsx_sample_stream(uint64_t* pcg, int64_t N, int64_t M, int64_t H, int32_t* samp_v, int32_t* samp_m);

/* Bind the scene: copies v and n into the handle's structure-of-arrays layout. */
int rsx_problem_set(rsx_handle* h, int C, int64_t N, int64_t M, const double* v, const double* n, void* stream);

/* calib_linear on each minimal sample (calib_ransac.py:51-131 called at :152) and its continuous score
 * (calib_ransac.py:154-180):  score = ||reprojection error||_2 / M + ||angular error||_2 / N.
 *   samp_v, samp_m  device int32 [H][10]
 *   score           device [H]
 *   Rt              device [H][C][12] (row-major [R | t] per camera) or NULL
 *   parts           device [H][2] = { sum e^2, sum e3d^2 } or NULL */
int rsx_fit_score(rsx_handle* h, const int32_t* samp_v, const int32_t* samp_m, int64_t H, double* score,
                  double* Rt, double* parts, void* stream);

/* Exact bail-out in the score pass of rsx_fit_score (no counterpart in the reference, which scores every hypothesis in
 * full at calib_ransac.py:154-181 and then keeps the minimum, :182-189).  The score only grows while rays are added, so
 * a hypothesis whose partial score already exceeds the best complete score published so far cannot be the first minimum:
 * it is dropped and reported as +inf.  The winner, its score and every output derived from it are bit-identical to the
 * unpruned pass; WHICH losers are cut short depends on timing.  on = 1 / 0; every call resets the running bound and the
 * count (call it once per RANSAC run; the bound then carries over the chunks of that run).  Ignored when rsx_fit_score
 * is asked for `parts`.  rsx_prune_stats: hypotheses dropped since the last rsx_prune and the best complete score so far
 * (+inf if none); synchronises. */
int rsx_prune(rsx_handle* h, int on, void* stream);
int rsx_prune_stats(rsx_handle* h, int64_t* pruned_host, double* best_host, void* stream);

/* First minimum of score[0..H) (strict '<' from +inf: NaN and +inf never win, calib_ransac.py:182).
 * best_host[0] = index or -1, score_host[0] = its score. Synchronises. */
int rsx_argmin(rsx_handle* h, const double* score, int64_t H, int64_t* best_host, double* score_host, void* stream);

/* The same search for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * No handle, and no cudaMalloc / cudaFree / allocating library call inside: the per-CTA partial winners live in a
 * file-scope static device array of RSX_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 *   device  CUDA device ordinal the pointers belong to
 *   score   device [H], 1 <= H <= RSX_STATIC_MAX_H; a larger H is rejected with BAX_E_ARG (use rsx_argmin)
 *   out     device [2] = { (double) index of the first minimum or -1, its score (+inf if none) }
 * Asynchronous on `stream`; does not synchronise.  The first call on a device looks the array's address up
 * (cudaGetSymbolAddress): make it once before any graph capture.  All calls share the one array per device: calls that
 * overlap on the same GPU race and must be ordered by the caller (one stream, or events). */
#define RSX_STATIC_CHUNK 4096                   /* scores per CTA = scores per scratch entry */
#define RSX_STATIC_MAX_H (1 << 24)              /* 16 777 216 scores = 4096 scratch entries */
#define RSX_STATIC_SCRATCH_BYTES 65792          /* 256 (arrival counter) + 4096 x (8 + 8) */
int rsx_argmin_static(int device, const double* score, int64_t H, double* out, void* stream);

/* Inlier masks of one model (calib_ransac.py:193-204):  mask2d[i] = all_c K00[c] * ||e_ci[:2]|| <= th2d,
 * mask3d[i] = e3d_i <= th3d.   Rt device [C][12]; K00 device [C]; mask2d uint8 [M], mask3d uint8 [N];
 * e_out [M][C][3] and e3d_out [N] optional (NULL).  counts_host[0..1] = inliers 2-D, 3-D. Synchronises. */
int rsx_winner_masks(rsx_handle* h, const double* Rt, const double* K00, double th2d, double th3d,
                     uint8_t* mask2d, uint8_t* mask3d, double* e_out, double* e3d_out, int64_t* counts_host,
                     void* stream);

/* calib_linear (calib_ransac.py:51-131) on the subsets selected by the masks (NULL = all): the final
 * inlier fit at calib_ransac.py:209 and, with NULL masks, calib_linear.py's one-shot model fit.
 * The 3(M_in + C) eigenproblem is never densified: rotations from the 3C x 3C Gram matrix; translations
 * from the 6 smallest eigenpairs of the arrow-structured normal matrix by block-eliminated inverse subspace
 * iteration with Rayleigh-Ritz.  Rt_out device [C][12]; X_out device [M_in][3] or NULL;
 * info_host[0..3] = { M_in, N_in, iterations, converged }; a run that does not settle still returns its model with
 * converged = 0 (the reference never fails here); BAX_E_NUMERIC only for a non-finite model. Synchronises. */
int rsx_calib_linear(rsx_handle* h, const uint8_t* mask3d, const uint8_t* mask2d, double* Rt_out, double* X_out,
                     int64_t* info_host, void* stream);

/* HOST: the 6 smallest eigenvalues (Ritz values, ascending) of the normal matrix from the last rsx_calib_linear on this
 * handle — what the reference gets from eigh(subset_by_index=(0, 5)) and tests as w[3] / w[4] > 1e-4 for its
 * "degenerate case" warning (calib_linear.py:94-100).  w[0..3] are converged (1e-9); w[4], w[5] are the Ritz values of
 * the 6-dimensional iteration space at that moment: upper bounds of the true eigenvalues, typically within 10-20 %. */
int rsx_calib_linear_eigs(rsx_handle* h, double* w_host /*[6]*/);

/* DLT re-triangulation of all M ray sets under one model: calib_linear.py:127-133, i.e. pycalib.calib.triangulate
 * (identical code vendored at util.py:240-258) per point.  Rt device [C][12]; X4_out device [M][4]: the
 * dehomogenised point [x, y, z, 1], or the unit eigenvector itself where np.isclose(x[3], 0) (point at infinity). */
int rsx_triangulate(rsx_handle* h, const double* Rt, double* X4_out, void* stream);

/* Solver statistics of the minimal-sample fits since rsx_create (no reference counterpart: the reference's dense eigh
 * at calib_ransac.py:96-98 never fails over).  out_host[0..3] = { hypotheses fitted, fits whose arrow-structured subspace
 * iteration did not settle and were finished by the Jacobi diagonalisation, how many of those ran in the deferred
 * second pass (k_rsx_fit_deferred), total sweeps of the arrow solver }.  Synchronises. */
int rsx_fit_stats(rsx_handle* h, int64_t* out_host /*[4]*/, void* stream);

/* timing hooks: CUDA-event ms and launch counts of { fit, score, final } since the last reset */
int rsx_timing_enable(rsx_handle* h, int on);
int rsx_timing_read(rsx_handle* h, double* ms_host /*[4]*/, int64_t* launches_host /*[4]*/);

#ifdef __cplusplus
}
#endif
#endif
