// One damped solve (J^T J + alpha I) p = -g through the whitened, augmented Schur formulation.
//
// For every point p with D_p = V_p + alpha I = L_p L_p^T (3x3, registers):
//   Y_p   = W_p L_p^-T       (6C x 3)   W_p = J_cam^T J_pt of the cameras that see p
//   Ut_p  = L_p^-1 U_p       (3 x B)    U_p = bone-length Jacobian columns touching p (<= 3 non-zero)
//   bt_p  = L_p^-1 (-g_p)    (3)
// Stack Yhat = [Y; Ut^T; bt^T]  ((6C+B+1) x 3P).  One SYRK Yhat Yhat^T yields every reduction the Schur
// complement with the rank-B Woodbury correction needs:
//   [ Y Y^T   G     rhs0 ]      S  = Hcc + alpha I - Y Y^T + G T^-1 G^T,   T = I + Ut^T Ut
//   [ G^T    T-I    h0   ]      pc = S^-1 (-gc - rhs0 + G T^-1 h0)
//   [ rhs0^T h0^T    .   ]      w  = T^-1 (h0 - G^T pc),  L_p^T p_p = bt_p - Y_p^T pc - Ut_p w
#include "bax_internal.cuh"
#include <stdlib.h>

__device__ __forceinline__ void bone_u_dev(const double* __restrict__ x, int64_t frame_base3,
                                           const int32_t* __restrict__ bones, int b, int sgn,
                                           const double* __restrict__ Lbar, double scale, double (&u)[3]) {
  const double* a = x + frame_base3 + 3 * bones[2 * b];
  const double* q = x + frame_base3 + 3 * bones[2 * b + 1];
  double d0 = a[0] - q[0], d1 = a[1] - q[1], d2 = a[2] - q[2];
  double L = sqrt(d0 * d0 + d1 * d1 + d2 * d2);
  double f = sgn * scale * (L - Lbar[b]) / L;
  u[0] = f * d0; u[1] = f * d1; u[2] = f * d2;
}

// (r02: routing the stores through a per-warp shared-memory strip so that they leave as 16-byte, fully contiguous
// 512 + 256 byte runs per row changed nothing — 0.460 ms before and after at the cfg4 shard.  The kernel is not limited
// by its store pattern: ~350 FP64 instructions per observation x 1.36e7 observations in 0.46 ms is ~60 % of the DFMA
// rate, on top of 2 GB of writes.)
__global__ void __launch_bounds__(128)
k_build_yhat(const double* __restrict__ campack, const double* __restrict__ x, const double2* __restrict__ y,
             const double* __restrict__ s, int64_t P, int C, int J, int B, const int32_t* __restrict__ bones,
             const int32_t* __restrict__ joint_bones, const double* __restrict__ Lbar, double scale,
             const double* __restrict__ Vp, const double* __restrict__ gp, double alpha,
             double* __restrict__ Yhat, int64_t ldy) {
  extern __shared__ double cams[];               // [C][CAMPACK]
  for (int i = threadIdx.x; i < C * CAMPACK; i += blockDim.x) cams[i] = campack[i];
  __syncthreads();
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const double x0 = x[3 * p], x1 = x[3 * p + 1], x2 = x[3 * p + 2];
  double V[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) V[k] = Vp[6 * p + k];
  const Chol3r L = chol3r(V, alpha);
  double* col = Yhat + 3 * p;
  // blockIdx.y = camera group: the cameras of a point are dealt over gridDim.y threads (of different CTAs) so that the
  // grid is many short CTAs instead of 2.2 waves of long ones (see bax_trial_begin); group 0 also writes bt and Ut
  const int c_lo = (int)(((int64_t)C * blockIdx.y) / gridDim.y), c_hi = (int)(((int64_t)C * (blockIdx.y + 1)) / gridDim.y);
  if (blockIdx.y == 0) {
  {  // bt = L^-1 (-g_p)
    double a = -gp[3 * p], b = -gp[3 * p + 1], d = -gp[3 * p + 2];
    chol3r_fwd(L, a, b, d);
    double* o = col + (size_t)(6 * C + B) * ldy;
    o[0] = a; o[1] = b; o[2] = d;
  }
  {  // Ut rows of the incident bones
    const int j = (int)(p % J); const int64_t fb = (p - j) * 3;
#pragma unroll
    for (int k = 0; k < BAX_MAX_JB; ++k) {
      int b = joint_bones[(j * BAX_MAX_JB + k) * 2];
      if (b >= 0) {
        double u[3];
        bone_u_dev(x, fb, bones, b, joint_bones[(j * BAX_MAX_JB + k) * 2 + 1], Lbar, scale, u);
        chol3r_fwd(L, u[0], u[1], u[2]);
        double* o = col + (size_t)(6 * C + b) * ldy;
        o[0] = u[0]; o[1] = u[1]; o[2] = u[2];
      }
    }
  }
  }
  // the (score, 2-D point) of the next BAX_PF cameras are in flight while this one is worked on:
  // a score-then-point load chain per camera left the kernel waiting on memory (ncu: long_scoreboard on top)
  double sc_q[BAX_PF]; double2 yy_q[BAX_PF];
#pragma unroll
  for (int k = 0; k < BAX_PF; ++k) {
    const int cc = c_lo + k < c_hi ? c_lo + k : c_hi - 1;
    sc_q[k] = s[(size_t)cc * P + p]; yy_q[k] = y[(size_t)cc * P + p];
  }
  for (int c0 = c_lo; c0 < c_hi; c0 += BAX_PF) {
#pragma unroll
   for (int k = 0; k < BAX_PF; ++k) {           // ring slot k is a compile-time index: the queue stays in registers
    const int c = c0 + k;
    if (c >= c_hi) break;
    const double sc = sc_q[k]; const double2 yy = yy_q[k];
    {
      const int cn = c + BAX_PF < c_hi ? c + BAX_PF : c_hi - 1;
      sc_q[k] = s[(size_t)cn * P + p]; yy_q[k] = y[(size_t)cn * P + p];
    }
    if (!(sc > 0)) continue;                     // rows of unseen cameras stay zero (buffer is zero-initialised)
    ObsLin o;
    obs_linearize<true>(cams + c * CAMPACK, x0, x1, x2, yy.x, yy.y, sc, o);
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      double j0 = i < 3 ? o.jr[0][i] : o.jt[0][i - 3];
      double j1 = i < 3 ? o.jr[1][i] : o.jt[1][i - 3];
      double w0 = j0 * o.jp[0][0] + j1 * o.jp[1][0];
      double w1 = j0 * o.jp[0][1] + j1 * o.jp[1][1];
      double w2 = j0 * o.jp[0][2] + j1 * o.jp[1][2];
      chol3r_fwd(L, w0, w1, w2);                  // row i of Y_p = (L^-1 W_i^T)^T
      double* q = col + (size_t)(6 * c + i) * ldy;
      q[0] = w0; q[1] = w1; q[2] = w2;
    }
   }
  }
}

// L_p^T p_p = bt_p - Y_p^T pc - Ut_p w ;  step written in theta order; bt_p <- L_p^-1 p_p for the
// second reduced solve; partial sums { |p_p|^2, g_p.p_p, |L^-1 p_p|^2 }.
__global__ void __launch_bounds__(128)
k_backsub(const double* __restrict__ s, int64_t P, int C, int J, int B, const int32_t* __restrict__ joint_bones,
          const double* __restrict__ Vp, const double* __restrict__ gp, double alpha,
          double* __restrict__ Yhat, int64_t ldy, const double* __restrict__ coef,
          double* __restrict__ step_pts, double* __restrict__ part, double* __restrict__ sums_out /*[3]*/,
          unsigned* __restrict__ ticket) {
  extern __shared__ double sm_all[];             // coef [6C+B] then reduction scratch [3*32]
  double* cf = sm_all;
  double* red = sm_all + ((6 * C + B + 1) & ~1);
  for (int i = threadIdx.x; i < 6 * C + B; i += blockDim.x) cf[i] = coef[i];
  __syncthreads();
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  double acc[3] = {0, 0, 0};
  if (p < P) {
    double* col = Yhat + 3 * p;
    double* bt = col + (size_t)(6 * C + B) * ldy;
    double a0 = bt[0], a1 = bt[1], a2 = bt[2];
    for (int c = 0; c < C; ++c) {
      if (!(s[(size_t)c * P + p] > 0)) continue;
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        const double* q = col + (size_t)(6 * c + i) * ldy;
        double f = cf[6 * c + i];
        a0 -= q[0] * f; a1 -= q[1] * f; a2 -= q[2] * f;
      }
    }
    const int j = (int)(p % J);
#pragma unroll
    for (int k = 0; k < BAX_MAX_JB; ++k) {
      int b = joint_bones[(j * BAX_MAX_JB + k) * 2];
      if (b >= 0) {
        const double* q = col + (size_t)(6 * C + b) * ldy;
        double f = cf[6 * C + b];
        a0 -= q[0] * f; a1 -= q[1] * f; a2 -= q[2] * f;
      }
    }
    double V[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) V[k] = Vp[6 * p + k];
    const Chol3r L = chol3r(V, alpha);
    chol3r_bwd(L, a0, a1, a2);                    // p_p
    step_pts[3 * p] = a0; step_pts[3 * p + 1] = a1; step_pts[3 * p + 2] = a2;
    acc[0] = a0 * a0 + a1 * a1 + a2 * a2;
    acc[1] = gp[3 * p] * a0 + gp[3 * p + 1] * a1 + gp[3 * p + 2] * a2;
    chol3r_fwd(L, a0, a1, a2);                    // L^-1 p_p
    bt[0] = a0; bt[1] = a1; bt[2] = a2;
    acc[2] = a0 * a0 + a1 * a1 + a2 * a2;
  }
  block_sum<3>(acc, red);
  if (threadIdx.x == 0) { part[3 * blockIdx.x] = acc[0]; part[3 * blockIdx.x + 1] = acc[1]; part[3 * blockIdx.x + 2] = acc[2]; }
  if (bax_last_block(ticket)) bax_fold_sum<128>(part, gridDim.x, 3, sums_out, red);   // was: k_reduce_partials after this kernel
}

// out[i] = sum_k Yhat[i][k] * Yhat[last][k], split over K chunks (deterministic two-stage sum)
__global__ void __launch_bounds__(256)
k_gemv_last(const double* __restrict__ Yhat, int64_t ldy, int64_t K, int last, int64_t chunk, int rows,
            double* __restrict__ part /*[nchunk][rows]*/) {
  __shared__ double sm[32];
  const int i = blockIdx.x;
  const int64_t k0 = (int64_t)blockIdx.y * chunk;
  int64_t k1 = k0 + chunk; if (k1 > K) k1 = K;
  const double* a = Yhat + (size_t)i * ldy;
  const double* b = Yhat + (size_t)last * ldy;
  double acc[1] = {0.0};
  for (int64_t k = k0 + threadIdx.x; k < k1; k += 256) acc[0] += a[k] * b[k];
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0) part[(size_t)blockIdx.y * rows + i] = acc[0];
}

// rhs2[0 .. 6C+B) = Yhat[0 .. 6C+B) . (L^-1 p)   (the row k_backsub left in place of bt)
static int launch_phi_rhs(bax_handle* h, cudaStream_t s) {
  const int C6 = 6 * h->C, B = h->B;
  int64_t K = 3 * h->P, chunk = 1 << 15;
  int nch = (int)((K + chunk - 1) / chunk);
  double* part2 = h->partials + (size_t)(1 << 21);
  if ((size_t)nch * (C6 + B) > (1u << 21)) { bax_set_error("partials overflow in gemv"); return BAX_E_STATE; }
  k_gemv_last<<<dim3(C6 + B, nch), 256, 0, s>>>(h->Yhat, h->ldy, K, C6 + B, chunk, C6 + B, part2);
  BAX_LAUNCH_CHECK();
  return bax_reduce_partials(h, part2, nch, C6 + B, h->buf[BAX_BUF_RHS2], s);
}

extern "C" int bax_trial_begin(bax_handle* h, double alpha, void* stream) {
  if (!h || !h->Yhat || !(alpha > 0)) { bax_set_error("bax_trial_begin: needs a linearised problem and alpha > 0"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  h->alpha = alpha;
  const int C = h->C;
  {
    BaxTimer t(h, BAX_T_BUILD, s, 1);
    int nb = (int)((h->P + 127) / 128);
    double scale = h->lam2 * 2.0 / (double)h->n_frames_global;
    // Every CTA does the same work and they all start together, so the grid finishes in ceil(CTAs / resident CTAs)
    // rounds: 1661 CTAs on 5 x 148 slots are 2.24 -> 3 rounds, a quarter of the last one idle (ncu: SMs active 85 %).
    // Dealing the cameras of a point over `groups` CTAs shortens the rounds, but every group repeats the point's loads
    // and Cholesky and writes its own rows: measured at 64 cameras 0.474 / 0.413 / 0.446 / 0.493 ms for 1 / 2 / 4 / 8
    // groups (0.456 before the change), at 16 cameras 0.138 / 0.126 / 0.135 / 0.146, at 31 and at 4 cameras flat.
    static const int forced = getenv("BAX_YHAT_GROUPS") ? atoi(getenv("BAX_YHAT_GROUPS")) : 0;   // tuning aid
    const int groups = forced > 0 && forced <= C ? forced : (C >= 16 ? 2 : 1);
    k_build_yhat<<<dim3(nb, groups), 128, C * CAMPACK * sizeof(double), s>>>(h->campack_lin, h->theta_lin + 6 * C, (const double2*)h->y2d,
        h->s2d, h->P, C, h->J, h->B, h->bones, h->joint_bones, h->Lbar, scale, h->Vp, h->gp, alpha, h->Yhat, h->ldy);
    BAX_LAUNCH_CHECK();
  }
  // camera rows: dense SYRK on the FP64 tensor pipe; the B + 1 bone-length / right-hand-side rows ride along
  return bax_syrk_aug(h, h->Yhat, 6 * C, h->B + 1, 3 * h->P, h->ldy, h->buf[BAX_BUF_SCHUR], h->R, s);
}

extern "C" int bax_trial_mid(bax_handle* h, double* step, int need_phi_prime, void* stream) {
  if (!h || !h->Yhat || !step) { bax_set_error("bax_trial_mid: bad argument"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  const int C = h->C, C6 = 6 * C, B = h->B;
  int rc = bax_dense_solve1(h, step, s); if (rc) return rc;      // camera part of the step included
  {
    BaxTimer t(h, BAX_T_BACKSUB, s, 1);
    int nb = (int)((h->P + 127) / 128);
    size_t sm = sizeof(double) * (((C6 + B + 1) & ~1) + 3 * 32);
    double* part = h->partials;
    if ((size_t)nb * 3 > (1u << 21)) { bax_set_error("partials overflow in backsub"); return BAX_E_STATE; }
    k_backsub<<<nb, 128, sm, s>>>(h->s2d, h->P, C, h->J, B, h->joint_bones, h->Vp, h->gp, h->alpha, h->Yhat, h->ldy,
                                  h->coef, step + C6, part, h->buf[BAX_BUF_RHS2] + C6 + B, h->tickets + 3);
    BAX_LAUNCH_CHECK();
    double* rhs2 = h->buf[BAX_BUF_RHS2];
    if (need_phi_prime) {
      rc = launch_phi_rhs(h, s); if (rc) return rc;
    } else {
      BAX_CUDA(cudaMemsetAsync(rhs2, 0, sizeof(double) * (C6 + B), s));
    }
  }
  h->t_launch[BAX_T_BACKSUB] += need_phi_prime ? 2 : 0;
  h->need_phi = need_phi_prime;
  return BAX_OK;
}

// phi'(alpha) on demand: after a bax_trial_mid(.., need_phi_prime = 0, ..) / bax_trial_finish pair whose step turned
// out NOT to be the interior one, the right-hand side of the second reduced solve is formed now (the whitened blocks
// and L^-1 p are still in Yhat); the caller all-reduces BAX_BUF_RHS2_HEAD (frame-sharded runs) and calls
// bax_trial_finish again.  The interior (Gauss-Newton-like) steps — most steps of a converging run — never pay for it.
extern "C" int bax_trial_phi_prime(bax_handle* h, void* stream) {
  if (!h || !h->Yhat) { bax_set_error("bax_trial_phi_prime: no damped solve to extend"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  {
    BaxTimer t(h, BAX_T_BACKSUB, s, 2);
    int rc = launch_phi_rhs(h, s); if (rc) return rc;
  }
  h->need_phi = 1;
  return BAX_OK;
}

extern "C" int bax_trial_finish(bax_handle* h, double* out_host, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  int rc = bax_dense_solve2(h, s); if (rc) return rc;
  BAX_CUDA(cudaMemcpyAsync(h->host_pinned, h->buf[BAX_BUF_STEP], sizeof(double) * 4, cudaMemcpyDeviceToHost, s));
  BAX_CUDA(bax_stream_wait(h, s));
  for (int i = 0; i < 4; ++i) out_host[i] = h->host_pinned[i];
  return BAX_OK;
}
