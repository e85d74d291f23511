// Internal declarations shared by the translation units of libbax.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>
#include "../../include/bax.h"

#define BAX_MAX_B 16          // bones (reference: 12, util.py:129-145)
#define BAX_MAX_JB 4          // bones incident to one joint (reference skeleton: <= 3)
#ifndef BAX_PF
#define BAX_PF 2              // cameras whose (score, 2-D point) loads are in flight ahead of the one worked on (4 measured no faster)
#endif
#define CAMPACK 48            // doubles per camera staged in shared memory (padded to 16 B multiples)
// CamPack layout: K[0..8] R[9..17] t[18..20] M[21..29] Mn[30..38]   (M, Mn: see k_cam_prep)

void bax_set_error(const char* fmt, ...);
#define BAX_CUDA(call)                                                                     \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      bax_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));  \
      return BAX_E_CUDA;                                                                   \
    }                                                                                      \
  } while (0)
#define BAX_LAUNCH_CHECK() BAX_CUDA(cudaGetLastError())
struct bax_handle;
cudaError_t bax_dev_alloc(bax_handle* h, void** p, size_t bytes);   // stream-ordered (bax_problem.cu)
void bax_dev_free(bax_handle* h, void* p);
#define BAX_ALLOC(h, ptr, bytes) BAX_CUDA(bax_dev_alloc((h), (void**)&(ptr), (bytes)))

struct bax_handle {
  int device;
  int sm_count;
  // ---- problem
  int C, N, J, J3, B;
  int64_t P;                 // N*J points
  double lam1, lam2;
  int64_t n_frames_global;
  int blocking_sync;         // host waits sleep on an event instead of spinning (sweeps with many worker threads)
  cudaEvent_t ev_block;
  void* comm;                // ncclComm_t of a frame-sharded run (borrowed; bax_allreduce_attach) or NULL
  int64_t n_collectives;     // collectives issued by the native driver since bax_create
  const double* K;           // [C][9]      (copied)
  const double* y2d;         // [C][P][2]   (caller-owned)
  const double* s2d;         // [C][P]      (caller-owned)
  int32_t bones_host[BAX_MAX_B * 2];
  int32_t* bones;            // dev [B][2]
  int32_t* joint_bones;      // dev [J][BAX_MAX_JB][2] : (bone index, +1/-1), bone = -1 terminates
  double* dhat;              // [C][Nv][3] unit camera-frame bone directions of valid rows
  int64_t Nv;
  int64_t n_vis;
  int nblk_obs;              // ceil(P/256)
  int64_t* blockbase;        // [C][nblk_obs] exclusive scan of visible counts (residual row order)
  // ---- theta-dependent
  const double* theta;       // last bound theta (device, caller-owned)
  double* campack;           // [C][CAMPACK]  for residual / cost evaluations
  const double* theta_lin;   // theta of the current linearisation (trial steps refer to it)
  double* campack_lin;       // [C][CAMPACK]  cameras at theta_lin
  // ---- linearisation state
  double* Vp;                // [P][6]  sum_c Jpt^T Jpt (xx,xy,xz,yy,yz,zz)
  double* gp;                // [P][3]  gradient wrt points
  double* Hcc;               // [6C][6C] camera block (internal camera-major order: 6c+{rx,ry,rz,tx,ty,tz})
  double* gc;                // [6C]
  double* Lbar;              // [B] mean bone lengths (global)
  double* rb;                // [B] lambda2 * Var_b
  double cost;
  // saved copy of the linearisation (bax_linearize_push / _pop): the same fields, swapped by pointer
  struct { double *Vp, *gp, *lin, *Lbar, *rb, *campack_lin; const double* theta_lin; double cost; int valid; } saved;
  // ---- trial state
  double alpha;
  int need_phi;               // last trial asked for phi'(alpha)
  int R;                     // 6C + B + 1 augmented rows
  int64_t ldy;               // leading dimension of Yhat (3P padded)
  double* Yhat;              // [R][ldy]
  double* Gv;                // [(3C+1)][ldv] orientation-row Jacobian (transposed) + residual row
  int64_t ldv;
  double* syrk_part;         // split-K partial tiles
  size_t syrk_part_bytes;
  double* lin_part;          // per-chunk camera sums of k_lin_fused [chunks][C][28]
  int lin_staged;            // k_gp_bone's last block already gathered the scalars bax_linearize_finish reads back
  unsigned* tickets;         // [16] arrival counters of the kernels that fold their own partials (zero between launches)
  size_t lin_part_bytes;
  double* dense_ws;          // S factor (6C x 6C), Z (B x 6C), LT (B x B), pc, w, misc
  double* coef;              // [6C + B] coefficients for back-substitution
  double* lin_stage;         // [<= 4096] scalars of a linearisation gathered for one device -> host copy
  // ---- reduce buffers (see bax_reduce_buffer)
  double* buf[8];
  int64_t buf_count[8];
  double* partials;          // generic per-block partial sums
  size_t partials_bytes;
  double* host_pinned;       // small pinned staging for scalars
  cudaStream_t mem_stream;   // stream of the last bax_problem_set: the handle's device memory is allocated and freed in its order
  // ---- timing
  int dense_coop;            // 1: cooperative multi-CTA reduced solve (env BAX_DENSE_COOP=0 selects the single-CTA kernel)
  int rj_variant;            // tuning knob (env BAX_RJ_VARIANT): -1 = auto [default: 5 or 4]; 0 = per-warp staged LSU stores; 1/2/3 = CTA-staged TMA bulk stores (256x4, 128x8, 256x3); 4..8 = persistent TMA pipeline (stages x CTAs/SM: 2x3, 4x2, 3x2, 5x2, 8x1)
  int timing;
  cudaEvent_t ev0, ev1;
  double t_ms[BAX_T_COUNT];
  int64_t t_launch[BAX_T_COUNT];
};


struct BaxTimer {            // RAII CUDA-event timer (only when h->timing) around a kernel family
  bax_handle* h; int which; cudaStream_t s; int nl;
  BaxTimer(bax_handle* h_, int w, cudaStream_t s_, int launches) : h(h_), which(w), s(s_), nl(launches) {
    if (h->timing) cudaEventRecord(h->ev0, s);
  }
  ~BaxTimer() {
    h->t_launch[which] += nl;
    if (h->timing) {
      cudaEventRecord(h->ev1, s); cudaEventSynchronize(h->ev1);
      float ms = 0; cudaEventElapsedTime(&ms, h->ev0, h->ev1); h->t_ms[which] += ms;
    }
  }
};

// ---------------------------------------------------------------- device helpers
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- kernels that fold their own per-block partials (saves the k_reduce_partials / k_reduce_max launch that followed
// them: a 4-camera problem is bound by the number of launches, not by their work).  Every block writes its partial as
// before, then takes a ticket; the block that draws the last one sums ALL partials in the fixed order k_reduce_partials
// uses (thread-strided over the blocks, warp tree, warps in order), so the result does not depend on which block is last.
// The ticket returns to zero for the next launch on the stream.
__device__ __forceinline__ bool bax_last_block(unsigned* ticket) {
  __shared__ bool last;
  __threadfence();                               // this block's partials are visible before its ticket is counted
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(ticket, 1u);
    last = (t == gridDim.x * gridDim.y - 1);
    if (last) *ticket = 0u;
  }
  __syncthreads();
  if (last) __threadfence();
  return last;
}
template <int NT>
__device__ __forceinline__ void bax_fold_sum(const double* part, int nblk, int nv, double* out, double* sm /*[32]*/) {
  for (int v = 0; v < nv; ++v) {
    double acc = 0.0;
    for (int b = threadIdx.x; b < nblk; b += NT) acc += __ldcg(part + (size_t)b * nv + v);
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
      double x = threadIdx.x < NT / 32 ? sm[threadIdx.x] : 0.0;
      x = warp_sum(x);
      if (threadIdx.x == 0) out[v] = x;
    }
    __syncthreads();
  }
}
template <int NT>
__device__ __forceinline__ void bax_fold_max(const double* v, int n, double* out, double* sm /*[32]*/) {
  double mx = 0.0;
  for (int i = threadIdx.x; i < n; i += NT) mx = fmax(mx, __ldcg(v + i));
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x == 0) { for (int w = 0; w < NT / 32; ++w) mx = fmax(mx, sm[w]); *out = mx; }
  __syncthreads();
}

// Deterministic block sum of NV values per thread; result valid in thread 0 (array `v`).
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* smem /* >= NV*32 doubles */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) smem[i * 32 + warp] = v[i];
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      double x = lane < nw ? smem[i * 32 + lane] : 0.0;
      v[i] = warp_sum(x);
    }
  }
}

// Projection residual and Jacobian blocks of one observation.  cam: CamPack in shared memory.
//   e = sqrt(2) s (y - pi(K (R x + t)))                         reference ba.py:44-61
//   Jp = de/dx (2x3),  Jt = de/dt (2x3),  Jr = de/drvec (2x3)   SURVEY.md §3.3
struct ObsLin { double e0, e1; double jp[2][3]; double jt[2][3]; double jr[2][3]; };

__device__ __forceinline__ void obs_residual(const double* cam, double x0, double x1, double x2,
                                             double u, double v, double s, double& e0, double& e1) {
  const double* K = cam; const double* R = cam + 9; const double* t = cam + 18;
  double P0 = R[0] * x0 + R[1] * x1 + R[2] * x2 + t[0];
  double P1 = R[3] * x0 + R[4] * x1 + R[5] * x2 + t[1];
  double P2 = R[6] * x0 + R[7] * x1 + R[8] * x2 + t[2];
  double q0 = K[0] * P0 + K[1] * P1 + K[2] * P2;
  double q1 = K[3] * P0 + K[4] * P1 + K[5] * P2;
  double q2 = K[6] * P0 + K[7] * P1 + K[8] * P2;
  double w = 1.4142135623730951 * s;
  e0 = (u - q0 / q2) * w;
  e1 = (v - q1 / q2) * w;
}

// EXACT_DIV: the materialised residual / Jacobian of k_rj keeps the reference's own expression (u - q0 / q2); the
// normal-equation kernels use one reciprocal instead of three divisions (differs in the last bit only).
template <bool WITH_CAM, bool EXACT_DIV = false>
__device__ __forceinline__ void obs_linearize(const double* cam, double x0, double x1, double x2,
                                              double u, double v, double s, ObsLin& o) {
  const double* K = cam; const double* R = cam + 9; const double* t = cam + 18; const double* M = cam + 21;
  double P0 = R[0] * x0 + R[1] * x1 + R[2] * x2 + t[0];
  double P1 = R[3] * x0 + R[4] * x1 + R[5] * x2 + t[1];
  double P2 = R[6] * x0 + R[7] * x1 + R[8] * x2 + t[2];
  double q0 = K[0] * P0 + K[1] * P1 + K[2] * P2;
  double q1 = K[3] * P0 + K[4] * P1 + K[5] * P2;
  double q2 = K[6] * P0 + K[7] * P1 + K[8] * P2;
  double w = 1.4142135623730951 * s;
  double yh0, yh1, f;
  if (EXACT_DIV) { yh0 = q0 / q2; yh1 = q1 / q2; f = -w / q2; }
  else {
    const double iq2 = 1.0 / q2;                // one FP64 division (~25 pipe instructions) instead of three
    yh0 = q0 * iq2; yh1 = q1 * iq2; f = -w * iq2;
  }
  o.e0 = (u - yh0) * w;
  o.e1 = (v - yh1) * w;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    o.jt[0][k] = f * (K[k] - yh0 * K[6 + k]);
    o.jt[1][k] = f * (K[3 + k] - yh1 * K[6 + k]);
  }
#pragma unroll
  for (int a = 0; a < 2; ++a) {
#pragma unroll
    for (int k = 0; k < 3; ++k)
      o.jp[a][k] = o.jt[a][0] * R[k] + o.jt[a][1] * R[3 + k] + o.jt[a][2] * R[6 + k];
  }
  if (WITH_CAM) {
#pragma unroll
    for (int a = 0; a < 2; ++a) {   // Jr_a = (x cross jp_a)^T M
      double c0 = x1 * o.jp[a][2] - x2 * o.jp[a][1];
      double c1 = x2 * o.jp[a][0] - x0 * o.jp[a][2];
      double c2 = x0 * o.jp[a][1] - x1 * o.jp[a][0];
#pragma unroll
      for (int k = 0; k < 3; ++k) o.jr[a][k] = c0 * M[k] + c1 * M[3 + k] + c2 * M[6 + k];
    }
  }
}

// 3x3 SPD Cholesky of D = V + alpha I (V packed xx,xy,xz,yy,yz,zz); returns inverse-ready factors.
struct Chol3 { double l00, l10, l11, l20, l21, l22; };
__device__ __forceinline__ Chol3 chol3(const double* V, double alpha) {
  Chol3 c;
  c.l00 = sqrt(V[0] + alpha);
  c.l10 = V[1] / c.l00;
  c.l20 = V[2] / c.l00;
  c.l11 = sqrt(V[3] + alpha - c.l10 * c.l10);
  c.l21 = (V[4] - c.l20 * c.l10) / c.l11;
  c.l22 = sqrt(V[5] + alpha - c.l20 * c.l20 - c.l21 * c.l21);
  return c;
}
__device__ __forceinline__ void chol3_fwd(const Chol3& c, double& a, double& b, double& d) {  // L^-1 v
  a = a / c.l00;
  b = (b - c.l10 * a) / c.l11;
  d = (d - c.l20 * a - c.l21 * b) / c.l22;
}
__device__ __forceinline__ void chol3_bwd(const Chol3& c, double& a, double& b, double& d) {  // L^-T v
  d = d / c.l22;
  b = (b - c.l21 * d) / c.l11;
  a = (a - c.l10 * b - c.l20 * d) / c.l00;
}

// The same factorisation keeping reciprocal diagonals: the damped solve applies L^-1 to 6 rows per (camera, point) —
// three divisions each in chol3_fwd, more FP64-pipe work than the projection Jacobian itself.
struct Chol3r { double i00, l10, i11, l20, l21, i22; };
__device__ __forceinline__ Chol3r chol3r(const double* V, double alpha) {
  Chol3r c;
  const double l00 = sqrt(V[0] + alpha);
  c.i00 = 1.0 / l00;
  c.l10 = V[1] * c.i00;
  c.l20 = V[2] * c.i00;
  const double l11 = sqrt(V[3] + alpha - c.l10 * c.l10);
  c.i11 = 1.0 / l11;
  c.l21 = (V[4] - c.l20 * c.l10) * c.i11;
  c.i22 = 1.0 / sqrt(V[5] + alpha - c.l20 * c.l20 - c.l21 * c.l21);
  return c;
}
__device__ __forceinline__ void chol3r_fwd(const Chol3r& c, double& a, double& b, double& d) {  // L^-1 v
  a = a * c.i00;
  b = (b - c.l10 * a) * c.i11;
  d = (d - c.l20 * a - c.l21 * b) * c.i22;
}
__device__ __forceinline__ void chol3r_bwd(const Chol3r& c, double& a, double& b, double& d) {  // L^-T v
  d = d * c.i22;
  b = (b - c.l21 * d) * c.i11;
  a = (a - c.l10 * b - c.l20 * d) * c.i00;
}

// ---------------------------------------------------------------- cross-TU host entry points
int bax_launch_cam_prep(bax_handle* h, const double* theta, cudaStream_t s, bool lin = false);
int bax_syrk(bax_handle* h, const double* X, int rows, int64_t K, int64_t ldx, double* out /*[rows][rows]*/,
             cudaStream_t s);
int bax_syrk_aug(bax_handle* h, const double* X, int rows, int ne, int64_t K, int64_t ldx, double* out /*[rows+ne][ldo]*/, int ldo,
                 cudaStream_t s);
int bax_dense_solve1(bax_handle* h, double* step, cudaStream_t s);
int bax_dense_solve2(bax_handle* h, cudaStream_t s);
int bax_scatter_cam_step(bax_handle* h, double* step, cudaStream_t s);
int bax_bone_pass1(bax_handle* h, const double* theta, cudaStream_t s);
int bax_bone_pass2(bax_handle* h, const double* theta, double* out, cudaStream_t s);
int bax_reduce_partials(bax_handle* h, const double* partials, int nblk, int nv, double* out, cudaStream_t s);
// Host wait for the call's stream: a spinning cudaStreamSynchronize, or — bax_set_blocking_sync — a sleep on a
// blocking-sync event, which frees the core (a sweep's 12 spinning workers per GPU starved each other: DESIGN.md §6).
inline cudaError_t bax_stream_wait(bax_handle* h, cudaStream_t s) {
  if (!h->blocking_sync) return cudaStreamSynchronize(s);
  cudaError_t e = cudaEventRecord(h->ev_block, s);
  return e != cudaSuccess ? e : cudaEventSynchronize(h->ev_block);
}
int bax_allreduce(bax_handle* h, int which, bool op_max, cudaStream_t s);           // bax_comm.cu; no-op without a communicator
int bax_allreduce_pair(bax_handle* h, int which_sum, int which_max, cudaStream_t s);
