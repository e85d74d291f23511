// Handle lifetime, problem binding (constant derivations), camera packing, small utilities.
#include "bax_internal.cuh"
#include <cub/device/device_scan.cuh>
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>

static thread_local char g_err[512] = "";
void bax_set_error(const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
extern "C" const char* bax_last_error(void) { return g_err; }
extern "C" int bax_version(void) { return 100; }

// ------------------------------------------------------------------ kernels
// theta -> per-camera pack {K, R, t, M, Mn}.
//   R  = Rodrigues(r)                              (cv2.Rodrigues at reference ba.py:29)
//   M  = (r r^T + (R^T - I) hat(r)) / |r|^2        so that d(R x)/dr = -R hat(x) M     
//   Mn = (r r^T - (R   - I) hat(r)) / |r|^2        so that d(R^T d)/dr = R^T hat(d) Mn
//   r -> 0: R = I, M = Mn = I.
__global__ void k_cam_prep(const double* __restrict__ theta, const double* __restrict__ K, int C,
                           double* __restrict__ pack) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double r0 = theta[3 * c], r1 = theta[3 * c + 1], r2 = theta[3 * c + 2];
  double* o = pack + (size_t)c * CAMPACK;
  for (int i = 0; i < 9; ++i) o[i] = K[9 * c + i];
  for (int i = 0; i < 3; ++i) o[18 + i] = theta[3 * C + 3 * c + i];
  double n2 = r0 * r0 + r1 * r1 + r2 * r2;
  double th = sqrt(n2);
  double R[9], M[9], Mn[9];
  if (th < 2.220446049250313e-16) {
    for (int i = 0; i < 9; ++i) { R[i] = (i % 4 == 0); M[i] = R[i]; Mn[i] = R[i]; }
  } else {
    double k0 = r0 / th, k1 = r1 / th, k2 = r2 / th;
    double cs = cos(th), sn = sin(th), c1 = 1.0 - cs;
    R[0] = cs + c1 * k0 * k0;      R[1] = c1 * k0 * k1 - sn * k2; R[2] = c1 * k0 * k2 + sn * k1;
    R[3] = c1 * k1 * k0 + sn * k2; R[4] = cs + c1 * k1 * k1;      R[5] = c1 * k1 * k2 - sn * k0;
    R[6] = c1 * k2 * k0 - sn * k1; R[7] = c1 * k2 * k1 + sn * k0; R[8] = cs + c1 * k2 * k2;
    double hr[9] = {0, -r2, r1, r2, 0, -r0, -r1, r0, 0};
    double r[3] = {r0, r1, r2};
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        double a = 0, b = 0;   // a = ((R^T - I) hr)_ij, b = ((R - I) hr)_ij
        for (int k = 0; k < 3; ++k) {
          a += (R[3 * k + i] - (i == k)) * hr[3 * k + j];
          b += (R[3 * i + k] - (i == k)) * hr[3 * k + j];
        }
        M[3 * i + j] = (r[i] * r[j] + a) / n2;
        Mn[3 * i + j] = (r[i] * r[j] - b) / n2;
      }
  }
  for (int i = 0; i < 9; ++i) { o[9 + i] = R[i]; o[21 + i] = M[i]; o[30 + i] = Mn[i]; }
  for (int i = 39; i < CAMPACK; ++i) o[i] = 0.0;
}

int bax_launch_cam_prep(bax_handle* h, const double* theta, cudaStream_t s, bool lin) {
  k_cam_prep<<<(h->C + 63) / 64, 64, 0, s>>>(theta, h->K, h->C, lin ? h->campack_lin : h->campack);
  BAX_LAUNCH_CHECK();
  if (lin) h->theta_lin = theta; else h->theta = theta;
  return BAX_OK;
}

// sum per-block partials [nblk][nv] -> out[nv] in a fixed order (deterministic).
__global__ void k_reduce_partials(const double* __restrict__ part, int nblk, int nv, double* __restrict__ out) {
  __shared__ double sm[32];
  int v = blockIdx.x;
  double acc = 0.0;
  for (int b = threadIdx.x; b < nblk; b += blockDim.x) acc += part[(size_t)b * nv + v];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    double x = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : 0.0;
    x = warp_sum(x);
    if (threadIdx.x == 0) out[v] = x;
  }
}
int bax_reduce_partials(bax_handle* h, const double* partials, int nblk, int nv, double* out, cudaStream_t s) {
  k_reduce_partials<<<nv, 256, 0, s>>>(partials, nblk, nv, out);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}

// orientation term constants: valid flag + unit direction per (camera, frame, bone)   ba.py:73-91
__global__ void k_bone_dir_flags(const double* __restrict__ p3d, const double* __restrict__ s3d,
                                 const int32_t* __restrict__ bones, int C, int N, int J3, int B,
                                 int32_t* __restrict__ flag) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)N * B) return;
  int n = (int)(i / B), b = (int)(i % B);
  int ja = bones[2 * b], jb = bones[2 * b + 1];
  int ok = 1;
  for (int c = 0; c < C; ++c) {
    size_t base = ((size_t)c * N + n) * J3;
    if (!(s3d[base + ja] > 0) || !(s3d[base + jb] > 0)) { ok = 0; break; }
    const double* a = p3d + (base + ja) * 3;
    const double* bb = p3d + (base + jb) * 3;
    double d0 = a[0] - bb[0], d1 = a[1] - bb[1], d2 = a[2] - bb[2];
    double nn = sqrt(d0 * d0 + d1 * d1 + d2 * d2);
    d0 /= nn; d1 /= nn; d2 /= nn;
    if (isnan(d0) || isnan(d1) || isnan(d2)) { ok = 0; break; }
  }
  flag[i] = ok;
}
__global__ void k_bone_dir_scatter(const double* __restrict__ p3d, const int32_t* __restrict__ bones,
                                   const int32_t* __restrict__ flag, const int64_t* __restrict__ pos,
                                   int C, int N, int J3, int B, int64_t Nv, double* __restrict__ dhat) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)N * B || !flag[i]) return;
  int n = (int)(i / B), b = (int)(i % B);
  int ja = bones[2 * b], jb = bones[2 * b + 1];
  int64_t row = pos[i];
  for (int c = 0; c < C; ++c) {
    size_t base = ((size_t)c * N + n) * J3;
    const double* a = p3d + (base + ja) * 3;
    const double* bb = p3d + (base + jb) * 3;
    double d0 = a[0] - bb[0], d1 = a[1] - bb[1], d2 = a[2] - bb[2];
    double nn = sqrt(d0 * d0 + d1 * d1 + d2 * d2);
    double* o = dhat + ((size_t)c * Nv + row) * 3;
    o[0] = d0 / nn; o[1] = d1 / nn; o[2] = d2 / nn;
  }
}
__global__ void k_flag_to_i64(const int32_t* __restrict__ f, int64_t n, int64_t* __restrict__ o) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = f[i];
}
// visible observations per (camera, 256-point block)
__global__ void k_count_visible(const double* __restrict__ s2d, int64_t P, int64_t* __restrict__ cnt) {
  int c = blockIdx.y;
  int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x;
  int vis = (p < P) && (s2d[(size_t)c * P + p] > 0);
  int tot = __syncthreads_count(vis);
  if (threadIdx.x == 0) cnt[(size_t)c * gridDim.x + blockIdx.x] = tot;
}

// ------------------------------------------------------------------ handle
extern "C" int bax_create(int device, bax_handle** out) {
  if (!out) { bax_set_error("bax_create: null out"); return BAX_E_ARG; }
  BAX_CUDA(cudaSetDevice(device));
  bax_handle* h = (bax_handle*)calloc(1, sizeof(bax_handle));
  h->device = device;
  BAX_CUDA(cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device));   // (cudaGetDeviceProperties costs ~1 ms)
  { const char* e = getenv("BAX_DENSE_COOP"); h->dense_coop = e ? atoi(e) : 1; }
  { const char* e = getenv("BAX_RJ_VARIANT"); h->rj_variant = e ? atoi(e) : -1; }
  BAX_CUDA(cudaEventCreate(&h->ev0));
  BAX_CUDA(cudaEventCreate(&h->ev1));
  BAX_CUDA(cudaEventCreateWithFlags(&h->ev_block, cudaEventBlockingSync | cudaEventDisableTiming));
  BAX_CUDA(cudaMallocHost((void**)&h->host_pinned, (64 + 4096) * sizeof(double)));   // [0,64): trial scalars; [64,..): linearisation scalars
  {  // freed blocks stay in the device's default pool: the next handle of a sweep gets them back without a driver call
    cudaMemPool_t pool; unsigned long long keep = ~0ULL;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  *out = h;
  return BAX_OK;
}

// Device memory of a handle is stream-ordered (cudaMallocAsync / cudaFreeAsync on the stream of the last
// bax_problem_set): creating and destroying a handle per problem of a multi-sequence sweep costs no cudaMalloc /
// cudaFree, and — unlike cudaFree — releasing one problem does not synchronise the streams of the others.
cudaError_t bax_dev_alloc(bax_handle* h, void** p, size_t bytes) { return cudaMallocAsync(p, bytes ? bytes : 8, h->mem_stream); }
void bax_dev_free(bax_handle* h, void* p) {
  if (!p) return;
  if (cudaFreeAsync(p, h->mem_stream) != cudaSuccess) { cudaGetLastError(); cudaFree(p); }   // the stream no longer exists
}

static void free_problem(bax_handle* h) {
  void* ptrs[] = {(void*)h->K, h->bones, h->joint_bones, h->dhat, h->blockbase, h->campack, h->campack_lin, h->Vp, h->gp,
                  h->Hcc, h->gc, h->Lbar, h->rb, h->lin_stage, h->Yhat, h->Gv, h->syrk_part, h->lin_part, h->tickets, h->dense_ws, h->coef,
                  h->partials, h->saved.Vp, h->saved.gp, h->saved.lin, h->saved.Lbar, h->saved.rb, h->saved.campack_lin};
  for (void* p : ptrs) bax_dev_free(h, p);
  for (int i = 0; i < 8; ++i) { bax_dev_free(h, h->buf[i]); h->buf[i] = nullptr; }
  h->K = nullptr; h->bones = nullptr; h->joint_bones = nullptr; h->dhat = nullptr; h->blockbase = nullptr;
  h->campack = nullptr; h->campack_lin = nullptr; h->Vp = h->gp = h->Hcc = h->gc = h->Lbar = h->rb = nullptr;
  h->Yhat = h->Gv = h->syrk_part = h->dense_ws = h->coef = h->partials = h->lin_stage = nullptr;
  h->saved.Vp = h->saved.gp = h->saved.lin = h->saved.Lbar = h->saved.rb = h->saved.campack_lin = nullptr; h->saved.valid = 0;
  h->syrk_part_bytes = 0;
  h->lin_part = nullptr; h->lin_part_bytes = 0; h->tickets = nullptr;
}

extern "C" int bax_set_blocking_sync(bax_handle* h, int on) {
  if (!h) { bax_set_error("bax_set_blocking_sync: bad argument"); return BAX_E_ARG; }
  h->blocking_sync = on ? 1 : 0;
  return BAX_OK;
}

extern "C" int bax_destroy(bax_handle* h) {
  if (!h) return BAX_OK;
  cudaSetDevice(h->device);
  free_problem(h);
  cudaEventDestroy(h->ev0); cudaEventDestroy(h->ev1); cudaEventDestroy(h->ev_block);
  cudaFreeHost(h->host_pinned);
  free(h);
  return BAX_OK;
}

extern "C" int bax_problem_set(bax_handle* h, int C, int N, int J, int J3, int B, const double* K,
                               const double* y2d, const double* s2d, const double* p3d, const double* s3d,
                               const int32_t* bones_host, double lambda1, double lambda2, void* stream) {
  if (!h || C < 1 || N < 1 || J < 1 || B < 0 || B >= BAX_MAX_B || !K || !y2d || !s2d || (B > 0 && !bones_host)) {
    bax_set_error("bax_problem_set: bad argument"); return BAX_E_ARG;
  }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  free_problem(h);
  h->mem_stream = s;
  h->C = C; h->N = N; h->J = J; h->J3 = J3; h->B = B; h->P = (int64_t)N * J;
  h->lam1 = lambda1; h->lam2 = lambda2; h->n_frames_global = N;
  h->y2d = y2d; h->s2d = s2d;
  h->R = 6 * C + B + 1;
  for (int b = 0; b < B; ++b) {
    int a0 = bones_host[2 * b], a1 = bones_host[2 * b + 1];
    if (a0 < 0 || a1 < 0 || a0 >= J || a1 >= J || (p3d && (a0 >= J3 || a1 >= J3))) {
      bax_set_error("bax_problem_set: bone %d = (%d,%d) out of range for J=%d / J3=%d", b, a0, a1, J, J3);
      return BAX_E_ARG;
    }
    h->bones_host[2 * b] = a0; h->bones_host[2 * b + 1] = a1;
  }
  double* Kd; BAX_ALLOC(h, Kd, sizeof(double) * 9 * C);
  BAX_CUDA(cudaMemcpyAsync(Kd, K, sizeof(double) * 9 * C, cudaMemcpyDeviceToDevice, s));
  h->K = Kd;
  BAX_ALLOC(h, h->bones, sizeof(int32_t) * 2 * (B > 0 ? B : 1));
  BAX_CUDA(cudaMemcpyAsync(h->bones, h->bones_host, sizeof(int32_t) * 2 * B, cudaMemcpyHostToDevice, s));
  {  // joint -> incident bones
    size_t n = (size_t)J * BAX_MAX_JB * 2;
    int32_t* jb = (int32_t*)malloc(n * sizeof(int32_t));
    for (size_t i = 0; i < n; ++i) jb[i] = -1;
    for (int b = 0; b < B; ++b)
      for (int e = 0; e < 2; ++e) {
        int j = h->bones_host[2 * b + e], k = 0;
        while (k < BAX_MAX_JB && jb[((size_t)j * BAX_MAX_JB + k) * 2] >= 0) ++k;
        if (k == BAX_MAX_JB) { free(jb); bax_set_error("joint %d touches more than %d bones", j, BAX_MAX_JB); return BAX_E_ARG; }
        jb[((size_t)j * BAX_MAX_JB + k) * 2] = b;
        jb[((size_t)j * BAX_MAX_JB + k) * 2 + 1] = e == 0 ? 1 : -1;
      }
    BAX_ALLOC(h, h->joint_bones, n * sizeof(int32_t));
    BAX_CUDA(cudaMemcpyAsync(h->joint_bones, jb, n * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    BAX_CUDA(bax_stream_wait(h, s));
    free(jb);
  }
  // ---- orientation-term constants
  h->Nv = 0;
  if (p3d && s3d && B > 0) {
    int64_t NB = (int64_t)N * B;
    int32_t* flag; int64_t *f64, *pos;
    BAX_ALLOC(h, flag, NB * sizeof(int32_t));
    BAX_ALLOC(h, f64, NB * sizeof(int64_t));
    BAX_ALLOC(h, pos, NB * sizeof(int64_t));
    int nb = (int)((NB + 255) / 256);
    k_bone_dir_flags<<<nb, 256, 0, s>>>(p3d, s3d, h->bones, C, N, J3, B, flag);
    k_flag_to_i64<<<nb, 256, 0, s>>>(flag, NB, f64);
    size_t tb = 0; cub::DeviceScan::ExclusiveSum(nullptr, tb, f64, pos, NB, s);
    void* tmp; BAX_ALLOC(h, tmp, tb ? tb : 8);
    cub::DeviceScan::ExclusiveSum(tmp, tb, f64, pos, NB, s);
    int64_t lastp = 0; int32_t lastf = 0;
    BAX_CUDA(cudaMemcpyAsync(&lastp, pos + NB - 1, 8, cudaMemcpyDeviceToHost, s));
    BAX_CUDA(cudaMemcpyAsync(&lastf, flag + NB - 1, 4, cudaMemcpyDeviceToHost, s));
    BAX_CUDA(bax_stream_wait(h, s));
    h->Nv = lastp + lastf;
    if (h->Nv > 0) {
      BAX_ALLOC(h, h->dhat, sizeof(double) * 3 * C * h->Nv);
      k_bone_dir_scatter<<<nb, 256, 0, s>>>(p3d, h->bones, flag, pos, C, N, J3, B, h->Nv, h->dhat);
      BAX_LAUNCH_CHECK();
    }
    BAX_CUDA(bax_stream_wait(h, s));
    bax_dev_free(h, flag); bax_dev_free(h, f64); bax_dev_free(h, pos); bax_dev_free(h, tmp);
  }
  // ---- residual row order: exclusive scan of visible counts over (camera, 256-point block)
  h->nblk_obs = (int)((h->P + 255) / 256);
  {
    int64_t nb = (int64_t)C * h->nblk_obs;
    int64_t* cnt; BAX_ALLOC(h, cnt, nb * sizeof(int64_t));
    BAX_ALLOC(h, h->blockbase, nb * sizeof(int64_t));
    k_count_visible<<<dim3(h->nblk_obs, C), 256, 0, s>>>(s2d, h->P, cnt);
    size_t tb = 0; cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt, h->blockbase, nb, s);
    void* tmp; BAX_ALLOC(h, tmp, tb ? tb : 8);
    cub::DeviceScan::ExclusiveSum(tmp, tb, cnt, h->blockbase, nb, s);
    int64_t lb = 0, lc = 0;
    BAX_CUDA(cudaMemcpyAsync(&lb, h->blockbase + nb - 1, 8, cudaMemcpyDeviceToHost, s));
    BAX_CUDA(cudaMemcpyAsync(&lc, cnt + nb - 1, 8, cudaMemcpyDeviceToHost, s));
    BAX_CUDA(bax_stream_wait(h, s));
    h->n_vis = lb + lc;
    bax_dev_free(h, cnt); bax_dev_free(h, tmp);
  }
  // ---- state buffers
  const int C6 = 6 * C;
  BAX_ALLOC(h, h->campack, sizeof(double) * CAMPACK * C);
  BAX_ALLOC(h, h->campack_lin, sizeof(double) * CAMPACK * C);
  BAX_ALLOC(h, h->Vp, sizeof(double) * 6 * h->P);
  BAX_ALLOC(h, h->gp, sizeof(double) * 3 * h->P);
  BAX_ALLOC(h, h->Lbar, sizeof(double) * BAX_MAX_B);
  BAX_ALLOC(h, h->rb, sizeof(double) * BAX_MAX_B);
  BAX_ALLOC(h, h->tickets, 16 * sizeof(unsigned));
  BAX_CUDA(cudaMemsetAsync(h->tickets, 0, 16 * sizeof(unsigned), s));
  BAX_CUDA(cudaMemsetAsync(h->Lbar, 0, sizeof(double) * BAX_MAX_B, s));
  BAX_CUDA(cudaMemsetAsync(h->rb, 0, sizeof(double) * BAX_MAX_B, s));
  BAX_ALLOC(h, h->coef, sizeof(double) * (C6 + BAX_MAX_B));
  BAX_ALLOC(h, h->lin_stage, sizeof(double) * 4096);
  BAX_ALLOC(h, h->dense_ws, sizeof(double) * ((size_t)C6 * C6 + (size_t)BAX_MAX_B * C6 * 2 + 4 * C6 + 1024));
  h->partials_bytes = sizeof(double) * (size_t)(1 << 22);
  BAX_ALLOC(h, h->partials, h->partials_bytes);
  h->buf_count[BAX_BUF_BONE] = B + 1;
  h->buf_count[BAX_BUF_COST] = B + 2;
  h->buf_count[BAX_BUF_LIN] = (int64_t)C6 * C6 + C6 + B + 2;
  h->buf_count[BAX_BUF_SCHUR] = (int64_t)h->R * h->R;
  h->buf_count[BAX_BUF_RHS2] = C6 + B + 4;
  h->buf_count[BAX_BUF_STEP] = 4;
  h->buf_count[BAX_BUF_NORM] = 4;
  h->buf_count[BAX_BUF_GMAX] = 2;
  for (int i = 0; i < 8; ++i) {
    BAX_ALLOC(h, h->buf[i], sizeof(double) * (h->buf_count[i] > 0 ? h->buf_count[i] : 1));
    BAX_CUDA(cudaMemsetAsync(h->buf[i], 0, sizeof(double) * (h->buf_count[i] > 0 ? h->buf_count[i] : 1), s));
  }
  // Yhat / Gv are allocated lazily by the first linearise (bax_lm.cu)
  h->ldy = ((3 * h->P + 31) / 32) * 32;
  h->ldv = ((h->Nv + 31) / 32) * 32;
  BAX_CUDA(bax_stream_wait(h, s));
  return BAX_OK;
}

// Speculative linearisation (a trust-region step is linearised at the trial point straight away; a rejected step
// restores the previous linearisation): push swaps the current state into the saved slot, pop swaps it back.
static void swap_lin_state(bax_handle* h) {
  auto sw = [](double*& a, double*& b) { double* t = a; a = b; b = t; };
  sw(h->Vp, h->saved.Vp); sw(h->gp, h->saved.gp); sw(h->buf[BAX_BUF_LIN], h->saved.lin);
  sw(h->Lbar, h->saved.Lbar); sw(h->rb, h->saved.rb); sw(h->campack_lin, h->saved.campack_lin);
  const double* t = h->theta_lin; h->theta_lin = h->saved.theta_lin; h->saved.theta_lin = t;
  double c = h->cost; h->cost = h->saved.cost; h->saved.cost = c;
}
extern "C" int bax_linearize_push(bax_handle* h) {
  if (!h || !h->campack) { bax_set_error("bax_linearize_push: no problem"); return BAX_E_ARG; }
  BAX_CUDA(cudaSetDevice(h->device));
  if (!h->saved.Vp) {
    BAX_ALLOC(h, h->saved.Vp, sizeof(double) * 6 * h->P);
    BAX_ALLOC(h, h->saved.gp, sizeof(double) * 3 * h->P);
    BAX_ALLOC(h, h->saved.lin, sizeof(double) * h->buf_count[BAX_BUF_LIN]);
    BAX_ALLOC(h, h->saved.Lbar, sizeof(double) * BAX_MAX_B);
    BAX_ALLOC(h, h->saved.rb, sizeof(double) * BAX_MAX_B);
    BAX_ALLOC(h, h->saved.campack_lin, sizeof(double) * CAMPACK * h->C);
    BAX_CUDA(cudaStreamSynchronize(h->mem_stream));
  }
  swap_lin_state(h);
  h->saved.valid = 1;
  return BAX_OK;
}
extern "C" int bax_linearize_pop(bax_handle* h) {
  if (!h || !h->saved.valid) { bax_set_error("bax_linearize_pop: nothing pushed"); return BAX_E_STATE; }
  swap_lin_state(h);
  h->saved.valid = 0;
  return BAX_OK;
}

extern "C" int bax_observations_rebind(bax_handle* h, const double* y2d, const double* s2d) {
  if (!h || !h->campack || !y2d || !s2d) { bax_set_error("bax_observations_rebind: bad argument / no problem"); return BAX_E_ARG; }
  h->y2d = y2d; h->s2d = s2d;
  return BAX_OK;
}

extern "C" int bax_problem_dims(bax_handle* h, int64_t* d) {
  if (!h || !d || !h->campack) { bax_set_error("bax_problem_dims: no problem bound"); return BAX_E_STATE; }
  d[0] = h->n_vis; d[1] = h->Nv; d[2] = 2 * h->n_vis + h->Nv + h->B; d[3] = 6 * (int64_t)h->C + 3 * h->P;
  d[4] = h->R; d[5] = h->P;
  return BAX_OK;
}

extern "C" int bax_reduce_buffer(bax_handle* h, int which, double** p, int64_t* n) {
  if (!h || which < 0 || which > BAX_BUF_RHS2_HEAD || !h->campack) { bax_set_error("bax_reduce_buffer: bad argument"); return BAX_E_ARG; }
  if (which == BAX_BUF_RHS2_HEAD) { *p = h->buf[BAX_BUF_RHS2]; *n = 6 * h->C + h->B; return BAX_OK; }   // without the already-summed tail
  *p = h->buf[which]; *n = h->buf_count[which];
  return BAX_OK;
}
extern "C" int bax_set_global_frames(bax_handle* h, int64_t n) {
  if (!h || n < 1) { bax_set_error("bax_set_global_frames: bad argument"); return BAX_E_ARG; }
  h->n_frames_global = n; return BAX_OK;
}
extern "C" int bax_timing_enable(bax_handle* h, int on) { h->timing = on; return BAX_OK; }
extern "C" int bax_timing_reset(bax_handle* h) {
  for (int i = 0; i < BAX_T_COUNT; ++i) { h->t_ms[i] = 0; h->t_launch[i] = 0; }
  return BAX_OK;
}
extern "C" int bax_timing_read(bax_handle* h, double* ms, int64_t* nl) {
  for (int i = 0; i < BAX_T_COUNT; ++i) { if (ms) ms[i] = h->t_ms[i]; if (nl) nl[i] = h->t_launch[i]; }
  return BAX_OK;
}
