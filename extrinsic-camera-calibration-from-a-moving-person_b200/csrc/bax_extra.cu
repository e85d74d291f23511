// Batched weighted triangulation (BA initialisation) and the prior-term Jacobians for the public API.
#include "bax_internal.cuh"

// util.triangulate_with_conf (reference util.py:325-354): per (frame, joint), rows
//   conf * [u P2 - P0 ; v P2 - P1] x = -conf * (4th column)   over cameras with score > 0,
// least squares; NaN unless sum(conf) > 0 and >= 2 views.  Solved by the normal equations in FP64 plus one
// round of iterative refinement on the true residual (corrected semi-normal equations), which restores
// QR-level accuracy for these 3-unknown systems.
__global__ void __launch_bounds__(128)
k_triangulate(const double* __restrict__ K, const double* __restrict__ R, const double* __restrict__ t,
              const double2* __restrict__ y, const double* __restrict__ s, int64_t P, int C,
              double* __restrict__ X, int* __restrict__ status) {
  extern __shared__ double Pm[];                 // [C][12] projection matrices K [R|t]
  for (int i = threadIdx.x; i < 12 * C; i += blockDim.x) {
    int c = i / 12, r = (i % 12) / 4, k = i % 4;
    const double* Kc = K + 9 * c;
    double v = 0;
    for (int j = 0; j < 3; ++j) v += Kc[3 * r + j] * (k < 3 ? R[9 * c + 3 * j + k] : t[3 * c + j]);
    Pm[i] = v;
  }
  __syncthreads();
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  double N0 = 0, N1 = 0, N2 = 0, N3 = 0, N4 = 0, N5 = 0, v0 = 0, v1 = 0, v2 = 0, csum = 0;
  int nv = 0;
  for (int c = 0; c < C; ++c) {
    double sc = s[(size_t)c * P + p];
    csum += sc;                                  // NaN scores poison the sum exactly as confi.sum() does
    if (!(sc > 0)) continue;
    ++nv;
    double2 yy = y[(size_t)c * P + p];
    const double* Q = Pm + 12 * c;
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      double uv = a == 0 ? yy.x : yy.y;
      const double* Pa = Q + 4 * a;
      double a0 = sc * (uv * Q[8] - Pa[0]), a1 = sc * (uv * Q[9] - Pa[1]), a2 = sc * (uv * Q[10] - Pa[2]);
      double b = -sc * (uv * Q[11] - Pa[3]);
      N0 += a0 * a0; N1 += a0 * a1; N2 += a0 * a2; N3 += a1 * a1; N4 += a1 * a2; N5 += a2 * a2;
      v0 += a0 * b; v1 += a1 * b; v2 += a2 * b;
    }
  }
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  if (!(csum > 0) || nv < 2) { X[3 * p] = nan; X[3 * p + 1] = nan; X[3 * p + 2] = nan; return; }
  double V[6] = {N0, N1, N2, N3, N4, N5};
  Chol3 L = chol3(V, 0.0);
  double tr = N0 + N3 + N5;
  double tol = 2.220446049250313e-16 * (2.0 * nv);
  if (!(L.l00 * L.l00 > tol * tol * tr) || !(L.l11 * L.l11 > tol * tol * tr) || !(L.l22 * L.l22 > tol * tol * tr)) {
    atomicExch(status, 1);                       // rank < 3: the reference raises "not enough constraint"
    X[3 * p] = nan; X[3 * p + 1] = nan; X[3 * p + 2] = nan; return;
  }
  double x0 = v0, x1 = v1, x2 = v2;
  chol3_fwd(L, x0, x1, x2); chol3_bwd(L, x0, x1, x2);
  // refinement: d = (A^T A)^-1 A^T (b - A x)
  double r0 = 0, r1 = 0, r2 = 0;
  for (int c = 0; c < C; ++c) {
    double sc = s[(size_t)c * P + p];
    if (!(sc > 0)) continue;
    double2 yy = y[(size_t)c * P + p];
    const double* Q = Pm + 12 * c;
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      double uv = a == 0 ? yy.x : yy.y;
      const double* Pa = Q + 4 * a;
      double a0 = sc * (uv * Q[8] - Pa[0]), a1 = sc * (uv * Q[9] - Pa[1]), a2 = sc * (uv * Q[10] - Pa[2]);
      double b = -sc * (uv * Q[11] - Pa[3]);
      double res = b - (a0 * x0 + a1 * x1 + a2 * x2);
      r0 += a0 * res; r1 += a1 * res; r2 += a2 * res;
    }
  }
  chol3_fwd(L, r0, r1, r2); chol3_bwd(L, r0, r1, r2);
  X[3 * p] = x0 + r0; X[3 * p + 1] = x1 + r1; X[3 * p + 2] = x2 + r2;
}

extern "C" int bax_triangulate(bax_handle* h, const double* R, const double* t, double* X, int32_t* status_host,
                               void* stream) {
  if (!h || !h->campack || !R || !t || !X) { bax_set_error("bax_triangulate: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  int* st = (int*)h->buf[BAX_BUF_STEP];
  BAX_CUDA(cudaMemsetAsync(st, 0, sizeof(int), s));
  int nb = (int)((h->P + 127) / 128);
  k_triangulate<<<nb, 128, 12 * h->C * sizeof(double), s>>>(h->K, R, t, (const double2*)h->y2d, h->s2d, h->P, h->C, X, st);
  BAX_LAUNCH_CHECK();
  int hs = 0;
  BAX_CUDA(cudaMemcpyAsync(&hs, st, sizeof(int), cudaMemcpyDeviceToHost, s));
  BAX_CUDA(bax_stream_wait(h, s));
  if (status_host) *status_host = hs;
  return BAX_OK;
}

// ---- reprojection error of every observation under given cameras and points (eval.calc_repro_error, eval.py:76-108;
// also the distance test of ba.save_mask, ba.py:248-256): e[c][p] = || pi(K_c (R_c X_p + t_c)) - y_cp ||_2 in pixels,
// NaN where the observation is invisible (score <= 0 or NaN) or the point is NaN.  One thread per (camera, point).
__global__ void __launch_bounds__(256)
k_reproject_error(const double* __restrict__ K, const double* __restrict__ R, const double* __restrict__ t,
                  const double2* __restrict__ y, const double* __restrict__ sc, const double* __restrict__ X,
                  int64_t P, double* __restrict__ e_out) {
  __shared__ double cam[21];
  const int c = blockIdx.y;
  if (threadIdx.x < 9) { cam[threadIdx.x] = K[9 * c + threadIdx.x]; cam[9 + threadIdx.x] = R[9 * c + threadIdx.x]; }
  if (threadIdx.x < 3) cam[18 + threadIdx.x] = t[3 * c + threadIdx.x];
  __syncthreads();
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const double s = sc[(size_t)c * P + p];
  double e = nan("");
  if (s > 0) {
    const double2 yy = y[(size_t)c * P + p];
    const double x0 = X[3 * p], x1 = X[3 * p + 1], x2 = X[3 * p + 2];
    const double* Km = cam; const double* Rm = cam + 9; const double* tm = cam + 18;
    const double P0 = Rm[0] * x0 + Rm[1] * x1 + Rm[2] * x2 + tm[0];
    const double P1 = Rm[3] * x0 + Rm[4] * x1 + Rm[5] * x2 + tm[1];
    const double P2 = Rm[6] * x0 + Rm[7] * x1 + Rm[8] * x2 + tm[2];
    const double q0 = Km[0] * P0 + Km[1] * P1 + Km[2] * P2;
    const double q1 = Km[3] * P0 + Km[4] * P1 + Km[5] * P2;
    const double q2 = Km[6] * P0 + Km[7] * P1 + Km[8] * P2;
    const double d0 = q0 / q2 - yy.x, d1 = q1 / q2 - yy.y;
    e = sqrt(d0 * d0 + d1 * d1);
  }
  e_out[(size_t)c * P + p] = e;
}

extern "C" int bax_reproject_error(bax_handle* h, const double* R, const double* t, const double* X, double* e_out,
                                   void* stream) {
  if (!h || !h->K || !R || !t || !X || !e_out) { bax_set_error("bax_reproject_error: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  k_reproject_error<<<dim3((unsigned)((h->P + 255) / 256), h->C), 256, 0, s>>>(h->K, R, t, (const double2*)h->y2d, h->s2d, X, h->P, e_out);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}

// ---- prior Jacobians for the public (SciPy-callback style) API
__global__ void k_lbar(const double* __restrict__ bone_buf, int B, double* __restrict__ Lbar) {
  int b = threadIdx.x;
  if (b < B) Lbar[b] = bone_buf[b] / bone_buf[B];
}
__global__ void k_jvb(const double* __restrict__ x, int N, int J, int B, const int32_t* __restrict__ bones,
                      const double* __restrict__ Lbar, double scale, double* __restrict__ Jvb) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)N * B) return;
  int n = (int)(i / B), b = (int)(i % B);
  const double* a = x + ((size_t)n * J + bones[2 * b]) * 3;
  const double* q = x + ((size_t)n * J + bones[2 * b + 1]) * 3;
  double d0 = a[0] - q[0], d1 = a[1] - q[1], d2 = a[2] - q[2];
  double L = sqrt(d0 * d0 + d1 * d1 + d2 * d2);
  double f = scale * (L - Lbar[b]) / L;
  Jvb[3 * i] = f * d0; Jvb[3 * i + 1] = f * d1; Jvb[3 * i + 2] = f * d2;
}

int bax_var3d_jacobian_rowmajor(bax_handle* h, double* Jv3d, cudaStream_t s);   // bax_lm.cu

extern "C" int bax_prior_jacobian(bax_handle* h, const double* theta, double* Jv3d, double* Jvb, void* stream) {
  if (!h || !h->campack || !theta) { bax_set_error("bax_prior_jacobian: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  int rc = bax_launch_cam_prep(h, theta, s); if (rc) return rc;
  if (Jv3d && h->Nv > 0) { rc = bax_var3d_jacobian_rowmajor(h, Jv3d, s); if (rc) return rc; }
  if (Jvb && h->B > 0) {
    rc = bax_bone_pass1(h, theta, s); if (rc) return rc;
    k_lbar<<<1, 32, 0, s>>>(h->buf[BAX_BUF_BONE], h->B, h->Lbar);
    int64_t nb = ((int64_t)h->N * h->B + 255) / 256;
    double scale = h->lam2 * 2.0 / (double)h->n_frames_global;
    k_jvb<<<(unsigned)nb, 256, 0, s>>>(theta + 6 * h->C, h->N, h->J, h->B, h->bones, h->Lbar, scale, Jvb);
    BAX_LAUNCH_CHECK();
  }
  return BAX_OK;
}
