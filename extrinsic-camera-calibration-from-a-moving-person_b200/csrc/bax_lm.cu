// Cost, linearisation and the damped Schur-complement trial step.
#include "bax_internal.cuh"


// ------------------------------------------------------------------ cost
// sum of squared reprojection residuals; grid (nb, C), grid-stride over points of one camera.
__global__ void __launch_bounds__(256)
k_cost_reproj(const double* __restrict__ campack, const double* __restrict__ x, const double2* __restrict__ y,
              const double* __restrict__ s, int64_t P, double* __restrict__ part) {
  __shared__ double cam[CAMPACK];
  __shared__ double sm[32];
  const int c = blockIdx.y;
  if (threadIdx.x < CAMPACK) cam[threadIdx.x] = campack[(size_t)c * CAMPACK + threadIdx.x];
  __syncthreads();
  double acc[1] = {0.0};
  for (int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x; p < P; p += (int64_t)gridDim.x * 256) {
    double sc = s[(size_t)c * P + p];
    if (sc > 0) {
      double2 yy = y[(size_t)c * P + p];
      double e0, e1;
      obs_residual(cam, x[3 * p], x[3 * p + 1], x[3 * p + 2], yy.x, yy.y, sc, e0, e1);
      acc[0] += e0 * e0 + e1 * e1;
    }
  }
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0) part[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = acc[0];
}

// orientation rows: residual^2 sum and (optionally) the transposed Jacobian rows for the SYRK.
//   r_i = lam1 (1 - |m|),  m = (1/C) sum_c R_c^T d_c
//   dr_i/drvec_c = -(lam1/C) ((R_c mhat) x d_c)^T Mn_c                      
// Output element (3c+k, i) at Gv[(3c+k)*sr + i*si]; LM use: Gv [(3C+1)][ldv] (sr=ldv, si=1) with the
// residual in the last row (rrow); API use: row-major [Nv][3C] (sr=1, si=3C).
template <bool JAC>
__global__ void __launch_bounds__(256)
k_var3d(const double* __restrict__ campack, const double* __restrict__ dhat, int C, int64_t Nv, double lam1,
        double* __restrict__ Gv, int64_t sr, int64_t si, double* __restrict__ rrow, double* __restrict__ part) {
  extern __shared__ double sRM[];                 // [C][18]: R, Mn
  __shared__ double sm[32];
  for (int i = threadIdx.x; i < 18 * C; i += blockDim.x) {
    int c = i / 18, k = i % 18;
    sRM[i] = campack[(size_t)c * CAMPACK + (k < 9 ? 9 + k : 30 + (k - 9))];
  }
  __syncthreads();
  double acc[1] = {0.0};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < Nv; i += (int64_t)gridDim.x * blockDim.x) {
    double m0 = 0, m1 = 0, m2 = 0;
    for (int c = 0; c < C; ++c) {
      const double* d = dhat + ((size_t)c * Nv + i) * 3;
      const double* R = sRM + 18 * c;
      double d0 = d[0], d1 = d[1], d2 = d[2];
      m0 += R[0] * d0 + R[3] * d1 + R[6] * d2;
      m1 += R[1] * d0 + R[4] * d1 + R[7] * d2;
      m2 += R[2] * d0 + R[5] * d1 + R[8] * d2;
    }
    m0 /= C; m1 /= C; m2 /= C;
    double nm = sqrt(m0 * m0 + m1 * m1 + m2 * m2);
    double r = lam1 * (1.0 - nm);
    acc[0] += r * r;
    if (JAC) {
      double h0 = m0 / nm, h1 = m1 / nm, h2 = m2 / nm, f = -lam1 / C;
      for (int c = 0; c < C; ++c) {
        const double* d = dhat + ((size_t)c * Nv + i) * 3;
        const double* R = sRM + 18 * c; const double* Mn = R + 9;
        double d0 = d[0], d1 = d[1], d2 = d[2];
        double a0 = R[0] * h0 + R[1] * h1 + R[2] * h2;      // R mhat
        double a1 = R[3] * h0 + R[4] * h1 + R[5] * h2;
        double a2 = R[6] * h0 + R[7] * h1 + R[8] * h2;
        double c0 = a1 * d2 - a2 * d1, c1 = a2 * d0 - a0 * d2, c2 = a0 * d1 - a1 * d0;
#pragma unroll
        for (int k = 0; k < 3; ++k)
          Gv[(size_t)(3 * c + k) * sr + (size_t)i * si] = f * (c0 * Mn[k] + c1 * Mn[3 + k] + c2 * Mn[6 + k]);
      }
      if (rrow) rrow[i] = r;
    }
  }
  block_sum<1>(acc, sm);
  if (threadIdx.x == 0) part[blockIdx.x] = acc[0];
}

static int obs_grid(bax_handle* h) {            // blocks per camera for grid-stride kernels
  int g = h->nblk_obs;
  int cap = (8 * h->sm_count + h->C - 1) / h->C;
  if (cap < 1) cap = 1;
  return g < cap ? g : cap;
}
static int row_grid(bax_handle* h, int64_t n) {
  int64_t g = (n + 255) / 256; int cap = 8 * h->sm_count;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}

extern "C" int bax_cost_begin(bax_handle* h, const double* theta, void* stream) {
  if (!h || !h->campack || !theta) { bax_set_error("bax_cost_begin: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  BaxTimer t(h, BAX_T_COST, s, 2);
  int rc = bax_launch_cam_prep(h, theta, s); if (rc) return rc;
  return bax_bone_pass1(h, theta, s);
}
extern "C" int bax_cost_mid(bax_handle* h, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  BaxTimer t(h, BAX_T_COST, s, 5);
  double* out = h->buf[BAX_BUF_COST];
  int rc = bax_bone_pass2(h, h->theta, out, s); if (rc) return rc;
  int g = obs_grid(h);
  double* part = h->partials;
  k_cost_reproj<<<dim3(g, h->C), 256, 0, s>>>(h->campack, h->theta + 6 * h->C, (const double2*)h->y2d, h->s2d, h->P, part);
  BAX_LAUNCH_CHECK();
  rc = bax_reduce_partials(h, part, g * h->C, 1, out + h->B, s); if (rc) return rc;
  if (h->Nv > 0) {
    int gv = row_grid(h, h->Nv);
    double* part2 = h->partials + (1 << 20);
    k_var3d<false><<<gv, 256, 18 * h->C * sizeof(double), s>>>(h->campack, h->dhat, h->C, h->Nv, h->lam1, nullptr, 0, 0, nullptr, part2);
    BAX_LAUNCH_CHECK();
    rc = bax_reduce_partials(h, part2, gv, 1, out + h->B + 1, s); if (rc) return rc;
  } else {
    BAX_CUDA(cudaMemsetAsync(out + h->B + 1, 0, sizeof(double), s));
  }
  return BAX_OK;
}
static double bone_cost_host(bax_handle* h, const double* ssq, double count) {
  double c = 0;
  for (int b = 0; b < h->B; ++b) { double r = h->lam2 * ssq[b] / count; c += r * r; }
  return c;
}
extern "C" int bax_cost_finish(bax_handle* h, double* cost_host, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  double* hp = h->host_pinned;
  BAX_CUDA(cudaMemcpyAsync(hp, h->buf[BAX_BUF_COST], sizeof(double) * (h->B + 2), cudaMemcpyDeviceToHost, s));
  BAX_CUDA(cudaMemcpyAsync(hp + 32, h->buf[BAX_BUF_BONE] + h->B, sizeof(double), cudaMemcpyDeviceToHost, s));
  BAX_CUDA(bax_stream_wait(h, s));
  *cost_host = 0.5 * (hp[h->B] + hp[h->B + 1] + bone_cost_host(h, hp, hp[32]));
  return BAX_OK;
}

// ------------------------------------------------------------------ linearisation
// per-camera 6x6 block sum_p Jc^T Jc (21 unique), gradient Jc^T e (6), sum e^2 (1).
template <int NT, int MINB>
__global__ void __launch_bounds__(NT, MINB)
k_lin_cam(const double* __restrict__ campack, const double* __restrict__ x, const double2* __restrict__ y,
          const double* __restrict__ s, int64_t P, int C, double* __restrict__ part /*[gridDim.x][C][28]*/) {
  __shared__ double cam[CAMPACK];
  __shared__ double sm[28 * 32];
  const int c = blockIdx.y;
  if (threadIdx.x < CAMPACK) cam[threadIdx.x] = campack[(size_t)c * CAMPACK + threadIdx.x];
  __syncthreads();
  double acc[28];
#pragma unroll
  for (int i = 0; i < 28; ++i) acc[i] = 0.0;
  for (int64_t p = (int64_t)blockIdx.x * NT + threadIdx.x; p < P; p += (int64_t)gridDim.x * NT) {
    double sc = s[(size_t)c * P + p];
    if (sc > 0) {
      double2 yy = y[(size_t)c * P + p];
      ObsLin o;
      obs_linearize<true>(cam, x[3 * p], x[3 * p + 1], x[3 * p + 2], yy.x, yy.y, sc, o);
      double j0[6] = {o.jr[0][0], o.jr[0][1], o.jr[0][2], o.jt[0][0], o.jt[0][1], o.jt[0][2]};
      double j1[6] = {o.jr[1][0], o.jr[1][1], o.jr[1][2], o.jt[1][0], o.jt[1][1], o.jt[1][2]};
      int k = 0;
#pragma unroll
      for (int a = 0; a < 6; ++a) {
#pragma unroll
        for (int b = a; b < 6; ++b) { acc[k] += j0[a] * j0[b] + j1[a] * j1[b]; ++k; }
      }
#pragma unroll
      for (int a = 0; a < 6; ++a) acc[21 + a] += j0[a] * o.e0 + j1[a] * o.e1;
      acc[27] += o.e0 * o.e0 + o.e1 * o.e1;
    }
  }
  block_sum<28>(acc, sm);
  if (threadIdx.x == 0) {
    double* o = part + ((size_t)blockIdx.x * C + c) * 28;
    for (int i = 0; i < 28; ++i) o[i] = acc[i];
  }
}

// The same 28 sums on the FP64 tensor pipe.  With A_c = [J_c | e] (two rows of 7 per observation) the camera block,
// the gradient and the cost are the 7 x 7 matrix A_c^T A_c — a SYRK with 8 (7 used) rows and K = 2 x visible
// observations.  Each lane linearises one observation and parks its 2 x 8 values in the warp's shared-memory slab
// X[col][k] (k = 2 lane + row; column stride 68 doubles: the 16-byte stores are contiguous across lanes, the operand
// loads X[lane>>2][4 kk + (lane&3)] hit every bank pair once); 16 mma.sync.m8n8k4.f64 then fold the 64 K-rows into the
// warp's 8 x 8 accumulator tile, both operands being the same element (C += X^T X).  Two accumulator registers per
// lane instead of 28: the kernel needs 70-odd registers where the DFMA version needs 166 (one CTA of 8 warps per SM),
// and the 28-value block reduction disappears.
#define LCM_WARPS 4
#define LCM_LD 68
__device__ __forceinline__ void lcm_dmma(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
__global__ void __launch_bounds__(32 * LCM_WARPS, 4)
k_lin_cam_mma(const double* __restrict__ campack, const double* __restrict__ x, const double2* __restrict__ y,
              const double* __restrict__ s, int64_t P, int C, double* __restrict__ part /*[gridDim.x][C][28]*/) {
  __shared__ double cam[CAMPACK];
  __shared__ __align__(16) double X[LCM_WARPS][8 * LCM_LD];
  const int c = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < CAMPACK) cam[threadIdx.x] = campack[(size_t)c * CAMPACK + threadIdx.x];
  __syncthreads();
  double acc[2] = {0.0, 0.0};
  double* Xw = X[warp];
  const int NT = 32 * LCM_WARPS;
  // every lane of a warp runs the same number of rounds (the DMMA is warp-wide); out-of-range lanes contribute zeros
  const int64_t stride = (int64_t)gridDim.x * NT;
  // inputs of the next round are loaded before this round's arithmetic (ncu: long_scoreboard on top without it)
  double sc_n = 0.0, xn0 = 0.0, xn1 = 0.0, xn2 = 0.0; double2 yy_n = make_double2(0.0, 0.0);
  {
    const int64_t p = (int64_t)blockIdx.x * NT + warp * 32 + lane;
    if (p < P) { sc_n = s[(size_t)c * P + p]; yy_n = y[(size_t)c * P + p]; xn0 = x[3 * p]; xn1 = x[3 * p + 1]; xn2 = x[3 * p + 2]; }
  }
  for (int64_t base = (int64_t)blockIdx.x * NT + warp * 32; base < P; base += stride) {
    const int64_t p = base + lane;
    const double sc = sc_n, px0 = xn0, px1 = xn1, px2 = xn2; const double2 yy = yy_n;
    {
      const int64_t pn = p + stride;
      sc_n = 0.0;
      if (pn < P) { sc_n = s[(size_t)c * P + pn]; yy_n = y[(size_t)c * P + pn]; xn0 = x[3 * pn]; xn1 = x[3 * pn + 1]; xn2 = x[3 * pn + 2]; }
    }
    double j0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, j1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (p < P) {
      if (sc > 0) {
        ObsLin o;
        obs_linearize<true>(cam, px0, px1, px2, yy.x, yy.y, sc, o);
#pragma unroll
        for (int k = 0; k < 3; ++k) { j0[k] = o.jr[0][k]; j0[3 + k] = o.jt[0][k]; j1[k] = o.jr[1][k]; j1[3 + k] = o.jt[1][k]; }
        j0[6] = o.e0; j1[6] = o.e1;
      }
    }
#pragma unroll
    for (int col = 0; col < 8; ++col)
      *reinterpret_cast<double2*>(Xw + col * LCM_LD + 2 * lane) = make_double2(j0[col], j1[col]);
    __syncwarp();
    const double* src = Xw + (lane >> 2) * LCM_LD + (lane & 3);
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      const double v = src[4 * kk];
      lcm_dmma(acc, v, v);
    }
    __syncwarp();
  }
  // acc = C[lane>>2][2 (lane&3) + {0,1}] of this warp; fixed-order sum over the warps, then the packed 28 values
  __syncthreads();
  double* red = &X[0][0];                       // [LCM_WARPS][64]
  red[warp * 64 + (lane >> 2) * 8 + 2 * (lane & 3)] = acc[0];
  red[warp * 64 + (lane >> 2) * 8 + 2 * (lane & 3) + 1] = acc[1];
  __syncthreads();
  if (threadIdx.x < 28) {
    int a, b;
    if (threadIdx.x < 21) {                     // k-th entry of the upper triangle of the 6 x 6 block, row-major
      int k = threadIdx.x; a = 0;
      while (k >= 6 - a) { k -= 6 - a; ++a; }
      b = a + k;
    } else if (threadIdx.x < 27) { a = threadIdx.x - 21; b = 6; }
    else { a = 6; b = 6; }
    double v = 0.0;
#pragma unroll
    for (int w = 0; w < LCM_WARPS; ++w) v += red[w * 64 + a * 8 + b];
    part[((size_t)blockIdx.x * C + c) * 28 + threadIdx.x] = v;
  }
}

// per-point 3x3 block V_p = sum_c Jp^T Jp and reprojection gradient g_p = sum_c Jp^T e.
__global__ void __launch_bounds__(128)
k_lin_pt(const double* __restrict__ campack, const double* __restrict__ x, const double2* __restrict__ y,
         const double* __restrict__ s, int64_t P, int C, double* __restrict__ Vp, double* __restrict__ gp) {
  extern __shared__ double cams[];               // [C][CAMPACK]
  for (int i = threadIdx.x; i < C * CAMPACK; i += blockDim.x) cams[i] = campack[i];
  __syncthreads();
  int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  double x0 = x[3 * p], x1 = x[3 * p + 1], x2 = x[3 * p + 2];
  double V[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
  double sc_q[BAX_PF]; double2 yy_q[BAX_PF];               // loads of the next BAX_PF cameras in flight (see k_build_yhat)
#pragma unroll
  for (int k = 0; k < BAX_PF; ++k) {
    const int cc = k < C ? k : C - 1;
    sc_q[k] = s[(size_t)cc * P + p]; yy_q[k] = y[(size_t)cc * P + p];
  }
  for (int c0 = 0; c0 < C; c0 += BAX_PF) {
#pragma unroll
   for (int k = 0; k < BAX_PF; ++k) {           // ring slot k is a compile-time index: the queue stays in registers
    const int c = c0 + k;
    if (c >= C) break;
    const double sc = sc_q[k]; const double2 yy = yy_q[k];
    {
      const int cn = c + BAX_PF < C ? c + BAX_PF : C - 1;
      sc_q[k] = s[(size_t)cn * P + p]; yy_q[k] = y[(size_t)cn * P + p];
    }
    if (sc > 0) {
      ObsLin o;
      obs_linearize<false>(cams + c * CAMPACK, x0, x1, x2, yy.x, yy.y, sc, o);
      V[0] += o.jp[0][0] * o.jp[0][0] + o.jp[1][0] * o.jp[1][0];
      V[1] += o.jp[0][0] * o.jp[0][1] + o.jp[1][0] * o.jp[1][1];
      V[2] += o.jp[0][0] * o.jp[0][2] + o.jp[1][0] * o.jp[1][2];
      V[3] += o.jp[0][1] * o.jp[0][1] + o.jp[1][1] * o.jp[1][1];
      V[4] += o.jp[0][1] * o.jp[0][2] + o.jp[1][1] * o.jp[1][2];
      V[5] += o.jp[0][2] * o.jp[0][2] + o.jp[1][2] * o.jp[1][2];
#pragma unroll
      for (int q = 0; q < 3; ++q) g[q] += o.jp[0][q] * o.e0 + o.jp[1][q] * o.e1;
    }
   }
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) Vp[6 * p + k] = V[k];
#pragma unroll
  for (int k = 0; k < 3; ++k) gp[3 * p + k] = g[k];
}

// k_lin_cam_mma and k_lin_pt in one pass: both linearise every observation (reading s, y and x again), one to sum over
// the points of a camera, the other over the cameras of a point.  Here a thread owns R points (as in k_lin_pt: V_p and
// g_p stay in its registers over the camera loop) and a CTA owns the 128 R points of a chunk; per camera each warp folds
// its 32 R observations into the 8 x 8 tile [J_c | e]^T [J_c | e] with DMMA exactly as k_lin_cam_mma does, the four warps'
// tiles meet in shared memory (double-buffered: one CTA barrier per camera) and the 28 packed values go to
// part[chunk][c][28].  The order of the additions inside V_p / g_p is that of k_lin_pt; the camera sums are cut into
// chunks of points instead of grid strides, so they differ from k_lin_cam_mma's in the last bits.
template <int R>
__global__ void __launch_bounds__(128, 4)
k_lin_fused(const double* __restrict__ campack, const double* __restrict__ x, const double2* __restrict__ y,
            const double* __restrict__ s, int64_t P, int C, double* __restrict__ Vp, double* __restrict__ gp,
            double* __restrict__ part /*[gridDim.x][C][28]*/, double* __restrict__ camsum /*[C][28]*/,
            unsigned* __restrict__ ticket /*null: the caller reduces `part` with a kernel of its own*/) {
  extern __shared__ __align__(16) double smf[];  // cams [C][CAMPACK] | X [4][8 LCM_LD] | red [2][4][64]
  double* cams = smf;
  double* Xall = smf + C * CAMPACK;
  double* red = Xall + 4 * 8 * LCM_LD;
  for (int i = threadIdx.x; i < C * CAMPACK; i += blockDim.x) cams[i] = campack[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double* Xw = Xall + warp * 8 * LCM_LD;
  int64_t pt[R]; bool in[R];
  double px[R][3], V[R][6], g[R][3];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    pt[r] = ((int64_t)blockIdx.x * R + r) * 128 + threadIdx.x;
    in[r] = pt[r] < P;
    const int64_t q = in[r] ? pt[r] : P - 1;
    px[r][0] = x[3 * q]; px[r][1] = x[3 * q + 1]; px[r][2] = x[3 * q + 2];
#pragma unroll
    for (int k = 0; k < 6; ++k) V[r][k] = 0.0;
    g[r][0] = g[r][1] = g[r][2] = 0.0;
    if (!in[r]) pt[r] = P - 1;                   // loads stay in range; the score is forced to zero below
  }
  // (score, 2-D point) of the next camera are in flight while this one is worked on
  double sc_n[R]; double2 yy_n[R];
#pragma unroll
  for (int r = 0; r < R; ++r) { sc_n[r] = s[pt[r]]; yy_n[r] = y[pt[r]]; }
  for (int c = 0; c < C; ++c) {
    double acc[2] = {0.0, 0.0};
    const int cn = c + 1 < C ? c + 1 : c;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const double sc = in[r] ? sc_n[r] : 0.0; const double2 yy = yy_n[r];
      sc_n[r] = s[(size_t)cn * P + pt[r]]; yy_n[r] = y[(size_t)cn * P + pt[r]];
      double j0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, j1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (sc > 0) {
        ObsLin o;
        obs_linearize<true>(cams + c * CAMPACK, px[r][0], px[r][1], px[r][2], yy.x, yy.y, sc, o);
        V[r][0] += o.jp[0][0] * o.jp[0][0] + o.jp[1][0] * o.jp[1][0];
        V[r][1] += o.jp[0][0] * o.jp[0][1] + o.jp[1][0] * o.jp[1][1];
        V[r][2] += o.jp[0][0] * o.jp[0][2] + o.jp[1][0] * o.jp[1][2];
        V[r][3] += o.jp[0][1] * o.jp[0][1] + o.jp[1][1] * o.jp[1][1];
        V[r][4] += o.jp[0][1] * o.jp[0][2] + o.jp[1][1] * o.jp[1][2];
        V[r][5] += o.jp[0][2] * o.jp[0][2] + o.jp[1][2] * o.jp[1][2];
#pragma unroll
        for (int q = 0; q < 3; ++q) g[r][q] += o.jp[0][q] * o.e0 + o.jp[1][q] * o.e1;
#pragma unroll
        for (int k = 0; k < 3; ++k) { j0[k] = o.jr[0][k]; j0[3 + k] = o.jt[0][k]; j1[k] = o.jr[1][k]; j1[3 + k] = o.jt[1][k]; }
        j0[6] = o.e0; j1[6] = o.e1;
      }
#pragma unroll
      for (int col = 0; col < 8; ++col)
        *reinterpret_cast<double2*>(Xw + col * LCM_LD + 2 * lane) = make_double2(j0[col], j1[col]);
      __syncwarp();
      const double* src = Xw + (lane >> 2) * LCM_LD + (lane & 3);
#pragma unroll
      for (int kk = 0; kk < 16; ++kk) {
        const double v = src[4 * kk];
        lcm_dmma(acc, v, v);
      }
      __syncwarp();
    }
    // acc = tile[lane>>2][2 (lane&3) + {0,1}] of this warp for camera c
    double* rd = red + (c & 1) * 256;
    *reinterpret_cast<double2*>(rd + warp * 64 + (lane >> 2) * 8 + 2 * (lane & 3)) = make_double2(acc[0], acc[1]);
    __syncthreads();                              // the other buffer is not written again before the next barrier
    if (threadIdx.x < 28) {
      int a, b;
      if (threadIdx.x < 21) {                     // k-th entry of the upper triangle of the 6 x 6 block, row-major
        int k = threadIdx.x; a = 0;
        while (k >= 6 - a) { k -= 6 - a; ++a; }
        b = a + k;
      } else if (threadIdx.x < 27) { a = threadIdx.x - 21; b = 6; }
      else { a = 6; b = 6; }
      double v = 0.0;
#pragma unroll
      for (int w = 0; w < 4; ++w) v += rd[w * 64 + a * 8 + b];
      part[((size_t)blockIdx.x * C + c) * 28 + threadIdx.x] = v;
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (!in[r]) continue;
#pragma unroll
    for (int k = 0; k < 6; ++k) Vp[6 * pt[r] + k] = V[r][k];
#pragma unroll
    for (int k = 0; k < 3; ++k) gp[3 * pt[r] + k] = g[r][k];
  }
  // small problems: the block that finishes last adds the chunks up, one value per thread, chunks in order
  if (ticket && bax_last_block(ticket)) {
    const int nv = C * 28;
    for (int v = threadIdx.x; v < nv; v += 128) {
      double acc = 0.0;
      for (unsigned b = 0; b < gridDim.x; ++b) acc += __ldcg(part + (size_t)b * nv + v);
      camsum[v] = acc;
    }
  }
}

// scatter the per-camera sums and the orientation-row SYRK into the packed LIN buffer
//   lin = [ Hcc (6C x 6C) | gc (6C) | ssq_b (B) | sum e^2 | sum r_v3d^2 ]
__global__ void k_lin_pack(const double* __restrict__ camsum /*[C][28]*/, const double* __restrict__ v3d /*[(3C+1)^2] or null*/,
                           int C, double* __restrict__ lin, int B) {
  const int C6 = 6 * C, n3 = 3 * C + 1;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < C6 * C6; idx += gridDim.x * blockDim.x) {
    int i = idx / C6, j = idx % C6;
    int ci = i / 6, a = i % 6, cj = j / 6, b = j % 6;
    double v = 0.0;
    if (ci == cj) {
      int lo = a < b ? a : b, hi = a < b ? b : a;
      int k = lo * 6 - lo * (lo - 1) / 2 + (hi - lo);
      v = camsum[ci * 28 + k];
    }
    if (v3d && a < 3 && b < 3) v += v3d[(size_t)(3 * ci + a) * n3 + (3 * cj + b)];
    lin[idx] = v;
  }
  if (blockIdx.x == 0) {
    double* gc = lin + (size_t)C6 * C6;
    for (int i = threadIdx.x; i < C6; i += blockDim.x) {
      int c = i / 6, a = i % 6;
      double v = camsum[c * 28 + 21 + a];
      if (v3d && a < 3) v += v3d[(size_t)(3 * c + a) * n3 + 3 * C];
      gc[i] = v;
    }
    if (threadIdx.x == 0) {
      double e = 0; for (int c = 0; c < C; ++c) e += camsum[c * 28 + 27];
      gc[C6 + B] = e;
      gc[C6 + B + 1] = v3d ? v3d[(size_t)(3 * C) * n3 + 3 * C] : 0.0;
    }
  }
}

// after the LIN reduction: Lbar_b, r_b = lam2 * ssq_b / Ntot
__global__ void k_bone_finalize(const double* __restrict__ bone_buf, const double* __restrict__ ssq, int B,
                                double lam2, double* __restrict__ Lbar, double* __restrict__ rb) {
  int b = threadIdx.x;
  if (b < B) { Lbar[b] = bone_buf[b] / bone_buf[B]; rb[b] = lam2 * ssq[b] / bone_buf[B]; }
}

// bone-length Jacobian entry of point p for incident bone (b, sgn):  u = sgn * lam2 (2/Ntot) (L - Lbar) dv / L
__device__ __forceinline__ void bone_u(const double* __restrict__ x, int64_t frame_base3, const int32_t* __restrict__ bones,
                                       int b, int sgn, const double* __restrict__ Lbar, double scale, double (&u)[3]) {
  const double* a = x + frame_base3 + 3 * bones[2 * b];
  const double* q = x + frame_base3 + 3 * bones[2 * b + 1];
  double d0 = a[0] - q[0], d1 = a[1] - q[1], d2 = a[2] - q[2];
  double L = sqrt(d0 * d0 + d1 * d1 + d2 * d2);
  double f = sgn * scale * (L - Lbar[b]) / L;
  u[0] = f * d0; u[1] = f * d1; u[2] = f * d2;
}

// g_p += sum_b u_pb r_b; partial sums of |g_p|^2, |x_p|^2 and max |g_p|.
__global__ void __launch_bounds__(256)
k_gp_bone(const double* __restrict__ x, int64_t P, int J, const int32_t* __restrict__ bones,
          const int32_t* __restrict__ joint_bones, const double* __restrict__ bone_buf /*[B] sums, [B] count*/,
          const double* __restrict__ ssq /*[B]*/, int B, double lam2, double* __restrict__ Lbar_out, double* __restrict__ rb_out,
          double scale, double* __restrict__ gp, const double* __restrict__ Vp,
          double* __restrict__ part /*[grid][2]*/, double* __restrict__ pmax /*[2][grid]*/,
          double* __restrict__ norm_out /*[2]*/, double* __restrict__ gmax_out /*[2]*/, unsigned* __restrict__ ticket,
          const double* __restrict__ stage_lin_tail /*null: no staging*/, const double* __restrict__ stage_theta, int C6,
          double* __restrict__ stage_out) {
  __shared__ double sm[2 * 32];
  __shared__ double smax[16];
  // Lbar_b and r_b = lam2 ssq_b / Ntot (was: k_bone_finalize before this kernel); block 0 publishes them
  __shared__ double Lbar[BAX_MAX_B], rb[BAX_MAX_B];
  if (threadIdx.x < BAX_MAX_B) {
    const int b = threadIdx.x;
    const double lb = b < B ? bone_buf[b] / bone_buf[B] : 0.0, r = b < B ? lam2 * ssq[b] / bone_buf[B] : 0.0;
    Lbar[b] = lb; rb[b] = r;
    if (blockIdx.x == 0 && b < B) { Lbar_out[b] = lb; rb_out[b] = r; }
  }
  __syncthreads();
  double acc[2] = {0, 0}, mx = 0, mv = 0;
  for (int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x; p < P; p += (int64_t)gridDim.x * 256) {
    int j = (int)(p % J); int64_t fb = (p - j) * 3;
    double g0 = gp[3 * p], g1 = gp[3 * p + 1], g2 = gp[3 * p + 2];
#pragma unroll
    for (int k = 0; k < BAX_MAX_JB; ++k) {
      int b = joint_bones[(j * BAX_MAX_JB + k) * 2];
      if (b >= 0) {
        double u[3];
        bone_u(x, fb, bones, b, joint_bones[(j * BAX_MAX_JB + k) * 2 + 1], Lbar, scale, u);
        double r = rb[b];
        g0 += u[0] * r; g1 += u[1] * r; g2 += u[2] * r;
      }
    }
    gp[3 * p] = g0; gp[3 * p + 1] = g1; gp[3 * p + 2] = g2;
    acc[0] += g0 * g0 + g1 * g1 + g2 * g2;
    acc[1] += x[3 * p] * x[3 * p] + x[3 * p + 1] * x[3 * p + 1] + x[3 * p + 2] * x[3 * p + 2];
    mx = fmax(mx, fmax(fabs(g0), fmax(fabs(g1), fabs(g2))));
    mv = fmax(mv, fmax(Vp[6 * p], fmax(Vp[6 * p + 3], Vp[6 * p + 5])));
  }
  for (int o = 16; o > 0; o >>= 1) {
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mv = fmax(mv, __shfl_xor_sync(0xffffffffu, mv, o));
  }
  if ((threadIdx.x & 31) == 0) { smax[threadIdx.x >> 5] = mx; smax[8 + (threadIdx.x >> 5)] = mv; }
  block_sum<2>(acc, sm);
  if (threadIdx.x == 0) {
    for (int w = 0; w < 8; ++w) { mx = fmax(mx, smax[w]); mv = fmax(mv, smax[8 + w]); }
    part[2 * blockIdx.x] = acc[0]; part[2 * blockIdx.x + 1] = acc[1];
    pmax[blockIdx.x] = mx; pmax[gridDim.x + blockIdx.x] = mv;
  }
  if (bax_last_block(ticket)) {                  // was: k_reduce_partials and two k_reduce_max after this kernel
    bax_fold_sum<256>(part, gridDim.x, 2, norm_out, sm);
    bax_fold_max<256>(pmax, gridDim.x, gmax_out, sm);
    bax_fold_max<256>(pmax + gridDim.x, gridDim.x, gmax_out + 1, sm);
    if (stage_lin_tail) {                        // was: k_lin_stage (runs only when nothing is all-reduced in between)
      const int n1 = C6 + B + 2;
      for (int i = threadIdx.x; i < n1; i += blockDim.x) stage_out[i] = stage_lin_tail[i];
      for (int i = threadIdx.x; i < C6; i += blockDim.x) stage_out[n1 + 6 + i] = stage_theta[i];
      if (threadIdx.x == 0) {
        stage_out[n1] = norm_out[0]; stage_out[n1 + 1] = norm_out[1]; stage_out[n1 + 2] = gmax_out[0];
        stage_out[n1 + 3] = bone_buf[B]; stage_out[n1 + 4] = gmax_out[1]; stage_out[n1 + 5] = 0.0;
      }
    }
  }
}
__global__ void k_reduce_max(const double* __restrict__ v, int n, double* __restrict__ out) {
  __shared__ double sm[32];
  double mx = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) mx = fmax(mx, v[i]);
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x == 0) { for (int w = 0; w < (int)(blockDim.x >> 5); ++w) mx = fmax(mx, sm[w]); *out = mx; }
}

static int ensure_trial_buffers(bax_handle* h) {
  if (!h->Yhat) {
    BAX_ALLOC(h, h->Yhat, sizeof(double) * (size_t)h->R * h->ldy);
    BAX_CUDA(cudaMemsetAsync(h->Yhat, 0, sizeof(double) * (size_t)h->R * h->ldy, h->mem_stream));
    BAX_CUDA(cudaStreamSynchronize(h->mem_stream));
  }
  if (!h->Gv && h->Nv > 0) {
    BAX_ALLOC(h, h->Gv, sizeof(double) * (size_t)(3 * h->C + 1) * h->ldv);
    BAX_CUDA(cudaMemsetAsync(h->Gv, 0, sizeof(double) * (size_t)(3 * h->C + 1) * h->ldv, h->mem_stream));
    BAX_CUDA(cudaStreamSynchronize(h->mem_stream));
  }
  return BAX_OK;
}

extern "C" int bax_linearize_begin(bax_handle* h, const double* theta, void* stream) {
  if (!h || !h->campack || !theta) { bax_set_error("bax_linearize_begin: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  int rc = ensure_trial_buffers(h); if (rc) return rc;
  BaxTimer t(h, BAX_T_LIN, s, 2);
  rc = bax_launch_cam_prep(h, theta, s, true); if (rc) return rc;
  return bax_bone_pass1(h, theta, s);
}

extern "C" int bax_linearize_mid(bax_handle* h, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  const int C = h->C, C6 = 6 * C;
  const double* x = h->theta_lin + C6;
  double* lin = h->buf[BAX_BUF_LIN];
  int rc;
  {
    BaxTimer t(h, BAX_T_LIN, s, 0);              // launches counted below
    // BAX_LINCAM: tuning knob — 5 (default) = fused camera + point kernel (0.944 -> 0.827 ms per linearisation at the
    // cfg4 shard, 0.351 -> 0.292 at cfg3, unchanged at 4 cameras); 4 = DMMA camera kernel + k_lin_pt;
    // 0..3 = DFMA camera kernel at 256x1, 128x3, 256x2, 128x4 CTAs/SM + k_lin_pt
    static const int lc_variant = getenv("BAX_LINCAM") ? atoi(getenv("BAX_LINCAM")) : 5;
    double* camsum = h->partials + (1 << 20);    // [C][28]
    bool lin_fold = false;
    const int nbp = (int)((h->P + 127) / 128);
    if (lc_variant >= 5) {
      // one round of 128 points per CTA: two rounds (half the partials, 18 more accumulator registers) measured
      // 0.842 against 0.827 ms per linearisation at 64 cameras, 0.312 against 0.292 at 31
      const int g = nbp;
      const size_t need = sizeof(double) * (size_t)g * C * 28;
      if (need > h->lin_part_bytes) {
        bax_dev_free(h, h->lin_part); h->lin_part = nullptr; h->lin_part_bytes = 0;
        BAX_ALLOC(h, h->lin_part, need);
        BAX_CUDA(cudaStreamSynchronize(h->mem_stream));
        h->lin_part_bytes = need;
      }
      const size_t smem = sizeof(double) * ((size_t)C * CAMPACK + 4 * 8 * LCM_LD + 2 * 256);
      BAX_CUDA(cudaFuncSetAttribute(k_lin_fused<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      const bool fold = (size_t)g * C * 28 <= 32768;             // a 4-camera problem: one launch fewer
      k_lin_fused<1><<<g, 128, smem, s>>>(h->campack_lin, x, (const double2*)h->y2d, h->s2d, h->P, C, h->Vp, h->gp, h->lin_part,
                                          camsum, fold ? h->tickets + 4 : nullptr);
      BAX_LAUNCH_CHECK();
      if (!fold) { rc = bax_reduce_partials(h, h->lin_part, g, C * 28, camsum, s); if (rc) return rc; }
      lin_fold = fold;
    } else {
      int g = obs_grid(h);
      double* part = h->partials;                  // [g][C][28]
      if ((size_t)g * C * 28 * sizeof(double) > h->partials_bytes / 4) { bax_set_error("partials overflow"); return BAX_E_STATE; }
      if (lc_variant >= 4) {                       // one wave of 4 resident CTAs per SM
        int gm = (4 * h->sm_count) / C; if (gm < 1) gm = 1;
        g = gm < nbp ? gm : nbp;
      }
      if (lc_variant == 0) k_lin_cam<256, 1><<<dim3(g, C), 256, 0, s>>>(h->campack_lin, x, (const double2*)h->y2d, h->s2d, h->P, C, part);
      else if (lc_variant == 1) k_lin_cam<128, 3><<<dim3(g, C), 128, 0, s>>>(h->campack_lin, x, (const double2*)h->y2d, h->s2d, h->P, C, part);
      else if (lc_variant == 2) k_lin_cam<256, 2><<<dim3(g, C), 256, 0, s>>>(h->campack_lin, x, (const double2*)h->y2d, h->s2d, h->P, C, part);
      else if (lc_variant == 3) k_lin_cam<128, 4><<<dim3(g, C), 128, 0, s>>>(h->campack_lin, x, (const double2*)h->y2d, h->s2d, h->P, C, part);
      else k_lin_cam_mma<<<dim3(g, C), 32 * LCM_WARPS, 0, s>>>(h->campack_lin, x, (const double2*)h->y2d, h->s2d, h->P, C, part);
      BAX_LAUNCH_CHECK();
      rc = bax_reduce_partials(h, part, g, C * 28, camsum, s); if (rc) return rc;
      k_lin_pt<<<nbp, 128, C * CAMPACK * sizeof(double), s>>>(h->campack_lin, x, (const double2*)h->y2d, h->s2d, h->P, C, h->Vp, h->gp);
      BAX_LAUNCH_CHECK();
    }
    rc = bax_bone_pass2(h, h->theta_lin, lin + (size_t)C6 * C6 + C6, s); if (rc) return rc;
    t.nl = (lc_variant >= 5 ? (lin_fold ? 1 : 2) : 3) + (h->B > 0 ? 1 : 0);
  }
  double* v3d = nullptr;
  if (h->Nv > 0) {
    int gv = row_grid(h, h->Nv);
    {
      BaxTimer t(h, BAX_T_LIN, s, 1);
      k_var3d<true><<<gv, 256, 18 * C * sizeof(double), s>>>(h->campack_lin, h->dhat, C, h->Nv, h->lam1, h->Gv, h->ldv, 1,
                                                               h->Gv + (size_t)(3 * C) * h->ldv, h->partials + (1 << 20) + 4096);
      BAX_LAUNCH_CHECK();
    }
    v3d = h->partials + (1 << 20) + 8192 + 65536;       // [(3C+1)^2]
    rc = bax_syrk_aug(h, h->Gv, 3 * C, 1, h->Nv, h->ldv, v3d, 3 * C + 1, s); if (rc) return rc;   // residual row rides along
  }
  k_lin_pack<<<64, 256, 0, s>>>(h->partials + (1 << 20), v3d, C, lin, h->B);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}

// after the LIN all-reduce: finish g_p with the bone-length rows, start the norm reductions
extern "C" int bax_linearize_post(bax_handle* h, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  const int C6 = 6 * h->C;
  BaxTimer t(h, BAX_T_LIN, s, 1);
  double* lin = h->buf[BAX_BUF_LIN];
  int g = row_grid(h, h->P);
  double* part = h->partials; double* pmax = h->partials + (1 << 16);
  // scale = lam2 * 2 / Ntot ; Ntot lives on device (buf[BONE][B]); single-GPU fast path reads host N,
  // multi-GPU callers set it through bax_set_global_frames before linearising.
  double scale = h->lam2 * 2.0 / (double)h->n_frames_global;
  // an unsharded problem (no communicator attached, all frames local: frame-sharded callers, whichever language issues
  // their collectives, announce the global frame count through bax_set_global_frames) has nothing all-reduced between
  // this kernel and bax_linearize_finish: the last block gathers the scalars for the read-back as well, and
  // h->lin_staged tells bax_linearize_finish to skip k_lin_stage
  h->lin_staged = h->comm == nullptr && h->n_frames_global == (int64_t)h->N;
  k_gp_bone<<<g, 256, 0, s>>>(h->theta_lin + C6, h->P, h->J, h->bones, h->joint_bones, h->buf[BAX_BUF_BONE], lin + (size_t)C6 * C6 + C6,
                              h->B, h->lam2, h->Lbar, h->rb, scale, h->gp, h->Vp, part, pmax,
                              h->buf[BAX_BUF_NORM], h->buf[BAX_BUF_GMAX], h->tickets + 2,
                              h->lin_staged ? lin + (size_t)C6 * C6 : nullptr, h->theta_lin, C6, h->lin_stage);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}

// scalars of a linearisation gathered into one contiguous block (layout of bax_linearize_finish's staging area)
__global__ void k_lin_stage(const double* __restrict__ lin_tail, const double* __restrict__ norm, const double* __restrict__ gmax,
                            const double* __restrict__ bone_n, const double* __restrict__ theta, int C6, int B,
                            double* __restrict__ out) {
  const int n1 = C6 + B + 2;
  for (int i = threadIdx.x; i < n1; i += blockDim.x) out[i] = lin_tail[i];
  for (int i = threadIdx.x; i < C6; i += blockDim.x) out[n1 + 6 + i] = theta[i];
  if (threadIdx.x == 0) {
    out[n1] = norm[0]; out[n1 + 1] = norm[1]; out[n1 + 2] = gmax[0]; out[n1 + 3] = *bone_n; out[n1 + 4] = gmax[1];
    out[n1 + 5] = 0.0;
  }
}

extern "C" int bax_linearize_finish(bax_handle* h, double* out_host, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  const int C6 = 6 * h->C, B = h->B;
  if (C6 + B + 2 + 8 + C6 > 4096) { bax_set_error("too many cameras for the scalar staging buffer"); return BAX_E_ARG; }
  // one gather kernel + one copy into pinned memory instead of six small pageable copies (each of those is a
  // separate driver round trip: ~40 us per linearisation, which a 4-camera problem feels)
  const int n_stage = C6 + B + 8 + C6;
  double* lin = h->buf[BAX_BUF_LIN];
  if (!h->lin_staged) {
    k_lin_stage<<<1, 256, 0, s>>>(lin + (size_t)C6 * C6, h->buf[BAX_BUF_NORM], h->buf[BAX_BUF_GMAX], h->buf[BAX_BUF_BONE] + B,
                                  h->theta_lin, C6, B, h->lin_stage);
    BAX_LAUNCH_CHECK();
    h->t_launch[BAX_T_LIN] += 1;
  }
  h->lin_staged = 0;
  double* stage = h->host_pinned + 64;
  BAX_CUDA(cudaMemcpyAsync(stage, h->lin_stage, sizeof(double) * n_stage, cudaMemcpyDeviceToHost, s));
  BAX_CUDA(bax_stream_wait(h, s));
  const double* gc = stage; const double* ssq = stage + C6;
  double ginf = stage[C6 + B + 4], g2 = stage[C6 + B + 2], x2 = stage[C6 + B + 3];
  for (int i = 0; i < C6; ++i) { ginf = fmax(ginf, fabs(gc[i])); g2 += gc[i] * gc[i]; x2 += stage[C6 + B + 8 + i] * stage[C6 + B + 8 + i]; }
  h->cost = 0.5 * (stage[C6 + B] + stage[C6 + B + 1] + bone_cost_host(h, ssq, stage[C6 + B + 5]));
  out_host[0] = h->cost; out_host[1] = ginf; out_host[2] = sqrt(g2); out_host[3] = sqrt(x2);
  out_host[4] = stage[C6 + B + 6];
  return BAX_OK;
}

// orientation-row Jacobian in row-major [Nv][3C] for bax_prior_jacobian
int bax_var3d_jacobian_rowmajor(bax_handle* h, double* Jv3d, cudaStream_t s) {
  int gv = row_grid(h, h->Nv);
  k_var3d<true><<<gv, 256, 18 * h->C * sizeof(double), s>>>(h->campack, h->dhat, h->C, h->Nv, h->lam1, Jv3d, 1,
                                                             3 * h->C, nullptr, h->partials + (1 << 20) + 4096);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}
