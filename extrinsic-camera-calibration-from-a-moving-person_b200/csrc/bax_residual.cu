// Residual vector (ba.objfun) and the fused residual + Jacobian kernel (one thread per observation).
#include "bax_internal.cuh"

// ------------------------------------------------------------------ compaction helper
// Rank of this thread's visible observation inside its 256-thread block (reference row order is
// camera-major, visible only: ba.py:44-61), by warp ballots + a 8-entry shared scan.
__device__ __forceinline__ int block_rank(bool vis, int* s_warp /*[8]*/) {
  unsigned m = __ballot_sync(0xffffffffu, vis);
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int r = __popc(m & ((1u << lane) - 1u));
  if (lane == 0) s_warp[warp] = __popc(m);
  __syncthreads();
  int base = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) base += (w < warp) ? s_warp[w] : 0;
  return base + r;
}

// ------------------------------------------------------------------ reprojection rows
__global__ void __launch_bounds__(256)
k_residual_reproj(const double* __restrict__ campack, const double* __restrict__ x,
                  const double2* __restrict__ y, const double* __restrict__ s, int64_t P,
                  const int64_t* __restrict__ blockbase, double* __restrict__ r_out) {
  __shared__ double cam[CAMPACK];
  __shared__ int s_warp[8];
  const int c = blockIdx.y;
  if (threadIdx.x < CAMPACK) cam[threadIdx.x] = campack[(size_t)c * CAMPACK + threadIdx.x];
  const int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x;
  double sc = 0.0; double2 yy = make_double2(0, 0);
  if (p < P) { sc = s[(size_t)c * P + p]; yy = y[(size_t)c * P + p]; }
  const bool vis = sc > 0;                       // NaN score -> not visible, as (ss2d > 0) in ba.py:124
  int rank = block_rank(vis, s_warp);            // contains the __syncthreads that publishes cam[]
  if (!vis) return;
  double e0, e1;
  obs_residual(cam, x[3 * p], x[3 * p + 1], x[3 * p + 2], yy.x, yy.y, sc, e0, e1);
  int64_t row = blockbase[(size_t)c * gridDim.x + blockIdx.x] + rank;
  reinterpret_cast<double2*>(r_out)[row] = make_double2(e0, e1);
}

// Fused residual + Jacobian blocks, materialised (the HBM-bound kernel of SURVEY §8d):
// algorithmic bytes per visible observation: read y 16 + s 8 + x 24/Cbar, write r 16 + Jcam 96 + Jpt 48.
// Output staging: each warp transposes its blocks through shared memory so every global store
// instruction writes 32 consecutive 16-byte words (full 128-byte lines).
__global__ void __launch_bounds__(256)
k_rj(const double* __restrict__ campack, const double* __restrict__ x, const double2* __restrict__ y,
     const double* __restrict__ s, int64_t P, const int64_t* __restrict__ blockbase,
     double2* __restrict__ r2, double2* __restrict__ Jcam, double2* __restrict__ Jpt,
     int32_t* __restrict__ obs_cam, int32_t* __restrict__ obs_pt) {
  __shared__ double cam[CAMPACK];
  __shared__ int s_warp[8];
  __shared__ double2 stage[8][32 * 9 + 1];       // per warp: 32 observations x 9 double2 (6 Jcam + 3 Jpt)
  const int c = blockIdx.y;
  if (threadIdx.x < CAMPACK) cam[threadIdx.x] = campack[(size_t)c * CAMPACK + threadIdx.x];
  const int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x;
  double sc = 0.0; double2 yy = make_double2(0, 0);
  if (p < P) { sc = s[(size_t)c * P + p]; yy = y[(size_t)c * P + p]; }
  const bool vis = sc > 0;
  const unsigned bal = __ballot_sync(0xffffffffu, vis);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wr = __popc(bal & ((1u << lane) - 1u));
  if (lane == 0) s_warp[warp] = __popc(bal);
  __syncthreads();
  int wbase = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) wbase += (w < warp) ? s_warp[w] : 0;
  const int nw = s_warp[warp];
  const int64_t row0 = blockbase[(size_t)c * gridDim.x + blockIdx.x] + wbase;   // first row of this warp
  if (vis) {
    ObsLin o;
    obs_linearize<true, true>(cam, x[3 * p], x[3 * p + 1], x[3 * p + 2], yy.x, yy.y, sc, o);
    r2[row0 + wr] = make_double2(o.e0, o.e1);
    if (obs_cam) { obs_cam[row0 + wr] = c; obs_pt[row0 + wr] = (int32_t)p; }
    double2* st = &stage[warp][wr * 9];
    st[0] = make_double2(o.jr[0][0], o.jr[0][1]);
    st[1] = make_double2(o.jr[0][2], o.jt[0][0]);
    st[2] = make_double2(o.jt[0][1], o.jt[0][2]);
    st[3] = make_double2(o.jr[1][0], o.jr[1][1]);
    st[4] = make_double2(o.jr[1][2], o.jt[1][0]);
    st[5] = make_double2(o.jt[1][1], o.jt[1][2]);
    st[6] = make_double2(o.jp[0][0], o.jp[0][1]);
    st[7] = make_double2(o.jp[0][2], o.jp[1][0]);
    st[8] = make_double2(o.jp[1][1], o.jp[1][2]);
  }
  __syncwarp();
  // coalesced write-out: Jcam rows row0..row0+nw-1 are 6*nw consecutive double2, Jpt 3*nw
  double2* jc = Jcam + row0 * 6;
  for (int i = lane; i < 6 * nw; i += 32) jc[i] = stage[warp][(i / 6) * 9 + (i % 6)];
  double2* jq = Jpt + row0 * 3;
  for (int i = lane; i < 3 * nw; i += 32) jq[i] = stage[warp][(i / 3) * 9 + 6 + (i % 3)];
}

// Residual and Jacobian blocks of one observation, parked in shared memory in their final global layout as soon as
// each exists (same arithmetic as obs_linearize<true>; keeps the live set near 18 doubles).
__device__ __forceinline__ void rj_park(const double* cam, const double x0, const double x1, const double x2,
                                        const double2 yy, const double sc, double2* stRk, double2* b, double2* aout) {
  const double* K = cam; const double* R = cam + 9; const double* t = cam + 18; const double* M = cam + 21;
  double jt[2][3];
  {
    const double P0 = R[0] * x0 + R[1] * x1 + R[2] * x2 + t[0];
    const double P1 = R[3] * x0 + R[4] * x1 + R[5] * x2 + t[1];
    const double P2 = R[6] * x0 + R[7] * x1 + R[8] * x2 + t[2];
    const double q0 = K[0] * P0 + K[1] * P1 + K[2] * P2;
    const double q1 = K[3] * P0 + K[4] * P1 + K[5] * P2;
    const double q2 = K[6] * P0 + K[7] * P1 + K[8] * P2;
    const double w = 1.4142135623730951 * sc;
    const double yh0 = q0 / q2, yh1 = q1 / q2;
    *stRk = make_double2((yy.x - yh0) * w, (yy.y - yh1) * w);
    const double f = -w / q2;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      jt[0][k] = f * (K[k] - yh0 * K[6 + k]);
      jt[1][k] = f * (K[3 + k] - yh1 * K[6 + k]);
    }
  }
  double jp[2][3];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int k = 0; k < 3; ++k) jp[a][k] = jt[a][0] * R[k] + jt[a][1] * R[3 + k] + jt[a][2] * R[6 + k];
  b[0] = make_double2(jp[0][0], jp[0][1]);
  b[1] = make_double2(jp[0][2], jp[1][0]);
  b[2] = make_double2(jp[1][1], jp[1][2]);
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    const double c0 = x1 * jp[a][2] - x2 * jp[a][1];
    const double c1 = x2 * jp[a][0] - x0 * jp[a][2];
    const double c2 = x0 * jp[a][1] - x1 * jp[a][0];
    const double r0 = c0 * M[0] + c1 * M[3] + c2 * M[6];
    const double r1 = c0 * M[1] + c1 * M[4] + c2 * M[7];
    const double r2v = c0 * M[2] + c1 * M[5] + c2 * M[8];
    aout[3 * a + 0] = make_double2(r0, r1);
    aout[3 * a + 1] = make_double2(r2v, jt[a][0]);
    aout[3 * a + 2] = make_double2(jt[a][1], jt[a][2]);
  }
}

// Bulk-store variant: the whole CTA stages its compacted rows in shared memory in their final global layout
// (the rows of one CTA are contiguous in r2 / Jcam / Jpt), then three elected threads hand the three runs to the
// TMA engine (cp.async.bulk shared::cta -> global).  The LSU issues no global stores for the Jacobian blocks, the
// register budget drops (values go to shared memory as soon as they exist) and NT*MINB threads stay resident.
__device__ __forceinline__ void bulk_store_s2g(void* gdst, const void* ssrc, unsigned bytes) {
  unsigned sa = (unsigned)__cvta_generic_to_shared(ssrc);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gdst), "r"(sa), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
}

template <int NT, int MINB>
__global__ void __launch_bounds__(NT, MINB)
k_rj_bulk(const double* __restrict__ campack, const double* __restrict__ x, const double2* __restrict__ y,
          const double* __restrict__ s, int64_t P, const int64_t* __restrict__ blockbase, int sub /*CTAs per 256-block*/,
          double2* __restrict__ r2, double2* __restrict__ Jcam, double2* __restrict__ Jpt,
          int32_t* __restrict__ obs_cam, int32_t* __restrict__ obs_pt) {
  extern __shared__ __align__(128) unsigned char smraw[];
  double2* stJc = reinterpret_cast<double2*>(smraw);          // [NT][6]
  double2* stJp = stJc + NT * 6;                              // [NT][3]
  double2* stR = stJp + NT * 3;                               // [NT]
  __shared__ double cam[CAMPACK];
  __shared__ int s_warp[NT / 32];
  const int c = blockIdx.y;
  if (threadIdx.x < CAMPACK) cam[threadIdx.x] = campack[(size_t)c * CAMPACK + threadIdx.x];
  // blockbase is per 256 observations; a CTA of NT < 256 threads covers a 1/sub slice of one such block
  const int blk256 = blockIdx.x / sub, part = blockIdx.x - blk256 * sub;
  const int64_t p0 = (int64_t)blk256 * 256;
  const int64_t p = p0 + part * NT + threadIdx.x;
  double sc = 0.0; double2 yy = make_double2(0, 0);
  if (p < P) { sc = s[(size_t)c * P + p]; yy = y[(size_t)c * P + p]; }
  const bool vis = sc > 0;
  const unsigned bal = __ballot_sync(0xffffffffu, vis);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wr = __popc(bal & ((1u << lane) - 1u));
  if (lane == 0) s_warp[warp] = __popc(bal);
  // visible observations of the earlier slices of this 256-block (sub > 1 only)
  int before = 0;
  if (sub > 1) {
    for (int q = 0; q < part; ++q) {
      int cnt = 0;
      for (int i = lane; i < NT; i += 32) { int64_t pp = p0 + q * NT + i; cnt += (pp < P && s[(size_t)c * P + pp] > 0) ? 1 : 0; }
      for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
      before += cnt;
    }
  }
  __syncthreads();
  int wbase = 0, ntot = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; ++w) { const int v = s_warp[w]; wbase += (w < warp) ? v : 0; ntot += v; }
  const int64_t row0 = blockbase[(size_t)c * (gridDim.x / sub) + blk256] + before;   // first row of this CTA
  if (vis) {
    const int rk = wbase + wr;
    if (obs_cam) { obs_cam[row0 + rk] = c; obs_pt[row0 + rk] = (int32_t)p; }
    rj_park(cam, x[3 * p], x[3 * p + 1], x[3 * p + 2], yy, sc, stR + rk, stJp + rk * 3, stJc + rk * 6);
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");   // generic-proxy writes -> visible to the async proxy
  }
  __syncthreads();
  if (ntot == 0) return;
  if (threadIdx.x == 0) bulk_store_s2g(Jcam + row0 * 6, stJc, (unsigned)ntot * 96u);
  else if (threadIdx.x == 32) bulk_store_s2g(Jpt + row0 * 3, stJp, (unsigned)ntot * 48u);
  else if (threadIdx.x == 64) bulk_store_s2g(r2 + row0, stR, (unsigned)ntot * 16u);
}


// Persistent, TMA-pipelined variant (BAX_RJ_VARIANT=4).  One CTA per SM slot walks the (camera, 256-observation)
// tiles with a fixed stride.  Every tile's inputs — 256 x (u,v) 4 KB, 256 x point 6 KB, 256 x score 2 KB and the
// camera pack 384 B, all contiguous in HBM — arrive by four cp.async.bulk copies that complete on an mbarrier, into
// a ring of STAGES shared-memory slots, so the loads of tile k+STAGES are in flight while tile k is computed and
// its compacted rows leave through the bulk-store path of k_rj_bulk.  No thread ever waits on a global load.
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load_g2s(void* sdst, const void* gsrc, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
               ::"r"(smem_u32(sdst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

constexpr int RJ_TILE = 256;
constexpr int RJ_IN_Y = 0, RJ_IN_X = RJ_TILE * 16, RJ_IN_S = RJ_IN_X + RJ_TILE * 24, RJ_IN_CAM = RJ_IN_S + RJ_TILE * 8;
constexpr int RJ_IN_BYTES = RJ_IN_CAM + CAMPACK * 8;      // 12 672 B per stage (a multiple of 128)
constexpr int RJ_OUT_BYTES = RJ_TILE * 160;
__host__ __device__ constexpr int rj_pipe_smem(int stages) { return RJ_OUT_BYTES + stages * RJ_IN_BYTES; }

template <int STAGES, int MINB>
__global__ void __launch_bounds__(RJ_TILE, MINB)
k_rj_pipe(const double* __restrict__ campack, const double* __restrict__ x, const double2* __restrict__ y,
          const double* __restrict__ s, int64_t P, const int64_t* __restrict__ blockbase, int nblk, int64_t ntiles,
          double2* __restrict__ r2, double2* __restrict__ Jcam, double2* __restrict__ Jpt,
          int32_t* __restrict__ obs_cam, int32_t* __restrict__ obs_pt) {
  extern __shared__ __align__(128) unsigned char smraw[];
  double2* stJc = reinterpret_cast<double2*>(smraw);          // [256][6]
  double2* stJp = stJc + RJ_TILE * 6;                         // [256][3]
  double2* stR = stJp + RJ_TILE * 3;                          // [256]
  unsigned char* ring = smraw + RJ_OUT_BYTES;
  __shared__ __align__(8) uint64_t full[STAGES];
  __shared__ int s_warp[RJ_TILE / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  auto issue_loads = [&](int64_t t, int stage) {              // thread 0 only
    const int c = (int)(t / nblk);
    const int64_t p0 = (t - (int64_t)c * nblk) * RJ_TILE;
    const int64_t rem = P - p0;
    const unsigned cnt = rem < RJ_TILE ? (unsigned)rem : (unsigned)RJ_TILE;   // even: P is even on this path
    unsigned char* st = ring + (size_t)stage * RJ_IN_BYTES;
    mbar_expect_tx(&full[stage], cnt * 48u + CAMPACK * 8u);
    bulk_load_g2s(st + RJ_IN_Y, y + (size_t)c * P + p0, cnt * 16u, &full[stage]);
    bulk_load_g2s(st + RJ_IN_X, x + 3 * p0, cnt * 24u, &full[stage]);
    bulk_load_g2s(st + RJ_IN_S, s + (size_t)c * P + p0, cnt * 8u, &full[stage]);
    bulk_load_g2s(st + RJ_IN_CAM, campack + (size_t)c * CAMPACK, CAMPACK * 8u, &full[stage]);
  };

  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < STAGES; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < STAGES; ++i) {
      const int64_t t = (int64_t)blockIdx.x + (int64_t)i * gridDim.x;
      if (t < ntiles) issue_loads(t, i);
    }
  }
  int stage = 0; unsigned parity = 0;
  int64_t row0 = blockIdx.x < ntiles ? blockbase[blockIdx.x] : 0;      // [C][nblk] is tile-indexed
  for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int64_t tnext = t + gridDim.x;
    const int64_t row0_next = tnext < ntiles ? blockbase[tnext] : 0;    // in flight under this tile's arithmetic
    const int c = (int)(t / nblk);
    const int blk = (int)(t - (int64_t)c * nblk);
    const int64_t p = (int64_t)blk * RJ_TILE + threadIdx.x;
    const unsigned char* st = ring + (size_t)stage * RJ_IN_BYTES;
    mbar_wait(&full[stage], parity);
    double sc = 0.0; double2 yy = make_double2(0, 0); double x0 = 0, x1 = 0, x2 = 0;
    if (p < P) {
      sc = reinterpret_cast<const double*>(st + RJ_IN_S)[threadIdx.x];
      yy = reinterpret_cast<const double2*>(st + RJ_IN_Y)[threadIdx.x];
      const double* xs = reinterpret_cast<const double*>(st + RJ_IN_X) + 3 * threadIdx.x;
      x0 = xs[0]; x1 = xs[1]; x2 = xs[2];
    }
    const double* cam = reinterpret_cast<const double*>(st + RJ_IN_CAM);
    const bool vis = sc > 0;
    const unsigned bal = __ballot_sync(0xffffffffu, vis);
    const int wr = __popc(bal & ((1u << lane) - 1u));
    if (lane == 0) s_warp[warp] = __popc(bal);
    // the staging area is free once the TMA engine has read the previous tile's rows out of it
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
    __syncthreads();
    int wbase = 0, ntot = 0;
#pragma unroll
    for (int w = 0; w < RJ_TILE / 32; ++w) { const int v = s_warp[w]; wbase += (w < warp) ? v : 0; ntot += v; }
    if (vis) {
      const int rk = wbase + wr;
      if (obs_cam) { obs_cam[row0 + rk] = c; obs_pt[row0 + rk] = (int32_t)p; }
      rj_park(cam, x0, x1, x2, yy, sc, stR + rk, stJp + rk * 3, stJc + rk * 6);
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    }
    __syncthreads();                       // rows parked; every thread is done with this input slot and s_warp
    if (threadIdx.x == 0) {
      if (ntot > 0) {
        const unsigned n = (unsigned)ntot;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(Jcam + row0 * 6), "r"(smem_u32(stJc)), "r"(n * 96u) : "memory");
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(Jpt + row0 * 3), "r"(smem_u32(stJp)), "r"(n * 48u) : "memory");
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(r2 + row0), "r"(smem_u32(stR)), "r"(n * 16u) : "memory");
        asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
      }
      const int64_t tn = t + (int64_t)STAGES * gridDim.x;
      if (tn < ntiles) issue_loads(tn, stage);
    }
    if (++stage == STAGES) { stage = 0; parity ^= 1u; }
    row0 = row0_next;
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");   // smem must outlive the last reads
}

// ------------------------------------------------------------------ orientation rows  (ba.py:67-96)
__global__ void k_var3d_residual(const double* __restrict__ campack, const double* __restrict__ dhat,
                                 int C, int64_t Nv, double lam1, double* __restrict__ r_out) {
  extern __shared__ double sR[];                 // [C][9]
  for (int i = threadIdx.x; i < 9 * C; i += blockDim.x) sR[i] = campack[(size_t)(i / 9) * CAMPACK + 9 + (i % 9)];
  __syncthreads();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Nv) return;
  double m0 = 0, m1 = 0, m2 = 0;
  for (int c = 0; c < C; ++c) {
    const double* d = dhat + ((size_t)c * Nv + i) * 3;
    const double* R = sR + 9 * c;
    double d0 = d[0], d1 = d[1], d2 = d[2];
    m0 += R[0] * d0 + R[3] * d1 + R[6] * d2;     // R^T d
    m1 += R[1] * d0 + R[4] * d1 + R[7] * d2;
    m2 += R[2] * d0 + R[5] * d1 + R[8] * d2;
  }
  m0 /= C; m1 /= C; m2 /= C;
  r_out[i] = lam1 * (1.0 - sqrt(m0 * m0 + m1 * m1 + m2 * m2));
}

// ------------------------------------------------------------------ bone-length rows  (ba.py:99-110)
// pass 1: per-bone sum of lengths over the local frames; pass 2: sum of squared deviations from
// the (global) mean.  Two passes keep the population variance accurate to FP64 round-off.
__global__ void __launch_bounds__(256)
k_bone_len(const double* __restrict__ x, int N, int J, int B, const int32_t* __restrict__ bones,
           const double* __restrict__ bone_buf /* null in pass 1: [B] sums, [B] = count */,
           double* __restrict__ part /*[gridDim.x][B]*/, double* __restrict__ out /*[B] (+ the frame count in pass 1)*/,
           unsigned* __restrict__ ticket) {
  __shared__ double sm[BAX_MAX_B * 32];
  double acc[BAX_MAX_B];
#pragma unroll
  for (int b = 0; b < BAX_MAX_B; ++b) acc[b] = 0.0;
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
    const double* xf = x + (size_t)n * J * 3;
#pragma unroll
    for (int b = 0; b < BAX_MAX_B; ++b) {
      if (b < B) {
        const double* a = xf + 3 * bones[2 * b];
        const double* q = xf + 3 * bones[2 * b + 1];
        double d0 = a[0] - q[0], d1 = a[1] - q[1], d2 = a[2] - q[2];
        double L = sqrt(d0 * d0 + d1 * d1 + d2 * d2);
        if (bone_buf) { double dl = L - bone_buf[b] / bone_buf[B]; acc[b] += dl * dl; }
        else acc[b] += L;
      }
    }
  }
  block_sum<BAX_MAX_B>(acc, sm);
  if (threadIdx.x == 0)
    for (int b = 0; b < B; ++b) part[(size_t)blockIdx.x * B + b] = acc[b];
  if (bax_last_block(ticket)) {                  // was: k_reduce_partials (+ k_set_scalar) after this kernel
    bax_fold_sum<256>(part, gridDim.x, B, out, sm);
    if (!bone_buf && threadIdx.x == 0) out[B] = (double)N;
  }
}
__global__ void k_bone_rows(const double* __restrict__ ssq, const double* __restrict__ bone_buf, int B,
                            double lam2, double* __restrict__ r_out) {
  int b = threadIdx.x;
  if (b < B) r_out[b] = lam2 * ssq[b] / bone_buf[B];
}

static int bone_grid(bax_handle* h) {
  int g = (h->N + 255) / 256;
  int cap = 4 * h->sm_count;
  return g < cap ? g : cap;
}
// fills buf[BONE] = { sum_n L_nb (b < B), local frame count }
int bax_bone_pass1(bax_handle* h, const double* theta, cudaStream_t s) {
  int g = bone_grid(h);
  const double* x = theta + 6 * h->C;
  k_bone_len<<<g, 256, 0, s>>>(x, h->N, h->J, h->B, h->bones, nullptr, h->partials, h->buf[BAX_BUF_BONE], h->tickets + 0);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}
// out[b] = sum_n (L_nb - Lbar_b)^2 over local frames (Lbar from the, possibly all-reduced, buf[BONE])
int bax_bone_pass2(bax_handle* h, const double* theta, double* out, cudaStream_t s) {
  if (h->B == 0) return BAX_OK;
  int g = bone_grid(h);
  const double* x = theta + 6 * h->C;
  double* part = h->partials + (size_t)(1 << 21);     // second half: pass-1 partials may still be live
  k_bone_len<<<g, 256, 0, s>>>(x, h->N, h->J, h->B, h->bones, h->buf[BAX_BUF_BONE], part, out, h->tickets + 1);
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}

// ------------------------------------------------------------------ API
extern "C" int bax_residual(bax_handle* h, const double* theta, double* r_out, void* stream) {
  if (!h || !h->campack || !theta || !r_out) { bax_set_error("bax_residual: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  int rc = bax_launch_cam_prep(h, theta, s); if (rc) return rc;
  const double* x = theta + 6 * h->C;
  k_residual_reproj<<<dim3(h->nblk_obs, h->C), 256, 0, s>>>(h->campack, x, (const double2*)h->y2d, h->s2d,
                                                            h->P, h->blockbase, r_out);
  BAX_LAUNCH_CHECK();
  if (h->Nv > 0) {
    k_var3d_residual<<<(unsigned)((h->Nv + 255) / 256), 256, 9 * h->C * sizeof(double), s>>>(
        h->campack, h->dhat, h->C, h->Nv, h->lam1, r_out + 2 * h->n_vis);
    BAX_LAUNCH_CHECK();
  }
  if (h->B > 0) {
    rc = bax_bone_pass1(h, theta, s); if (rc) return rc;
    rc = bax_bone_pass2(h, theta, h->buf[BAX_BUF_COST], s); if (rc) return rc;
    k_bone_rows<<<1, 32, 0, s>>>(h->buf[BAX_BUF_COST], h->buf[BAX_BUF_BONE], h->B, h->lam2,
                                 r_out + 2 * h->n_vis + h->Nv);
    BAX_LAUNCH_CHECK();
  }
  return BAX_OK;
}

extern "C" int bax_residual_jacobian(bax_handle* h, const double* theta, double* r2, double* Jcam, double* Jpt,
                                     int32_t* obs_cam, int32_t* obs_pt, void* stream) {
  if (!h || !h->campack || !theta || !r2 || !Jcam || !Jpt || ((obs_cam == nullptr) != (obs_pt == nullptr))) {
    bax_set_error("bax_residual_jacobian: bad argument / no problem"); return BAX_E_ARG;
  }
  cudaStream_t s = (cudaStream_t)stream;
  BAX_CUDA(cudaSetDevice(h->device));
  int rc = bax_launch_cam_prep(h, theta, s); if (rc) return rc;
  {
    BaxTimer t(h, BAX_T_RJ, s, 1);
    const double* x = theta + 6 * h->C;
#define RJ_ARGS h->campack, x, (const double2*)h->y2d, h->s2d, h->P, h->blockbase
#define RJ_OUT (double2*)r2, (double2*)Jcam, (double2*)Jpt, obs_cam, obs_pt
    // default (BAX_RJ_VARIANT unset): the persistent TMA pipeline, 2 CTAs/SM x 4 input stages when (nearly) every
    // observation is visible, 3 CTAs/SM x 2 stages when compaction leaves threads idle (more warps hide the arithmetic)
    int variant = h->rj_variant;
    if (variant < 0) variant = (double)h->n_vis >= 0.9 * (double)h->C * (double)h->P ? 5 : 4;
    switch (variant) {
      case 0:
        k_rj<<<dim3(h->nblk_obs, h->C), 256, 0, s>>>(RJ_ARGS, RJ_OUT);
        break;
      case 2: {
        const int smem = 128 * 160;
        k_rj_bulk<128, 8><<<dim3(h->nblk_obs * 2, h->C), 128, smem, s>>>(RJ_ARGS, 2, RJ_OUT);
        break;
      }
      case 4: case 5: case 6: case 7: case 8: {
        // persistent TMA pipeline; its bulk loads need 16-byte aligned score rows (P even), else variant 2
        if ((h->P & 1) == 0 && (((uintptr_t)x | (uintptr_t)h->y2d | (uintptr_t)h->s2d) & 15) == 0) {
          const int64_t ntiles = (int64_t)h->nblk_obs * h->C;
#define RJ_PIPE(ST, MB)                                                                                              \
  {                                                                                                                  \
    constexpr int smem = rj_pipe_smem(ST);                                                                           \
    static bool once = false;                                                                                        \
    if (!once) { BAX_CUDA(cudaFuncSetAttribute(k_rj_pipe<ST, MB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); once = true; } \
    int64_t g = (int64_t)h->sm_count * MB; if (g > ntiles) g = ntiles;                                               \
    k_rj_pipe<ST, MB><<<(unsigned)g, RJ_TILE, smem, s>>>(RJ_ARGS, h->nblk_obs, ntiles, RJ_OUT);                      \
  }
          switch (variant) {
            case 4: RJ_PIPE(2, 3) break;
            case 6: RJ_PIPE(3, 2) break;
            case 7: RJ_PIPE(5, 2) break;
            case 8: RJ_PIPE(8, 1) break;
            default: RJ_PIPE(4, 2) break;
          }
#undef RJ_PIPE
          break;
        }
        const int smem = 128 * 160;
        k_rj_bulk<128, 8><<<dim3(h->nblk_obs * 2, h->C), 128, smem, s>>>(RJ_ARGS, 2, RJ_OUT);
        break;
      }
      case 3: {
        const int smem = 256 * 160;
        static bool once = false;
        if (!once) { BAX_CUDA(cudaFuncSetAttribute(k_rj_bulk<256, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); once = true; }
        k_rj_bulk<256, 3><<<dim3(h->nblk_obs, h->C), 256, smem, s>>>(RJ_ARGS, 1, RJ_OUT);
        break;
      }
      default: {
        const int smem = 256 * 160;
        static bool once = false;
        if (!once) { BAX_CUDA(cudaFuncSetAttribute(k_rj_bulk<256, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); once = true; }
        k_rj_bulk<256, 4><<<dim3(h->nblk_obs, h->C), 256, smem, s>>>(RJ_ARGS, 1, RJ_OUT);
        break;
      }
    }
#undef RJ_ARGS
#undef RJ_OUT
  }
  BAX_LAUNCH_CHECK();
  return BAX_OK;
}
