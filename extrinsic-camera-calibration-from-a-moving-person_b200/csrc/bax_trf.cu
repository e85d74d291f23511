// Trust-region driver on the native side of the C ABI: the scalar logic of SciPy's trf_no_bounds
// (scipy/optimize/_lsq/trf.py:415-587; solve_lsq_trust_region, update_tr_radius, check_termination in
// _lsq/common.py:57-168,222-248,705-717) — the solver reference ba.py:187-206 runs — around the linearise / damped
// Schur solve phases of this library.  A frame-sharded run attaches an NCCL communicator (bax_allreduce_attach,
// bax_comm.cu) and the packed buffers are all-reduced here, between the phases, on the compute stream; the host-language
// twin (calib_b200/ba.py TRFSolver, which this file mirrors statement by statement) remains for instrumented passes.
// What it buys: a whole outer iteration — or a whole solve — is ONE call across the ABI, so a sweep over many small
// problems (BASELINE configs[4]) runs its worker streams without the host interpreter between the kernels.
#include "bax_internal.cuh"

__global__ void k_axpy_out(int64_t n, const double* __restrict__ x, const double* __restrict__ p, double kappa,
                           double* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = x[i] + kappa * p[i];
}

namespace {

struct Lin { double cost, g_inf, g_2, x_norm, h_scale; };

int linearize(bax_handle* h, const double* theta, Lin& o, cudaStream_t s) {
  int rc;
  // frame-sharded runs: the packed buffers are summed over the ranks between the phases (no-ops without a communicator)
  if ((rc = bax_linearize_begin(h, theta, s))) return rc;
  if ((rc = bax_allreduce(h, BAX_BUF_BONE, false, s))) return rc;
  if ((rc = bax_linearize_mid(h, s))) return rc;
  if ((rc = bax_allreduce(h, BAX_BUF_LIN, false, s))) return rc;
  if ((rc = bax_linearize_post(h, s))) return rc;
  if ((rc = bax_allreduce_pair(h, BAX_BUF_NORM, BAX_BUF_GMAX, s))) return rc;
  double out[8];
  if ((rc = bax_linearize_finish(h, out, s))) return rc;
  o.cost = out[0]; o.g_inf = out[1]; o.g_2 = out[2]; o.x_norm = out[3]; o.h_scale = out[4];
  return BAX_OK;
}

// one damped solve: p -> st->step; out = { |p|^2, g.p, p^T (J^T J + alpha I)^-1 p }
int trial(bax_handle* h, bax_trf_state* st, double alpha, int need_phi, double out[4], cudaStream_t s) {
  int rc;
  st->n_trials += 1;
  if ((rc = bax_trial_begin(h, alpha, s))) return rc;
  if ((rc = bax_allreduce(h, BAX_BUF_SCHUR, false, s))) return rc;
  if ((rc = bax_trial_mid(h, st->step, need_phi, s))) return rc;
  if ((rc = bax_allreduce(h, BAX_BUF_RHS2, false, s))) return rc;
  if ((rc = bax_trial_finish(h, out, s))) return rc;
  if (!(isfinite(out[0]) && isfinite(out[1]))) {
    bax_set_error("non-finite trial step at alpha=%g: reduced camera system not SPD", alpha);
    return BAX_E_STATE;
  }
  return BAX_OK;
}

struct TrStep { double alpha; int n_it; double p_norm, gdotp; bool on_boundary; };

// More'-Sorensen iteration on ||p(alpha)|| = Delta (common.py:57-168) with alpha floored at 1e-13 x the Hessian scale
// (the exact Jacobian has a gauge null space: DESIGN.md §5); same statements as calib_b200/ba.py::_solve_tr.
int solve_tr(bax_handle* h, bax_trf_state* st, double Delta, double g_norm2, double initial_alpha, TrStep& r,
             cudaStream_t s, double rtol = 0.01, int max_iter = 10) {
  const double alpha_floor = 1e-13 * st->h_scale;
  double alpha_upper = g_norm2 / Delta, alpha_lower = 0.0, alpha;
  double o[4]; int rc;
  auto reset = [&]() { return fmax(fmax(0.001 * alpha_upper, pow(alpha_lower * alpha_upper, 0.5)), alpha_floor); };
  if (initial_alpha == 0.0) {
    if ((rc = trial(h, st, alpha_floor, 0, o, s))) return rc;
    if (sqrt(o[0]) <= Delta) { r = {alpha_floor, 0, sqrt(o[0]), o[1], false}; return BAX_OK; }
    alpha = reset();
  } else {
    alpha = fmax(initial_alpha, alpha_floor);
  }
  int it = 0;
  for (it = 0; it < max_iter; ++it) {
    if (alpha < alpha_lower || alpha > alpha_upper) alpha = reset();
    alpha = fmax(alpha, alpha_floor);
    // phi' costs another pass over the whitened blocks and a second reduced solve: at the floor the step is usually
    // the interior one and returns before phi' is looked at, so it is computed only once that is ruled out
    const bool lazy = alpha <= alpha_floor;
    if ((rc = trial(h, st, alpha, lazy ? 0 : 1, o, s))) return rc;
    const double p_norm = sqrt(o[0]), phi = p_norm - Delta;
    if (phi < 0 && alpha <= alpha_floor) { r = {alpha, it + 1, p_norm, o[1], false}; return BAX_OK; }
    if (lazy) {
      if ((rc = bax_trial_phi_prime(h, s))) return rc;
      if ((rc = bax_allreduce(h, BAX_BUF_RHS2_HEAD, false, s))) return rc;
      if ((rc = bax_trial_finish(h, o, s))) return rc;
    }
    if (!(o[2] > 0 && p_norm > 0)) break;
    const double phi_prime = -o[2] / p_norm;
    if (phi < 0) alpha_upper = alpha;
    const double ratio = phi / phi_prime;
    alpha_lower = fmax(alpha_lower, alpha - ratio);
    alpha -= (phi + Delta) * ratio / Delta;
    if (fabs(phi) < rtol * Delta) break;
  }
  const int n_it = (it < max_iter ? it : max_iter - 1) + 1;      // Python: `it` keeps its last value after the loop
  alpha = fmax(alpha, alpha_floor);
  if ((rc = trial(h, st, alpha, 0, o, s))) return rc;
  r = {alpha, n_it, sqrt(o[0]), o[1], true};
  return BAX_OK;
}

void update_tr_radius(double& Delta, double actual, double predicted, double step_norm, bool bound_hit, double& ratio) {
  if (predicted > 0) ratio = actual / predicted;
  else if (predicted == 0 && actual == 0) ratio = 1;
  else ratio = 0;
  if (ratio < 0.25) Delta = 0.25 * step_norm;
  else if (ratio > 0.75 && bound_hit) Delta *= 2.0;
}

int check_termination(double dF, double F, double dx_norm, double x_norm, double ratio, double ftol, double xtol) {
  const bool f_ok = dF < ftol * F && ratio > 0.25, x_ok = dx_norm < xtol * (xtol + x_norm);
  if (f_ok && x_ok) return 4;
  if (f_ok) return 2;
  if (x_ok) return 3;
  return -1;
}

bool done(bax_trf_state* st) {
  if (st->status < 0 && st->g_inf < st->gtol) st->status = 1;
  return st->status >= 0 || st->nfev >= st->max_nfev;
}

}  // namespace

extern "C" int bax_trf_begin(bax_handle* h, bax_trf_state* st, void* stream) {
  if (!h || !h->campack || !st || !st->x || !st->x_new || !st->step) { bax_set_error("bax_trf_begin: bad argument / no problem"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  Lin l; int rc = linearize(h, st->x, l, s); if (rc) return rc;
  st->cost = l.cost; st->g_inf = l.g_inf; st->g_2 = l.g_2; st->x_norm = l.x_norm; st->h_scale = l.h_scale;
  st->initial_cost = l.cost;
  st->nfev = 1; st->njev = 1; st->iteration = 0; st->n_trials = 0;
  if (st->max_nfev <= 0) st->max_nfev = 100 * (6 * (int64_t)h->C + 3 * h->P);
  st->Delta = l.x_norm > 0 ? l.x_norm : 1.0;
  st->alpha = 0.0;
  st->status = -1; st->n_hist = 0;
  if (!isfinite(l.cost)) { bax_set_error("Residuals are not finite in the initial point."); return BAX_E_STATE; }
  return BAX_OK;
}

// One pass of SciPy's outer loop body: trust-region solves and (speculative) linearisations at the trial points until
// a step is accepted, evaluations run out or a termination test fires.
extern "C" int bax_trf_iterate(bax_handle* h, bax_trf_state* st, void* stream) {
  if (!h || !h->campack || !st) { bax_set_error("bax_trf_iterate: bad argument"); return BAX_E_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t n = 6 * (int64_t)h->C + 3 * h->P;
  double actual = -1.0;
  Lin ln{}; ln.cost = st->cost;
  int rc;
  st->n_hist = 0;
  while (actual <= 0 && st->nfev < st->max_nfev) {
    TrStep tr;
    if ((rc = solve_tr(h, st, st->Delta, st->g_2, st->alpha, tr, s))) return rc;
    st->alpha = tr.alpha;
    const double alpha = tr.alpha, p_norm = tr.p_norm;
    const double kappa = tr.on_boundary ? st->Delta / p_norm : 1.0;
    const double predicted = -(0.5 * kappa * kappa * (-tr.gdotp - alpha * p_norm * p_norm) + kappa * tr.gdotp);
    k_axpy_out<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(n, st->x, st->step, kappa, st->x_new);
    BAX_LAUNCH_CHECK();
    h->t_launch[BAX_T_LIN] += 1;
    if ((rc = bax_linearize_push(h))) return rc;
    if ((rc = linearize(h, st->x_new, ln, s))) return rc;
    st->nfev += 1;
    const double step_norm = kappa * p_norm;
    if (!isfinite(ln.cost)) {
      if ((rc = bax_linearize_pop(h))) return rc;
      st->Delta = 0.25 * step_norm;
      continue;
    }
    actual = st->cost - ln.cost;
    if (!(actual > 0)) { if ((rc = bax_linearize_pop(h))) return rc; }
    double Delta_new = st->Delta, ratio;
    update_tr_radius(Delta_new, actual, predicted, step_norm, step_norm > 0.95 * st->Delta, ratio);
    {
      double* hrow = st->hist[st->n_hist < BAX_TRF_MAX_HIST ? st->n_hist : BAX_TRF_MAX_HIST - 1];
      hrow[0] = alpha; hrow[1] = st->Delta; hrow[2] = step_norm; hrow[3] = ln.cost; hrow[4] = predicted; hrow[5] = actual;
      hrow[6] = (double)tr.n_it; hrow[7] = actual > 0 ? 1.0 : 0.0;
      if (st->n_hist < BAX_TRF_MAX_HIST) st->n_hist += 1;
    }
    st->status = check_termination(actual, st->cost, step_norm, st->x_norm, ratio, st->ftol, st->xtol);
    if (st->status >= 0) break;
    if (!tr.on_boundary) st->alpha = 0.0;        // SciPy's alpha after an interior step (see ba.py TRFSolver.iterate)
    st->alpha *= st->Delta / Delta_new;
    st->Delta = Delta_new;
  }
  if (actual > 0) {
    double* t = st->x; st->x = st->x_new; st->x_new = t;
    st->cost = ln.cost; st->g_inf = ln.g_inf; st->g_2 = ln.g_2; st->x_norm = ln.x_norm; st->h_scale = ln.h_scale;
    st->njev += 1;
  }
  st->iteration += 1;
  return BAX_OK;
}

// Iterate a begun solve to termination (gtol / ftol / xtol / max_nfev) without returning to the caller in between.
extern "C" int bax_trf_solve(bax_handle* h, bax_trf_state* st, void* stream) {
  if (!h || !st) { bax_set_error("bax_trf_solve: bad argument"); return BAX_E_ARG; }
  int rc;
  while (!done(st)) { if ((rc = bax_trf_iterate(h, st, stream))) return rc; }
  if (st->status < 0) st->status = 0;
  return BAX_OK;
}
