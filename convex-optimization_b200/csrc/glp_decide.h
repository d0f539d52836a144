// glp_decide.h -- the scalar control flow of the solvers as one host/device function.
//
// Restates, as a state machine advanced once per forward stream, the control flow of
//   gl_ProxGD_primal.py:53,76,103-132   and   gl_FProxGD_primal.py:55,68-69,79,108-147 :
// sufficient-decrease test of the pending trial, step shrink / untested last step, commit,
// objective + sparsity bookkeeping (f_hist, f_hist_best), the two-ratio stop rule with
// stable_len, the `while inner_iter < maxit` bound, the mu-continuation stage switch (v and t reset
// for FProxGD) and the step rule of the next iteration.
// Compiled for the device (decide kernel, persistent kernel) and for the host (CPU unit test
// against the golden vectors: tests/test_host_logic.py builds it through tests/host_shim with g++ -ffp-contract=off).
#pragma once
#include <math.h>
#include <stdint.h>

#include "glp_types.cuh"

#if defined(__CUDA_ARCH__)
#define GLP_XMUL(a, b) __dmul_rn((a), (b))
#define GLP_XADD(a, b) __dadd_rn((a), (b))
#define GLP_XSUB(a, b) __dsub_rn((a), (b))
#define GLP_XDIV(a, b) __ddiv_rn((a), (b))
#define GLP_XSQRT(a) __dsqrt_rn((a))
#else
#define GLP_XMUL(a, b) ((a) * (b))
#define GLP_XADD(a, b) ((a) + (b))
#define GLP_XSUB(a, b) ((a) - (b))
#define GLP_XDIV(a, b) ((a) / (b))
#define GLP_XSQRT(a) sqrt((a))
#endif
#if defined(__CUDACC__)
#define GLP_HDI __host__ __device__ inline
#else
#define GLP_HDI inline
#endif

namespace glp {

// Globally reduced scalars of the pending candidate.
struct Reduced {
  double rr;    // sum((A xn - b)^2)
  double gd;    // FProxGD <grad, xn - y>      | ProxGD <grad, gt>
  double dd;    // FProxGD ||xn - y||^2        | ProxGD ||gt||^2
  double reg;   // sum_i ||xn_i||_2
  double cnt;   // #{ |xn| > 1e-6 * max|xn| }
  double gbase; // sum(r^2) at the base point of the running iteration (from the previous decide)
  double sx;    // FGD: sum_i sqrt(||xn_i||^2 + delta^2) - delta   (gl_FGD_primal.py:64-67)
  double sy;    // FGD: the same at the base point y
};

enum DecideAction { GLP_ACT_NONE = 0, GLP_ACT_RETRY = 1, GLP_ACT_ITERATE = 2, GLP_ACT_FINISH = 3 };

struct DecideOut {
  int32_t action;
  int32_t reset_v;      // stage start: v <- x (un-thresholded), Av <- Ax
  int32_t first_commit; // the committed candidate was x0
  int32_t new_cur;
  double theta_old;     // theta the committed candidate was built with (Av update)
};

GLP_HDI void ctrl_init(Ctrl& c, const SolveConst& sc) {
  c.done = 0; c.stage = 0; c.inner = 0; c.k = 0; c.stable = 0; c.trial = 0; c.cur = 0; c.retry = 0;
  c.tested = 0; c.first_commit = 1; c.ran_any = 0; c.fault = 0;
  c.stage_end[0] = c.stage_end[1] = c.stage_end[2] = 0; c.gy_sel = 0;
  c.t = sc.alpha0; c.t_k = sc.alpha0; c.theta = 1.0; c.mu = 100.0 * sc.mu0;
  c.f_cur = 0.0; c.s_cur = 0.0; c.f_prev = 0.0; c.s_prev = 0.0;
  c.f_best = INFINITY; c.g_base = 0.0; c.n_fwd = 0; c.n_tpass = 0;
}

GLP_HDI double stage_mu(int stage, double mu0) {
  return stage == 0 ? GLP_XMUL(100.0, mu0) : (stage == 1 ? GLP_XMUL(10.0, mu0) : mu0);
}

// Opt-in Barzilai-Borwein step (step_type "bb", gl_ProxGD_primal only; the reference has no such rule):
// first trial step of the next line search = ||d||^2 / ||A d||^2 for the displacement d = xn - x just
// accepted.  Both norms come from scalars the line search already reduced: ||d||^2 = t^2 ||gt||^2, and
// since g is quadratic, g(xn) = g(x) + <grad, d> + 0.5 ||A d||^2 with <grad, d> = -t <grad, gt>, i.e.
// ||A d||^2 = rr_new - rr_base + 2 t <grad, gt>: no extra stream and no extra reduction.
// Safeguards: a non-positive / non-finite curvature falls back to alpha0; the step is kept within
// [alpha0 * 1e-3, alpha0 * 1e3]; the usual sufficient-decrease backtracking still runs on it.
GLP_HDI double bb_step(const SolveConst& sc, double t, const Reduced& red) {
  const double num = GLP_XMUL(GLP_XMUL(t, t), red.dd);
  const double den = GLP_XADD(GLP_XSUB(red.rr, red.gbase), GLP_XMUL(GLP_XMUL(2.0, t), red.gd));
  if (!(den > 0.0) || !(num > 0.0)) return sc.alpha0;
  const double bb = GLP_XDIV(num, den);
  const double lo = GLP_XMUL(sc.alpha0, 1e-3), hi = GLP_XMUL(sc.alpha0, 1e3);
  if (!(bb == bb)) return sc.alpha0;
  return bb < lo ? lo : (bb > hi ? hi : bb);
}

// f_hist / f_hist_best / s_hist may be null (then values are not stored).  size_total = n_total * l.
GLP_HDI void decide_step(const SolveConst& sc, const Ctrl& in, const Reduced& red, double size_total,
                         Ctrl& out, DecideOut& act, double* f_hist, double* f_hist_best, double* s_hist = nullptr) {
  out = in;
  act.action = GLP_ACT_NONE; act.reset_v = 0; act.first_commit = in.first_commit;
  act.new_cur = in.cur; act.theta_old = in.theta;
  if (in.done) return;
  const bool accel = sc.algo == ALGO_FPROXGD || sc.algo == ALGO_FGD;
  const bool plain = sc.algo == ALGO_GD || sc.algo == ALGO_SGD;   // no line search, no early stage break

  // ---- sufficient-decrease test of the pending trial
  if (in.tested) {
    const double g_x = GLP_XMUL(0.5, red.rr);
    const double g_y = GLP_XMUL(0.5, red.gbase);
    bool ok;
    if (sc.algo == ALGO_FGD) {  // g = 0.5||A.-b||^2 + mu*smoothed regulariser      gl_FGD_primal.py:64-67,98-101
      const double gx = GLP_XADD(g_x, GLP_XMUL(in.mu, red.sx)), gy = GLP_XADD(g_y, GLP_XMUL(in.mu, red.sy));
      ok = gx <= GLP_XADD(GLP_XADD(gy, red.gd), GLP_XDIV(red.dd, GLP_XMUL(2.0, in.t)));
    } else if (accel)  // g(x) <= g(y) + <grad, x-y> + ||x-y||^2/(2t)          gl_FProxGD_primal.py:99
      ok = g_x <= GLP_XADD(GLP_XADD(g_y, red.gd), GLP_XDIV(red.dd, GLP_XMUL(2.0, in.t)));
    else        // g(x - t*gt) <= g(x) - t*<grad,gt> + 0.5*t*||gt||^2    gl_ProxGD_primal.py:91-92
      ok = g_x <= GLP_XADD(GLP_XSUB(g_y, GLP_XMUL(in.t, red.gd)), GLP_XMUL(GLP_XMUL(0.5, in.t), red.dd));
    if (!ok) {
      out.t = GLP_XMUL(in.t, sc.atten);
      out.trial = in.trial + 1;
      out.tested = out.trial < sc.ls_cap ? 1 : 0;  // the (ls_cap+1)-th step is used untested
      out.retry = 1;
      out.n_fwd = in.n_fwd + 1;
      act.action = GLP_ACT_RETRY;
      return;
    }
  }

  // ---- commit the candidate
  out.cur = 1 - in.cur;
  act.new_cur = out.cur;
  out.retry = 0;
  out.first_commit = 0;
  if (accel && !in.first_commit) out.t_k = in.t;
  if (sc.step_type == STEP_BB && !in.first_commit) out.t_k = bb_step(sc, in.t, red);
  out.f_cur = GLP_XADD(GLP_XMUL(0.5, red.rr), GLP_XMUL(sc.mu0, red.reg));
  out.s_cur = GLP_XDIV(red.cnt, size_total);

  // ---- loop tops and stage switches
  bool fresh = in.first_commit != 0;
  for (;;) {
    if (fresh) {
      out.inner = 0; out.stable = 0;
      out.mu = stage_mu(out.stage, sc.mu0);
      if (accel) { out.t_k = sc.alpha0; act.reset_v = 1; }
      fresh = false;
    }
    bool advance = false;
    if (out.inner >= sc.maxit) {
      advance = true;
    } else {
      // SGD records the objective with the STAGE mu (gl_SGD_primal.py:50-53,78), everything else with mu_0
      if (sc.algo == ALGO_SGD) out.f_cur = GLP_XADD(GLP_XMUL(0.5, red.rr), GLP_XMUL(out.mu, red.reg));
      if (f_hist != nullptr && out.k < sc.hist_cap) f_hist[out.k] = out.f_cur;
      out.f_best = out.f_cur < out.f_best ? out.f_cur : out.f_best;  // min(f_best, f_now)
      if (f_hist_best != nullptr && out.k < sc.hist_cap) f_hist_best[out.k] = out.f_best;
      if (s_hist != nullptr && out.k < sc.hist_cap) s_hist[out.k] = out.s_cur;   // sparsity_hist (:116 / :119)
      out.k += 1;
      out.inner += 1;
      bool stable_now = false;
      if (out.k > 1) {
        const double df = GLP_XDIV(fabs(GLP_XSUB(out.f_cur, out.f_prev)), fabs(out.f_prev));
        const double ds = GLP_XDIV(fabs(GLP_XSUB(out.s_cur, out.s_prev)), fabs(out.s_prev));
        stable_now = (df < sc.ftol) && (ds < sc.ftol);
      }
      out.f_prev = out.f_cur;
      out.s_prev = out.s_cur;
      out.stable = stable_now ? out.stable + 1 : 0;
      if (out.stable > sc.stable_cap && !plain) advance = true;   // GD / SGD: the break is commented out (:90-91)
    }
    if (!advance) break;
    out.stage_end[out.stage] = out.k;
    out.stage += 1;
    if (out.stage == 3) {
      out.done = 1;
      act.action = GLP_ACT_FINISH;
      return;
    }
    fresh = true;
  }

  // ---- step rule and extrapolation weight of the iteration that starts now
  out.ran_any = 1;
  out.theta = accel ? GLP_XDIV(2.0, double(out.inner + 1)) : 1.0;
  out.trial = 0;
  if (plain && (sc.step_type == 1 || out.mu > sc.mu0)) {   // gl_GD_primal.py:67-70 / gl_SGD_primal.py:65-68
    out.t = sc.alpha0;
    out.tested = 0;
  } else if (sc.step_type == 0 || sc.step_type == STEP_BB) {  // line_search (ProxGD restarts at alpha0 or the BB step)
    out.t = (accel || sc.step_type == STEP_BB) ? out.t_k : sc.alpha0;
    out.tested = sc.ls_cap > 0 ? 1 : 0;
  } else {
    const int it = (out.inner > 1000 ? out.inner : 1000) - 999;
    if (sc.step_type == 1) out.t = sc.alpha0;
    else if (sc.step_type == 2) out.t = GLP_XDIV(sc.alpha0, GLP_XSQRT(double(it)));
    else out.t = GLP_XDIV(sc.alpha0, double(it));
    out.tested = 0;
  }
  out.n_fwd = in.n_fwd + 1;
  out.n_tpass = in.n_tpass + 1;
  act.action = GLP_ACT_ITERATE;
}

}  // namespace glp
