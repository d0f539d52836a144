// glp_alm.cuh -- gl_Alm_dual (code/gl_ALM_dual.py:10-158) on the two stream kernels.
//
// One OUTER iteration of the reference (:119-147) is
//   u  = argmin over the row-wise mu-ball of the augmented Lagrangian with z eliminated       :122  solve_sub_problem
//   z  = (I + rho A A^T)^-1 (A (x - rho u) - b)                                               :123
//   x' = x - tau rho (u + A^T z) ;  r = u + A^T z ,  s = A (u_prev - u)                       :124-127
// and the inner solver (:10-63) forms T = G^-1 (G = I + rho A A^T), F = rho T A, the n x n matrices
// Gm = I - A^T F, Q = F^T F + rho Gm^T Gm and J = rho Gm^T H - F^T E - F^T b explicitly, then runs 500 accelerated
// projected-gradient steps with gradient Q u + J.  Here nothing n x n is formed.  With G T = I (so rho A A^T T = I - T):
//   Gm^2 = I - rho A^T T A - rho A^T T^2 A      =>   Q = rho Gm = rho (I - rho A^T T A)
//   A H  = A (A^T E - x / rho) = -(E + b) / rho =>   J = rho H            (E = T (A x - b), H = A^T E - x / rho)
// i.e. grad(y) = rho (y - rho A^T T (A y)) + rho H: ONE forward stream (A u of the new inner iterate; A y follows by
// linearity from the carried A u, A v), the m x m solve as two streams of Y = L^-T, ONE transposed stream and a
// row-local kernel per inner step -- the same pipeline as an iteration of gl_Admm_dual.  The outer step needs no
// further forward stream either: A (x - rho u) - b = (A x - b) - rho A u.  A x' for the objective is the forward stream
// that opens the next outer iteration.  The identities are exact in exact arithmetic; measured against the unmodified
// reference the objective history agrees to ~1e-15 (NumPy prototype) -- the inner loop is a contraction.
// fp64 only (1 / (1 + rho lambda) ~ 1e-5 of the leading digits cancel in y - rho A^T T A y).
#pragma once
#include "glp_admm.cuh"

namespace glp {

constexpr int ALM_INNER_STEPS = 500;      // gl_ALM_dual.py:50
constexpr double ALM_INNER_T = 1e-2;      // gl_ALM_dual.py:54
constexpr int ALM_THREADS = 256;

struct AlmConst {
  int32_t maxit, converge_len, hist_cap, pad;
  double thres, tau, rho, mu;
};

// ---- outer evaluation (single CTA): record / test the outer iteration whose A x just arrived, advance the control,
//      D = A x - b -> the right-hand side of the solve for E, and a copy for the z step
struct AlmEvalArgs {
  const Ctrl* cin;
  Ctrl* cout;
  AlmConst c;
  const double* P;            // [l][ldm] = A x
  const double* b;            // [l][ldm]
  const double* Au;           // [l][ldm] = A u of the outer iteration being evaluated
  double* Au_prev;            // [l][ldm] = A u of the one before; <- Au
  const double* rec; int32_t n_rec;   // [n_rec][ADMM_GSTRIDE]: Gram of r, sum ||x_i||, max|x| from the outer update
  const double* cnt_partial;  int32_t n_cnt;
  double* f_hist;
  double* f_hist_best;
  double size_total;
  double* D;                  // [l][ldm]
  double* w_out; int64_t w_ld;
};

__global__ void __launch_bounds__(ADMM_THREADS) glp_alm_eval_kernel(const AlmEvalArgs a, const Geom g) {
  constexpr int NV = 2 + 2 * ADMM_NPAIR;            // rr, reg, Gram of s (upper), Gram of r (upper)
  __shared__ double s_red[ADMM_THREADS / 32][NV];
  __shared__ double s_tot[NV];
  __shared__ double s_cnt[ADMM_THREADS / 32];
  __shared__ int s_done;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const Ctrl cin = *a.cin;
  if (cin.done) {
    if (tid == 0) *a.cout = cin;
    return;
  }
  const int l = g.l, m = g.m, npair = l * (l + 1) / 2;
  const int64_t ldm = g.ldm;
  double v[NV], cnt = 0.0;
#pragma unroll
  for (int q = 0; q < NV; ++q) v[q] = 0.0;
  for (int i = tid; i < m; i += ADMM_THREADS) {
    double sv[ADMM_MAX_L];
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c) {
      sv[c] = 0.0;
      if (c < l) {
        const double d = xsub(a.P[c * ldm + i], a.b[c * ldm + i]);
        v[0] += d * d;
        sv[c] = xsub(a.Au_prev[c * ldm + i], a.Au[c * ldm + i]);      // s_k = A @ (u_k - u)            :127
      }
    }
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c)
#pragma unroll
      for (int d = c; d < ADMM_MAX_L; ++d)
        if (d < l) v[2 + admm_pair(c, d, l)] += sv[c] * sv[d];
  }
  for (int i = tid; i < a.n_rec; i += ADMM_THREADS) {
    const double* p = a.rec + ADMM_GSTRIDE * int64_t(i);
#pragma unroll
    for (int q = 0; q < ADMM_NPAIR; ++q)
      if (q < npair) v[2 + ADMM_NPAIR + q] += __ldcg(p + q);
    v[1] += __ldcg(p + ADMM_NPAIR);
  }
  for (int i = tid; i < a.n_cnt; i += ADMM_THREADS) cnt += __ldcg(&a.cnt_partial[i]);
#pragma unroll
  for (int q = 0; q < NV; ++q) {
    const double t = warp_sum(v[q]);
    if (lane == 0) s_red[warp][q] = t;
  }
  cnt = warp_sum(cnt);
  if (lane == 0) s_cnt[warp] = cnt;
  __syncthreads();
  if (tid < NV) {
    double t = 0.0;
    for (int w = 0; w < ADMM_THREADS / 32; ++w) t += s_red[w][tid];
    s_tot[tid] = t;
  }
  __syncthreads();
  if (tid == 0) {
    double ctot = 0.0;
    for (int w = 0; w < ADMM_THREADS / 32; ++w) ctot += s_cnt[w];
    Ctrl out = cin;
    const double f_now = xadd(xmul(0.5, s_tot[0]), xmul(a.c.mu, s_tot[1]));
    out.f_cur = f_now;
    out.s_cur = ctot / a.size_total;
    if (cin.k > 0) {   // outer iteration cin.k is complete: record and test it (:130-147)
      if (a.f_hist != nullptr && cin.k - 1 < a.c.hist_cap) a.f_hist[cin.k - 1] = f_now;
      out.f_best = f_now < cin.f_best ? f_now : cin.f_best;
      if (a.f_hist_best != nullptr && cin.k - 1 < a.c.hist_cap) a.f_hist_best[cin.k - 1] = out.f_best;
      const double* sgt = s_tot + 2;
      const double* rgt = s_tot + 2 + ADMM_NPAIR;
      const double rn = l <= 2 ? admm_norm2(rgt[0], l > 1 ? rgt[1] : 0.0, l > 1 ? rgt[2] : 0.0, l) : admm_norm2_jacobi(rgt, l);
      const double sn = l <= 2 ? admm_norm2(sgt[0], l > 1 ? sgt[1] : 0.0, l > 1 ? sgt[2] : 0.0, l) : admm_norm2_jacobi(sgt, l);
      out.stable = (rn < a.c.thres && sn < a.c.thres) ? cin.stable + 1 : 0;
      if (out.stable >= a.c.converge_len) out.done = 1;
    }
    if (!out.done) {
      if (cin.k < a.c.maxit) out.k = cin.k + 1;
      else out.done = 1;
    }
    out.n_fwd = cin.n_fwd + (out.done ? 0 : ALM_INNER_STEPS + 1);
    out.n_tpass = cin.n_tpass + (out.done ? 0 : ALM_INNER_STEPS + 1);
    s_done = out.done;
    *a.cout = out;
  }
  __syncthreads();
  if (s_done) return;
  for (int i = tid; i < m; i += ADMM_THREADS)
    for (int c = 0; c < l; ++c) {
      const double d = xsub(a.P[c * ldm + i], a.b[c * ldm + i]);     // D = A @ x_k - b                 :31
      a.D[c * ldm + i] = d;
      a.w_out[c * a.w_ld + i] = d;
      a.Au_prev[c * ldm + i] = a.Au[c * ldm + i];
    }
}

// ---- m-space element kernels -----------------------------------------------------------------------------------
// mode 0 (inner step kk >= 2): fold the product of the previous inner iterate into the carried pair
//        A v <- A u + (P - A u) / gamma_prev ,  A u <- P        (v = u_k + (u - u_k) / gamma, :59), gamma_prev = 2 / kk
//        and hand  A y = (1 - gamma) A u + gamma A v  to the solve (y = (1 - gamma) u + gamma v, :56), gamma = 2/(kk+1)
// mode 1 (after the inner loop): A u <- P ;  w = D - rho A u   ( = A (x - rho u) - b, :123 )
struct AlmMArgs {
  const Ctrl* ctrl;
  int32_t mode, kk;
  const double* P;      // [l][ldm] product of the latest inner iterate
  double* Au;
  double* Av;
  const double* D;
  double rho;
  double* w_out; int64_t w_ld;
};

__global__ void __launch_bounds__(ALM_THREADS) glp_alm_m_kernel(const AlmMArgs a, const Geom g) {
  if (a.ctrl->done) return;
  const int64_t ldm = g.ldm, total = int64_t(g.l) * ldm;
  const double gp = 2.0 / double(a.kk), gm = 2.0 / double(a.kk + 1);
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int c = int(e / ldm);
    const int64_t i = e - int64_t(c) * ldm;
    if (i >= g.m) continue;
    const double p = a.P[e];
    if (a.mode == 0) {
      const double au_old = a.kk == 2 ? 0.0 : a.Au[e];
      const double av = xadd(au_old, xdiv(xsub(p, au_old), gp));
      a.Au[e] = p;
      a.Av[e] = av;
      a.w_out[c * a.w_ld + i] = xadd(xmul(xsub(1.0, gm), p), xmul(gm, av));
    } else {
      a.Au[e] = p;
      a.w_out[c * a.w_ld + i] = xsub(a.D[e], xmul(a.rho, p));
    }
  }
}

// ---- n-space row kernels -----------------------------------------------------------------------------------------
// mode 0: start of an outer iteration: J = rho H = rho (A^T E - x / rho)  (:36, :40 with the identities above),
//         u = v = 0 (:47-48)
// mode 1: inner step kk: y = (1 - gamma) u + gamma v ; w = y - t (rho (y - rho grad) + J) ; u' = P_mu(w) ;
//         v = u + (u' - u) / gamma ; operand <- u'   (:53-60; grad = A^T T A y, taken as 0 for kk = 1 where y = 0)
// mode 2: outer step: r = u + grad (grad = A^T z) ; x <- x - tau rho r ; operand <- x ; per-block record
// mode 3: before the first outer iteration: operand <- x0 and the per-block record of x0 (objective / sparsity of k = 0)
struct AlmNArgs {
  const Ctrl* ctrl;
  int32_t mode, kk;
  double* x;            // [l][ldn]
  double* u;            // [l][ldn] inner iterate u_k
  double* v;            // [l][ldn]
  double* J;            // [l][ldn]
  const double* grad;   // [l][ldn]
  double* xop;          // [l][ldn] forward operand
  double* rec;          // [ncb][ADMM_GSTRIDE]
  double* scal_partial;
  AlmConst c;
};

__global__ void glp_alm_n_kernel(const AlmNArgs a, const Geom g) {
  __shared__ double s_red[2][ADMM_NPAIR + 2];
  if (a.ctrl->done) return;
  constexpr int TC = Elem<double>::TILE_COLS;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t j = int64_t(blockIdx.x) * TC + tid;
  const int l = g.l;
  const int64_t ldn = g.ldn;
  const bool live = j < g.n;
  if (a.mode == 0) {
    if (!live) return;
    for (int c = 0; c < l; ++c) {
      const double h = xsub(__ldcg(&a.grad[c * ldn + j]), xdiv(a.x[c * ldn + j], a.c.rho));
      a.J[c * ldn + j] = xmul(a.c.rho, h);
      a.u[c * ldn + j] = 0.0;
      a.v[c * ldn + j] = 0.0;
    }
    return;
  }
  if (a.mode == 1) {
    if (!live) return;
    const double gm = 2.0 / double(a.kk + 1);
    double uu[ADMM_MAX_L], wv[ADMM_MAX_L], nn = 0.0;
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c) {
      uu[c] = wv[c] = 0.0;
      if (c >= l) continue;
      uu[c] = a.u[c * ldn + j];
      const double y = xadd(xmul(xsub(1.0, gm), uu[c]), xmul(gm, a.v[c * ldn + j]));
      const double gr = a.kk == 1 ? 0.0 : __ldcg(&a.grad[c * ldn + j]);
      const double gf = xadd(xmul(a.c.rho, xsub(y, xmul(a.c.rho, gr))), a.J[c * ldn + j]);
      wv[c] = xsub(y, xmul(ALM_INNER_T, gf));
      nn = xadd(nn, xmul(wv[c], wv[c]));
    }
    const double rn = __dsqrt_rn(nn);
    const double den = rn < a.c.mu ? a.c.mu : rn;              // np.clip(row_norms, a_min=mu)           :45
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c) {
      if (c >= l) break;
      const double un = xdiv(xmul(a.c.mu, wv[c]), den);
      a.v[c * ldn + j] = xadd(uu[c], xdiv(xsub(un, uu[c]), gm));
      a.u[c * ldn + j] = un;
      a.xop[c * ldn + j] = un;
    }
    return;
  }
  // modes 2 and 3
  double gr[ADMM_NPAIR], reg = 0.0, mx = 0.0;
#pragma unroll
  for (int q = 0; q < ADMM_NPAIR; ++q) gr[q] = 0.0;
  if (live) {
    const double taurho = xmul(a.c.tau, a.c.rho);
    double rr[ADMM_MAX_L], n2 = 0.0;
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c) {
      rr[c] = 0.0;
      if (c >= l) continue;
      double xn = a.x[c * ldn + j];
      if (a.mode == 2) {
        rr[c] = xadd(a.u[c * ldn + j], __ldcg(&a.grad[c * ldn + j]));        // r_k = u + A.T @ z            :126
        xn = xsub(xn, xmul(taurho, rr[c]));                                  // x_k - tau * rho * (u + A.T @ z)  :124
      }
      a.x[c * ldn + j] = xn;
      a.xop[c * ldn + j] = xn;
      n2 = xadd(n2, xmul(xn, xn));
      mx = fmax(mx, fabs(xn));
    }
    reg = __dsqrt_rn(n2);
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c)
#pragma unroll
      for (int d = c; d < ADMM_MAX_L; ++d) gr[admm_pair(c, d, ADMM_MAX_L)] = rr[c] * rr[d];
  }
#pragma unroll
  for (int q = 0; q < ADMM_NPAIR; ++q) {
    const double vr = warp_sum(gr[q]);
    if (lane == 0) s_red[warp][q] = vr;
  }
  reg = warp_sum(reg);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) { s_red[warp][ADMM_NPAIR] = reg; s_red[warp][ADMM_NPAIR + 1] = mx; }
  __syncthreads();
  double* out = a.rec + ADMM_GSTRIDE * int64_t(blockIdx.x);
  if (tid < ADMM_NPAIR) {
    int c = 0, rem = tid;
    while (rem >= ADMM_MAX_L - c) { rem -= ADMM_MAX_L - c; ++c; }
    const int d = c + rem;
    if (d < l) out[admm_pair(c, d, l)] = s_red[0][tid] + s_red[1][tid];
  } else if (tid == ADMM_NPAIR) {
    out[ADMM_NPAIR] = s_red[0][ADMM_NPAIR] + s_red[1][ADMM_NPAIR];
  } else if (tid == ADMM_NPAIR + 1) {
    const double t = fmax(s_red[0][ADMM_NPAIR + 1], s_red[1][ADMM_NPAIR + 1]);
    out[ADMM_NPAIR + 1] = t;
    a.scal_partial[SCAL_STRIDE * int64_t(blockIdx.x) + 3] = t;
  }
}

}  // namespace glp
