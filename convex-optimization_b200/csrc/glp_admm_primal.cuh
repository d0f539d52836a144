// glp_admm_primal.cuh -- gl_Admm_primal (code/gl_ADMM_primal.py:10-117) on the two stream kernels.
//
// One iteration of the reference is
//   y  = (rho I + A^T A)^-1 (A^T b - z + rho x)            :78   n x n Cholesky factor computed once (:62)
//   x' = prox_{eta mu}(x - eta rho (x - y - z / rho))      :80   prox_tf :47-51, guarded denominator
//   z' = z - tau rho (x' - y)                              :81
//   r = x' - y ,  s = y - y_prev                           :83-84  judged by their matrix 2-norms (:98-99)
//   f = 0.5 ||A x' - b||^2 + mu sum_i ||x'_i||             :88
// The reference calls the n x n system "factorisation-bound".  Here it is never formed: with v = A^T b - z + rho x,
//   (rho I + A^T A)^-1 v = (v - A^T (rho I + A A^T)^-1 A v) / rho = (v - A^T G^-1 (A v) / rho) / rho ,
//   G = I + A A^T / rho   (m x m),
// so the iteration is ONE forward stream over the 2l operand columns [x | v] (A x for the objective, A v for the
// solve), the m x m solve as two streams of Y = L^-T (the blocked set-up of glp_dense.cuh, shared with gl_Admm_dual),
// ONE transposed stream, and a row-local kernel.  fp64 only: the identity cancels cond(rho I + A^T A) ~ 1e5 of the
// leading digits, which fp32 streams cannot afford (the Python layer widens fp32 inputs).
// The reference iteration amplifies ANY re-rounding of y to ~4e-10 of the objective history (eta rho = 1 and z / rho
// = 100 z: its own triangular solves against numpy.linalg.solve differ by that much, tests/test_admm2_cpu.py), so
// parity is held to 5e-9, not 1e-10.
#pragma once
#include "glp_admm.cuh"

namespace glp {

// per-column-block record of the update kernel: [0, 36) Gram matrix of r (upper triangle in admm_pair order for the
// solve's l), [36, 72) Gram matrix of s, [72] sum_i ||x_i||, [73] max|x|
constexpr int ADMMP_GSTRIDE = 76;

struct AdmmPConst {
  int32_t maxit, converge_len, hist_cap, step_type;   // step_type: GLP_STEP_FIXED / _DIMINISHING / _DIMINISHING2
  double thres, tau, rho, mu, eta0;
};

__device__ __forceinline__ double admmp_eta(const AdmmPConst& c, int k) {
  if (c.step_type == 1) return xdiv(c.eta0, __dsqrt_rn(double(k)));   // eta_0 / np.sqrt(k)        :70
  if (c.step_type == 2) return xdiv(c.eta0, double(k));               // eta_0 / k                 :72
  return c.eta0;                                                       // fixed                     :68
}

struct AdmmPUpdArgs {
  const Ctrl* ctrl;        // done => no-op; k = the iteration being computed
  int32_t init;            // 1: x = x0 already in place; y_prev = z = x0 (:53-55), operands and partials only
  int32_t pad;
  double* x;               // [l][ldn]
  double* z;               // [l][ldn]
  double* yprev;           // [l][ldn]
  const double* atb;       // [l][ldn] = A^T b
  const double* grad;      // [l][ldn] = A^T G^-1 A v
  double* xop;             // [2l][ldn]: [x | v]
  double* rec;             // [ncb][ADMMP_GSTRIDE]
  double* scal_partial;    // [ncb][SCAL_STRIDE]: slot 3 = max|x| of the block, where K1's sparsity count reads it
  AdmmPConst c;
};

__global__ void glp_admm_primal_update_kernel(const AdmmPUpdArgs a, const Geom g) {
  __shared__ double s_red[2][2 * ADMM_NPAIR + 2];
  if (a.ctrl != nullptr && a.ctrl->done) return;
  constexpr int TC = Elem<double>::TILE_COLS;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t j = int64_t(blockIdx.x) * TC + tid;
  const int l = g.l;
  const int64_t ldn = g.ldn;
  double gr[ADMM_NPAIR], gs[ADMM_NPAIR], reg = 0.0, mx = 0.0;
#pragma unroll
  for (int q = 0; q < ADMM_NPAIR; ++q) { gr[q] = 0.0; gs[q] = 0.0; }
  if (j < g.n) {
    double rr[ADMM_MAX_L], ss[ADMM_MAX_L];
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c) { rr[c] = 0.0; ss[c] = 0.0; }
    double n2 = 0.0;
    if (a.init) {
#pragma unroll
      for (int c = 0; c < ADMM_MAX_L; ++c) {
        if (c >= l) break;
        const double xv = a.x[c * ldn + j];
        a.z[c * ldn + j] = xv;
        a.yprev[c * ldn + j] = xv;
        a.xop[c * ldn + j] = xv;
        a.xop[(l + c) * ldn + j] = xadd(xsub(a.atb[c * ldn + j], xv), xmul(a.c.rho, xv));   // AT_b - z_k + rho * x_k
        n2 = xadd(n2, xmul(xv, xv));
        mx = fmax(mx, fabs(xv));
      }
    } else {
      const double eta = admmp_eta(a.c, a.ctrl->k);
      const double etarho = xmul(eta, a.c.rho), taurho = xmul(a.c.tau, a.c.rho), tmu = xmul(eta, a.c.mu);
      double xv[ADMM_MAX_L], zv[ADMM_MAX_L], yv[ADMM_MAX_L], wv[ADMM_MAX_L];
      double nn = 0.0;
#pragma unroll
      for (int c = 0; c < ADMM_MAX_L; ++c) {
        xv[c] = zv[c] = yv[c] = wv[c] = 0.0;
        if (c >= l) continue;
        xv[c] = a.x[c * ldn + j];
        zv[c] = a.z[c * ldn + j];
        const double v = xadd(xsub(a.atb[c * ldn + j], zv[c]), xmul(a.c.rho, xv[c]));
        yv[c] = xdiv(xsub(v, xdiv(__ldcg(&a.grad[c * ldn + j]), a.c.rho)), a.c.rho);   // (rho I + A^T A)^-1 v       :78
        wv[c] = xsub(xv[c], xmul(etarho, xsub(xsub(xv[c], yv[c]), xdiv(zv[c], a.c.rho))));   // :80
        nn = xadd(nn, xmul(wv[c], wv[c]));
      }
      const double rn = __dsqrt_rn(nn);
      const double shr = fmax(xsub(rn, tmu), 0.0);                      // np.clip(row_norms - t_mu, a_min=0)      :50
      const double den = xadd(rn < a.c.thres ? 1.0 : 0.0, rn);           // (row_norms < thres) + row_norms
#pragma unroll
      for (int c = 0; c < ADMM_MAX_L; ++c) {
        if (c >= l) break;
        const double xn = xdiv(xmul(wv[c], shr), den);
        rr[c] = xsub(xn, yv[c]);                                         // r_k = x - y                              :83
        ss[c] = xsub(yv[c], a.yprev[c * ldn + j]);                       // s_k = y - y_k                            :84
        const double zn = xsub(zv[c], xmul(taurho, rr[c]));              // z_k - tau * rho * (x - y)                :81
        a.x[c * ldn + j] = xn;
        a.z[c * ldn + j] = zn;
        a.yprev[c * ldn + j] = yv[c];
        a.xop[c * ldn + j] = xn;
        a.xop[(l + c) * ldn + j] = xadd(xsub(a.atb[c * ldn + j], zn), xmul(a.c.rho, xn));
        n2 = xadd(n2, xmul(xn, xn));
        mx = fmax(mx, fabs(xn));
      }
    }
    reg = __dsqrt_rn(n2);
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c)
#pragma unroll
      for (int d = c; d < ADMM_MAX_L; ++d) {
        gr[admm_pair(c, d, ADMM_MAX_L)] = rr[c] * rr[d];
        gs[admm_pair(c, d, ADMM_MAX_L)] = ss[c] * ss[d];
      }
  }
#pragma unroll
  for (int q = 0; q < ADMM_NPAIR; ++q) {
    const double vr = warp_sum(gr[q]), vs = warp_sum(gs[q]);
    if (lane == 0) { s_red[warp][q] = vr; s_red[warp][ADMM_NPAIR + q] = vs; }
  }
  reg = warp_sum(reg);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) { s_red[warp][2 * ADMM_NPAIR] = reg; s_red[warp][2 * ADMM_NPAIR + 1] = mx; }
  __syncthreads();
  double* out = a.rec + ADMMP_GSTRIDE * int64_t(blockIdx.x);
  if (tid < ADMM_NPAIR) {
    // fixed index (pair of an 8 x 8 triangle) -> (c, d) -> index in the l x l triangle the mid kernel reads
    int c = 0, rem = tid;
    while (rem >= ADMM_MAX_L - c) { rem -= ADMM_MAX_L - c; ++c; }
    const int d = c + rem;
    if (d < l) {
      out[admm_pair(c, d, l)] = s_red[0][tid] + s_red[1][tid];
      out[ADMM_NPAIR + admm_pair(c, d, l)] = s_red[0][ADMM_NPAIR + tid] + s_red[1][ADMM_NPAIR + tid];
    }
  } else if (tid == ADMM_NPAIR) {
    out[2 * ADMM_NPAIR] = s_red[0][2 * ADMM_NPAIR] + s_red[1][2 * ADMM_NPAIR];
  } else if (tid == ADMM_NPAIR + 1) {
    const double t = fmax(s_red[0][2 * ADMM_NPAIR + 1], s_red[1][2 * ADMM_NPAIR + 1]);
    out[2 * ADMM_NPAIR + 1] = t;
    a.scal_partial[SCAL_STRIDE * int64_t(blockIdx.x) + 3] = t;
  }
}

// ---- mid: evaluate the iteration whose products just arrived, advance the control, hand A v to the m x m solve ----
struct AdmmPMidArgs {
  const Ctrl* cin;
  Ctrl* cout;
  AdmmPConst c;
  const double* P;            // [2l][ldm] = A [x | v]
  const double* b;            // [l][ldm]
  const double* rec; int32_t n_rec;            // [n_rec][ADMMP_GSTRIDE] from the update kernel
  const double* cnt_partial;  int32_t n_cnt;
  double* f_hist;
  double* f_hist_best;
  double size_total;          // n * l
  double* w_out;              // [l][w_ld]: A v, the right-hand side of the streams of Y
  int64_t w_ld;
};

__global__ void __launch_bounds__(ADMM_THREADS) glp_admm_primal_mid_kernel(const AdmmPMidArgs a, const Geom g) {
  constexpr int NV = 2 + 2 * ADMM_NPAIR;            // rr, reg, Gram of r (upper), Gram of s (upper)
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
  const double* P1 = a.P;
  const double* P2 = a.P + int64_t(l) * ldm;
  double v[NV], cnt = 0.0;
#pragma unroll
  for (int q = 0; q < NV; ++q) v[q] = 0.0;
  for (int i = tid; i < m; i += ADMM_THREADS)
    for (int c = 0; c < l; ++c) {
      const double d = xsub(P1[c * ldm + i], a.b[c * ldm + i]);
      v[0] += d * d;
    }
  for (int i = tid; i < a.n_rec; i += ADMM_THREADS) {
    const double* p = a.rec + ADMMP_GSTRIDE * int64_t(i);
#pragma unroll
    for (int q = 0; q < ADMM_NPAIR; ++q)
      if (q < npair) { v[2 + q] += __ldcg(p + q); v[2 + ADMM_NPAIR + q] += __ldcg(p + ADMM_NPAIR + q); }
    v[1] += __ldcg(p + 2 * ADMM_NPAIR);
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
    if (cin.k > 0) {   // iteration cin.k has just been completed by the forward stream: record and test it (:87-107)
      if (a.f_hist != nullptr && cin.k - 1 < a.c.hist_cap) a.f_hist[cin.k - 1] = f_now;
      out.f_best = f_now < cin.f_best ? f_now : cin.f_best;
      if (a.f_hist_best != nullptr && cin.k - 1 < a.c.hist_cap) a.f_hist_best[cin.k - 1] = out.f_best;
      const double* rgt = s_tot + 2;
      const double* sgt = s_tot + 2 + ADMM_NPAIR;
      const double rn = l <= 2 ? admm_norm2(rgt[0], l > 1 ? rgt[1] : 0.0, l > 1 ? rgt[2] : 0.0, l) : admm_norm2_jacobi(rgt, l);
      const double sn = l <= 2 ? admm_norm2(sgt[0], l > 1 ? sgt[1] : 0.0, l > 1 ? sgt[2] : 0.0, l) : admm_norm2_jacobi(sgt, l);
      out.stable = (rn < a.c.thres && sn < a.c.thres) ? cin.stable + 1 : 0;
      if (out.stable >= a.c.converge_len) out.done = 1;
    }
    if (!out.done) {
      if (cin.k < a.c.maxit) { out.k = cin.k + 1; out.n_fwd = cin.n_fwd + 1; out.n_tpass = cin.n_tpass + 1; }
      else out.done = 1;
    }
    s_done = out.done;
    *a.cout = out;
  }
  __syncthreads();
  if (s_done) return;
  for (int i = tid; i < m; i += ADMM_THREADS)
    for (int c = 0; c < l; ++c) a.w_out[c * a.w_ld + i] = P2[c * ldm + i];
}

}  // namespace glp
