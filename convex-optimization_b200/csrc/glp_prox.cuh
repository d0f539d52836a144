// glp_prox.cuh -- the row-group epilogue of the transposed pass (K2's fused tail), also launched
// stand-alone for line-search retries and for the initial iterate.
//
// For every row group j of x (a column j of A) it evaluates, in the reference's operation order,
//   FProxGD  y   = (1-theta)*xt + theta*v                          gl_FProxGD_primal.py:139
//            xn  = prox_th(y - t*grad, t)                           :71-77, :144
//            vn  = xt + (xn - xt)/theta                             :145
//            gd  = sum(grad*(xn-y)),  dd = sum((xn-y)**2)           :96-99
//   ProxGD   xn  = prox_th(xt - t*grad, t)                          gl_ProxGD_primal.py:65-71, :132
//            gt  = (xt - xn)/t ; gd = sum(grad*gt), dd = sum(gt**2) :73-74, :89-92
//   both     xnt = xn with |.|<thres zeroed                         :127 / :136 (applied next iteration)
//            reg = sum_i ||xn_i||_2 , mx = max|xn|                  :24-27, :22
// prox_th keeps the reference's guard: z*max(rn-t*mu,0) / ((rn<thres)+rn), multiply before divide.
#pragma once
#include "glp_comm.cuh"
#include "glp_device.cuh"
#include "glp_types.cuh"

namespace glp {

struct ProxScal {
  double gd, dd, reg, mx;
  double sx, sy;  // FGD: smoothed regulariser at the candidate / at the base point
};

__device__ __forceinline__ double nmax(double a, double b) {
  return (a != a) ? a : ((b != b) ? b : fmax(a, b));
}

struct ProxResolved {
  int algo, cur, init;
  double theta, t, mu, thres, delta;
};

__device__ __forceinline__ ProxResolved resolve(const ProxArgs& p) {
  ProxResolved r;
  r.algo = p.algo;
  r.init = p.init;
  r.thres = p.thres;
  r.delta = p.delta;
  if (p.ctrl != nullptr) {
    r.cur = p.ctrl->cur; r.theta = p.ctrl->theta; r.t = p.ctrl->t; r.mu = p.ctrl->mu;
  } else {
    r.cur = p.cur; r.theta = p.theta; r.t = p.t; r.mu = p.mu;
  }
  return r;
}

// SMEM = true: every operand (state, grad) lives in shared memory (persistent kernel); the gradient is
// then read with a plain load (ld.global.cg on a shared address would fault).
template <typename T, bool SMEM = false>
__device__ __forceinline__ void prox_group(const ProxArgs& p, const ProxResolved& r, const Geom& g,
                                           int64_t j, ProxScal& sc) {
  const int l = g.l;
  const int64_t ldn = g.ldn;
  const double* S = p.state[r.cur];
  double* C = p.state[1 - r.cur];
  const double* xt = S + int64_t(l) * ldn;
  const double* v = S + 2 * int64_t(l) * ldn;
  double* xn = C;
  double* xnt = C + int64_t(l) * ldn;
  double* vn = C + 2 * int64_t(l) * ldn;
  T* xop = static_cast<T*>(p.xop);
  const bool accel = r.algo == ALGO_FPROXGD || r.algo == ALGO_FGD;

  if (r.init) {
    double nn = 0.0;
    for (int c = 0; c < l; ++c) {
      const double x = xn[c * ldn + j];
      const double xz = fabs(x) < r.thres ? 0.0 : x;
      xnt[c * ldn + j] = xz;
      if (xop != nullptr) { xop[c * ldn + j] = T(x); xop[(l + c) * ldn + j] = T(xz); }
      nn = xadd(nn, xmul(x, x));
      sc.mx = nmax(sc.mx, fabs(x));
    }
    sc.reg = xadd(sc.reg, __dsqrt_rn(nn));
    return;
  }

  const double omt = xsub(1.0, r.theta);
  if (r.algo >= ALGO_FGD) {
    // ---- smoothed / subgradient solvers: prox = identity, the regulariser enters
    // through its gradient at the base point.  Operation order of gl_FGD_primal.py:69-72,139-145,
    // gl_GD_primal.py:59-63,92-94 and gl_SGD_primal.py:56-61,90-92.
    double nb = 0.0;
    for (int c = 0; c < l; ++c) {
      const double xtc = xt[c * ldn + j];
      const double base = accel ? xadd(xmul(omt, xtc), xmul(r.theta, v[c * ldn + j])) : xtc;
      nb = xadd(nb, xmul(base, base));
    }
    double den;
    if (r.algo == ALGO_SGD) {
      const double rn = __dsqrt_rn(nb);
      den = xadd(rn < r.thres ? 1.0 : 0.0, rn);
    } else {
      den = __dsqrt_rn(xadd(nb, xmul(r.delta, r.delta)));
    }
    double n2 = 0.0;
    for (int c = 0; c < l; ++c) {
      const double xtc = xt[c * ldn + j];
      const double base = accel ? xadd(xmul(omt, xtc), xmul(r.theta, v[c * ldn + j])) : xtc;
      const double gr = SMEM ? p.grad[c * ldn + j] : __ldcg(&p.grad[c * ldn + j]);
      const double gtot = xadd(gr, xmul(r.mu, xdiv(base, den)));
      const double o = xsub(base, xmul(r.t, gtot));
      const double oz = fabs(o) < r.thres ? 0.0 : o;
      xn[c * ldn + j] = o;
      xnt[c * ldn + j] = oz;
      if (xop != nullptr) { xop[c * ldn + j] = T(o); xop[(l + c) * ldn + j] = T(oz); }
      if (accel) vn[c * ldn + j] = xadd(xtc, xdiv(xsub(o, xtc), r.theta));
      const double d = xsub(o, base);
      sc.gd = xadd(sc.gd, xmul(gtot, d));
      sc.dd = xadd(sc.dd, xmul(d, d));
      n2 = xadd(n2, xmul(o, o));
      sc.mx = nmax(sc.mx, fabs(o));
    }
    sc.reg = xadd(sc.reg, __dsqrt_rn(n2));
    if (r.algo == ALGO_FGD) {
      const double d2 = xmul(r.delta, r.delta);
      sc.sx = xadd(sc.sx, xsub(__dsqrt_rn(xadd(n2, d2)), r.delta));
      sc.sy = xadd(sc.sy, xsub(__dsqrt_rn(xadd(nb, d2)), r.delta));
    }
    return;
  }
  const double tmu = xmul(r.t, r.mu);
  double nn = 0.0;
  for (int c = 0; c < l; ++c) {
    const double xtc = xt[c * ldn + j];
    const double base = accel ? xadd(xmul(omt, xtc), xmul(r.theta, v[c * ldn + j])) : xtc;
    const double z = xsub(base, xmul(r.t, SMEM ? p.grad[c * ldn + j] : __ldcg(&p.grad[c * ldn + j])));
    nn = xadd(nn, xmul(z, z));
  }
  const double rn = __dsqrt_rn(nn);
  double shrink = xsub(rn, tmu);
  shrink = shrink < 0.0 ? 0.0 : shrink;  // np.clip(.., a_min=0): NaN stays NaN
  const double den = xadd(rn < r.thres ? 1.0 : 0.0, rn);
  double n2 = 0.0;
  for (int c = 0; c < l; ++c) {
    const double xtc = xt[c * ldn + j];
    const double base = accel ? xadd(xmul(omt, xtc), xmul(r.theta, v[c * ldn + j])) : xtc;
    const double gr = SMEM ? p.grad[c * ldn + j] : __ldcg(&p.grad[c * ldn + j]);
    const double z = xsub(base, xmul(r.t, gr));
    const double o = xdiv(xmul(z, shrink), den);
    const double oz = fabs(o) < r.thres ? 0.0 : o;
    xn[c * ldn + j] = o;
    xnt[c * ldn + j] = oz;
    if (xop != nullptr) { xop[c * ldn + j] = T(o); xop[(l + c) * ldn + j] = T(oz); }
    double d;
    if (accel) {
      d = xsub(o, base);
      vn[c * ldn + j] = xadd(xtc, xdiv(xsub(o, xtc), r.theta));
    } else {
      d = xdiv(xsub(base, o), r.t);
    }
    sc.gd = xadd(sc.gd, xmul(gr, d));
    sc.dd = xadd(sc.dd, xmul(d, d));
    n2 = xadd(n2, xmul(o, o));
    sc.mx = nmax(sc.mx, fabs(o));
  }
  sc.reg = xadd(sc.reg, __dsqrt_rn(n2));
}

// Column block cb (TILE_COLS row groups) handled by the 256 threads tid = 0..255 of one CTA that
// all call this function; the scalar partials land in scal_partial[cb][0..3] in a fixed order.
template <typename T>
__device__ __forceinline__ void prox_block(const ProxArgs& p, const ProxResolved& r, const Geom& g,
                                           int64_t cb, int tid, double (*s_red)[6]) {
  constexpr int TC = Elem<T>::TILE_COLS;
  ProxScal sc = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  for (int jj = tid; jj < TC; jj += CONSUMER_THREADS) {
    const int64_t j = cb * TC + jj;
    if (j < g.n) prox_group<T>(p, r, g, j, sc);
  }
  sc.gd = warp_sum(sc.gd);
  sc.dd = warp_sum(sc.dd);
  sc.reg = warp_sum(sc.reg);
  sc.sx = warp_sum(sc.sx);
  sc.sy = warp_sum(sc.sy);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sc.mx = nmax(sc.mx, __shfl_xor_sync(0xffffffffu, sc.mx, o));
  const int warp = tid >> 5, lane = tid & 31;
  if (lane == 0) {
    s_red[warp][0] = sc.gd; s_red[warp][1] = sc.dd; s_red[warp][2] = sc.reg; s_red[warp][3] = sc.mx;
    s_red[warp][4] = sc.sx; s_red[warp][5] = sc.sy;
  }
  consumer_bar();
  if (tid == 0) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0, a4 = 0.0, a5 = 0.0;
#pragma unroll
    for (int w = 0; w < NUM_CONSUMER_WARPS; ++w) {
      a0 += s_red[w][0]; a1 += s_red[w][1]; a2 += s_red[w][2]; a3 = nmax(a3, s_red[w][3]);
      a4 += s_red[w][4]; a5 += s_red[w][5];
    }
    double* out = p.scal_partial + SCAL_STRIDE * cb;
    out[0] = a0; out[1] = a1; out[2] = a2; out[3] = a3; out[4] = a4; out[5] = a5;
  }
  consumer_bar();
}

// ---- wide row groups (l >= 16): the same epilogue with the l columns of a row group spread over the 8
// warps of the CTA.  prox_group walks the columns of one row group serially (two dependent passes of l
// loads: 288 us for n = 65536, l = 128, latency-bound); here warp w takes the columns c = w, w+8, ... of
// 32 row groups at a time, the two row norms are combined through shared memory in warp order, and the
// loads of a pass are independent.  The summation order over the columns differs from NumPy's
// left-to-right add.reduce by rounding only (wide problems are judged at 1e-10 / 1e-5, not bit for bit).
template <typename T>
__device__ __forceinline__ void prox_block_wide(const ProxArgs& p, const ProxResolved& r, const Geom& g,
                                                int64_t cb, int tid, double (*s_red)[6], double (*s_row)[32]) {
  constexpr int TC = Elem<T>::TILE_COLS;
  constexpr int NW = NUM_CONSUMER_WARPS;
  const int warp = tid >> 5, lane = tid & 31;
  const int l = g.l;
  const int64_t ldn = g.ldn, ll = int64_t(l) * ldn;
  const double* S = p.state[r.cur];
  double* C = p.state[1 - r.cur];
  const double* xt = S + ll;
  const double* v = S + 2 * ll;
  double* xn = C;
  double* xnt = C + ll;
  double* vn = C + 2 * ll;
  T* xop = static_cast<T*>(p.xop);
  const bool accel = r.algo == ALGO_FPROXGD || r.algo == ALGO_FGD;
  const bool smooth = r.algo >= ALGO_FGD;
  const double omt = xsub(1.0, r.theta), tmu = xmul(r.t, r.mu), d2 = xmul(r.delta, r.delta);
  ProxScal sc = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  // fixed-order sum over the 8 warps of one per-row value; every thread of a row's lane gets the result
  auto row_sum = [&](double part) {
    consumer_bar();
    s_row[warp][lane] = part;
    consumer_bar();
    double t = s_row[0][lane];
#pragma unroll
    for (int w = 1; w < NW; ++w) t = xadd(t, s_row[w][lane]);
    return t;
  };
  for (int rb = 0; rb < TC / 32; ++rb) {
    const int64_t j = cb * TC + rb * 32 + lane;
    const bool live = j < g.n;
    if (r.init) {
      double nn = 0.0;
      if (live)
#pragma unroll 4
        for (int c = warp; c < l; c += NW) {
          const double x = xn[c * ldn + j];
          const double xz = fabs(x) < r.thres ? 0.0 : x;
          xnt[c * ldn + j] = xz;
          if (xop != nullptr) { xop[c * ldn + j] = T(x); xop[(l + c) * ldn + j] = T(xz); }
          nn = xadd(nn, xmul(x, x));
          sc.mx = nmax(sc.mx, fabs(x));
        }
      nn = row_sum(nn);
      if (warp == 0 && live) sc.reg = xadd(sc.reg, __dsqrt_rn(nn));
      continue;
    }
    // pass 1: squared norm of the prox argument (prox solvers) or of the base point (smooth solvers)
    double n1 = 0.0;
    if (live)
#pragma unroll 4
      for (int c = warp; c < l; c += NW) {
        const double xtc = xt[c * ldn + j];
        const double base = accel ? xadd(xmul(omt, xtc), xmul(r.theta, v[c * ldn + j])) : xtc;
        const double z = smooth ? base : xsub(base, xmul(r.t, __ldcg(&p.grad[c * ldn + j])));
        n1 = xadd(n1, xmul(z, z));
      }
    n1 = row_sum(n1);
    double shrink = 0.0, den;
    if (smooth) {
      if (r.algo == ALGO_SGD) {
        const double rn = __dsqrt_rn(n1);
        den = xadd(rn < r.thres ? 1.0 : 0.0, rn);
      } else {
        den = __dsqrt_rn(xadd(n1, d2));
      }
    } else {
      const double rn = __dsqrt_rn(n1);
      shrink = xsub(rn, tmu);
      shrink = shrink < 0.0 ? 0.0 : shrink;
      den = xadd(rn < r.thres ? 1.0 : 0.0, rn);
    }
    // pass 2: the new iterate, its thresholded copy, the extrapolation point and the scalar partials
    double n2 = 0.0;
    if (live)
#pragma unroll 4
      for (int c = warp; c < l; c += NW) {
        const double xtc = xt[c * ldn + j];
        const double base = accel ? xadd(xmul(omt, xtc), xmul(r.theta, v[c * ldn + j])) : xtc;
        const double gr = __ldcg(&p.grad[c * ldn + j]);
        double o, d, gsc;
        if (smooth) {
          gsc = xadd(gr, xmul(r.mu, xdiv(base, den)));
          o = xsub(base, xmul(r.t, gsc));
          d = xsub(o, base);
        } else {
          gsc = gr;
          o = xdiv(xmul(xsub(base, xmul(r.t, gr)), shrink), den);
          d = accel ? xsub(o, base) : xdiv(xsub(base, o), r.t);
        }
        const double oz = fabs(o) < r.thres ? 0.0 : o;
        xn[c * ldn + j] = o;
        xnt[c * ldn + j] = oz;
        if (xop != nullptr) { xop[c * ldn + j] = T(o); xop[(l + c) * ldn + j] = T(oz); }
        if (accel) vn[c * ldn + j] = xadd(xtc, xdiv(xsub(o, xtc), r.theta));
        sc.gd = xadd(sc.gd, xmul(gsc, d));
        sc.dd = xadd(sc.dd, xmul(d, d));
        n2 = xadd(n2, xmul(o, o));
        sc.mx = nmax(sc.mx, fabs(o));
      }
    n2 = row_sum(n2);
    if (warp == 0 && live) {
      sc.reg = xadd(sc.reg, __dsqrt_rn(n2));
      if (r.algo == ALGO_FGD) {
        sc.sx = xadd(sc.sx, xsub(__dsqrt_rn(xadd(n2, d2)), r.delta));
        sc.sy = xadd(sc.sy, xsub(__dsqrt_rn(xadd(n1, d2)), r.delta));
      }
    }
  }
  sc.gd = warp_sum(sc.gd);
  sc.dd = warp_sum(sc.dd);
  sc.reg = warp_sum(sc.reg);
  sc.sx = warp_sum(sc.sx);
  sc.sy = warp_sum(sc.sy);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sc.mx = nmax(sc.mx, __shfl_xor_sync(0xffffffffu, sc.mx, o));
  if (lane == 0) {
    s_red[warp][0] = sc.gd; s_red[warp][1] = sc.dd; s_red[warp][2] = sc.reg; s_red[warp][3] = sc.mx;
    s_red[warp][4] = sc.sx; s_red[warp][5] = sc.sy;
  }
  consumer_bar();
  if (tid == 0) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0, a4 = 0.0, a5 = 0.0;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      a0 += s_red[w][0]; a1 += s_red[w][1]; a2 += s_red[w][2]; a3 = nmax(a3, s_red[w][3]);
      a4 += s_red[w][4]; a5 += s_red[w][5];
    }
    double* out = p.scal_partial + SCAL_STRIDE * cb;
    out[0] = a0; out[1] = a1; out[2] = a2; out[3] = a3; out[4] = a4; out[5] = a5;
  }
  consumer_bar();
}

// Stand-alone launch: retries (ctrl->retry), the initial iterate (init) and the chunked generic-l path.
// only_if_retry = 1: return unless ctrl->retry is set (the scheduler enqueues it unconditionally).
template <typename T>
__global__ void __launch_bounds__(CONSUMER_THREADS)
glp_prox_kernel(const ProxArgs p, const Geom g, const int only_if_retry) {
  __shared__ double s_red[NUM_CONSUMER_WARPS][6];
  __shared__ double s_row[NUM_CONSUMER_WARPS][32];
  pdl_wait();
  pdl_launch();
  if (p.ctrl != nullptr) {
    if (p.ctrl->done) return;
    if (only_if_retry && !p.ctrl->retry) return;
  }
  const ProxResolved r = resolve(p);
  if (g.l >= 16) {
    for (int64_t cb = blockIdx.x; cb < g.ncb; cb += gridDim.x) prox_block_wide<T>(p, r, g, cb, threadIdx.x, s_red, s_row);
  } else {
    for (int64_t cb = blockIdx.x; cb < g.ncb; cb += gridDim.x) prox_block<T>(p, r, g, cb, threadIdx.x, s_red);
  }
  if (threadIdx.x < 32) prox_post_arrive(p.post, p.scal_partial, g.ncb, threadIdx.x);
}

}  // namespace glp
