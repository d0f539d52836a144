// glp_decide.cuh -- K3: the decide kernel.  Runs between the forward stream of one trial and the
// transposed stream of the next iteration.  Every CTA (a) reduces the per-block partial scalars of
// the two streams in a fixed order, (b) advances the control state machine (glp_decide.h) -- all
// CTAs compute the identical decision, CTA 0 publishes it -- and (c) performs its slice of the
// O((m+n) l) vector work that the decision implies:
//   Av   <- Axt_old + (Axn - Axt_old)/theta          (carried product of v, FProxGD :145 by linearity)
//   r    <- (1-theta')*Axnt + theta'*Av - b          (residual at the next base point, :139-140)
//   gy   <- sum(r^2)                                 (g(y) of the next sufficient-decrease test, :92)
//   v    <- x , Av <- Ax  at a continuation-stage start (:68)
#pragma once
#include "glp_decide.h"
#include "glp_device.cuh"
#include "glp_types.cuh"

namespace glp {

struct DecideArgs {
  const Ctrl* cin;
  Ctrl* cout;
  SolveConst sc;
  const double* rr_partial;   int32_t n_rr;
  const double* scal_partial; int32_t n_scal;   // [n_scal][SCAL_STRIDE]
  const double* cnt_partial;  int32_t n_cnt;
  double* gy_partial[2];      int32_t n_gy;     // [gridDim] each
  double* AX[2];              // [2l][ldm] : {A x | A xt} of the two state buffers
  double* Av;                 // [l][ldm]
  const double* b;            // [l][ldm]
  void* r;                    // [l][ldm] dtype T
  double* state[2];           // [3][l][ldn]
  double* f_hist;
  double* f_hist_best;
  double* s_hist;             // nullable: sparsity of every recorded iterate (the reference's sparsity_hist)
  unsigned* r_amax;           // nullable (fp16-pair wide mode): float bits of max|r|, zeroed before the launch
};

constexpr int DECIDE_THREADS = 256;

__device__ __forceinline__ double dec_nmax(double a, double b) {
  return (a != a) ? a : ((b != b) ? b : fmax(a, b));
}

// Shared-memory scratch of the decide phase (one instance per CTA).
struct DecideSmem {
  double part[DECIDE_THREADS / 32][8];
  Ctrl out;
  DecideOut act;
  double gy[DECIDE_THREADS / 32];
  unsigned rmx[DECIDE_THREADS / 32];
};

// The decide phase proper, executed by every CTA of a grid of DECIDE_THREADS-wide CTAs (the stand-alone decide
// kernel, or the second half of the fused gather + decide kernel of the multi-GPU path).
template <typename T>
__device__ __forceinline__ void decide_phase(const DecideArgs& a, const Geom& g, const Ctrl& cin, DecideSmem& sm) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // ---- (a) fixed-order reductions: strided per thread, shuffle tree per warp, warps in order
  double v_rr = 0.0, v_gd = 0.0, v_dd = 0.0, v_reg = 0.0, v_cnt = 0.0, v_gy = 0.0, v_sx = 0.0, v_sy = 0.0;
  for (int i = tid; i < a.n_rr; i += DECIDE_THREADS) v_rr += __ldcg(&a.rr_partial[i]);
  for (int i = tid; i < a.n_scal; i += DECIDE_THREADS) {
    v_gd += __ldcg(&a.scal_partial[SCAL_STRIDE * i + 0]);
    v_dd += __ldcg(&a.scal_partial[SCAL_STRIDE * i + 1]);
    v_reg += __ldcg(&a.scal_partial[SCAL_STRIDE * i + 2]);
    v_sx += __ldcg(&a.scal_partial[SCAL_STRIDE * i + 4]);
    v_sy += __ldcg(&a.scal_partial[SCAL_STRIDE * i + 5]);
  }
  for (int i = tid; i < a.n_cnt; i += DECIDE_THREADS) v_cnt += __ldcg(&a.cnt_partial[i]);
  const double* gy_in = a.gy_partial[cin.gy_sel];
  for (int i = tid; i < a.n_gy; i += DECIDE_THREADS) v_gy += __ldcg(&gy_in[i]);
  v_rr = warp_sum(v_rr); v_gd = warp_sum(v_gd); v_dd = warp_sum(v_dd);
  v_reg = warp_sum(v_reg); v_cnt = warp_sum(v_cnt); v_gy = warp_sum(v_gy);
  v_sx = warp_sum(v_sx); v_sy = warp_sum(v_sy);
  if (lane == 0) {
    sm.part[warp][0] = v_rr; sm.part[warp][1] = v_gd; sm.part[warp][2] = v_dd;
    sm.part[warp][3] = v_reg; sm.part[warp][4] = v_cnt; sm.part[warp][5] = v_gy;
    sm.part[warp][6] = v_sx; sm.part[warp][7] = v_sy;
  }
  __syncthreads();

  // ---- (b) state machine (thread 0 of every CTA; identical inputs => identical decision)
  if (tid == 0) {
    Reduced red = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    for (int w = 0; w < DECIDE_THREADS / 32; ++w) {
      red.rr += sm.part[w][0]; red.gd += sm.part[w][1]; red.dd += sm.part[w][2];
      red.reg += sm.part[w][3]; red.cnt += sm.part[w][4]; red.gbase += sm.part[w][5];
      red.sx += sm.part[w][6]; red.sy += sm.part[w][7];
    }
    const bool writer = blockIdx.x == 0;
    decide_step(a.sc, cin, red, double(g.n_total) * double(g.l), sm.out, sm.act,
                writer ? a.f_hist : nullptr, writer ? a.f_hist_best : nullptr, writer ? a.s_hist : nullptr);
    if (sm.act.action == GLP_ACT_ITERATE) sm.out.gy_sel = 1 - cin.gy_sel;
    if (writer) *a.cout = sm.out;
  }
  __syncthreads();
  if (sm.act.action != GLP_ACT_ITERATE) return;

  // ---- (c) vector work of the commit, sliced over the grid
  const int nc = sm.act.new_cur;
  const bool accel = a.sc.algo == ALGO_FPROXGD || a.sc.algo == ALGO_FGD;
  const double th_old = sm.act.theta_old, th = sm.out.theta, omt = xsub(1.0, th);
  const int64_t lenm = int64_t(g.l) * g.ldm, lenn = int64_t(g.l) * g.ldn;
  const double* Axn = a.AX[nc];
  const double* Axnt = a.AX[nc] + lenm;
  const double* Axt_old = a.AX[1 - nc] + lenm;
  T* r = static_cast<T*>(a.r);
  double gy = 0.0;
  unsigned rmx = 0u;   // max|r| as float bits: the scale of the next transposed product's fp16 operand
  const int64_t stride = int64_t(gridDim.x) * DECIDE_THREADS;
  for (int64_t i = int64_t(blockIdx.x) * DECIDE_THREADS + tid; i < lenm; i += stride) {
    double base;
    if (accel) {
      double av;
      if (sm.act.reset_v) {
        av = __ldcg(&Axn[i]);
      } else {
        const double o = __ldcg(&Axt_old[i]);
        av = xadd(o, xdiv(xsub(__ldcg(&Axn[i]), o), th_old));
      }
      a.Av[i] = av;
      base = xadd(xmul(omt, __ldcg(&Axnt[i])), xmul(th, av));
    } else {
      base = __ldcg(&Axnt[i]);
    }
    const double res = xsub(base, a.b[i]);
    r[i] = T(res);
    gy += res * res;
    const unsigned u = __float_as_uint(fabsf(float(T(res))));
    rmx = u > rmx ? u : rmx;
  }
  // max|r| of the CTA first, then ONE atomic per CTA: same-address L2 atomics serialise (~27 cycles each), and one per
  // warp (4.7 k of them at C5) was 80 of this kernel's 116 us (profiles/r02_launches_c5_l128.csv)
  rmx = __reduce_max_sync(0xffffffffu, rmx);
  if (a.r_amax != nullptr && lane == 0) sm.rmx[warp] = rmx;
  if (accel && sm.act.reset_v) {
    const double* x = a.state[nc];
    double* v = a.state[nc] + 2 * lenn;
    for (int64_t i = int64_t(blockIdx.x) * DECIDE_THREADS + tid; i < lenn; i += stride) v[i] = x[i];
  }
  gy = warp_sum(gy);
  if (lane == 0) sm.gy[warp] = gy;
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
    for (int w = 0; w < DECIDE_THREADS / 32; ++w) tot += sm.gy[w];
    a.gy_partial[sm.out.gy_sel][blockIdx.x] = tot;
    if (a.r_amax != nullptr) {
      unsigned mx = 0u;
      for (int w = 0; w < DECIDE_THREADS / 32; ++w) mx = sm.rmx[w] > mx ? sm.rmx[w] : mx;
      if (mx != 0u) atomicMax(a.r_amax, mx);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(DECIDE_THREADS)
glp_decide_kernel(const DecideArgs a, const Geom g) {
  __shared__ DecideSmem sm;
  pdl_wait();
  pdl_launch();
  const Ctrl cin = *a.cin;
  if (cin.done) {  // keep both control blocks identical once finished
    if (blockIdx.x == 0 && threadIdx.x == 0) *a.cout = cin;
    return;
  }
  decide_phase<T>(a, g, cin, sm);
}

}  // namespace glp
