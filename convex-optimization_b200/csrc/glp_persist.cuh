// glp_persist.cuh -- K4: the whole solve as ONE persistent kernel for problems whose A fits in the
// shared memory of one thread-block cluster (the homework fixture: 256 x 512 fp64 = 1 MiB over 8 CTAs).
//
// Replaces the bodies of gl_ProxGD_primal / gl_FProxGD_primal (gl_ProxGD_primal.py:49-146 ;
// gl_FProxGD_primal.py:50-161) for small problems, where the streaming path is launch-bound.
//
// CTA c of the cluster keeps the column block A[:, J_c] resident in its shared memory for the whole
// solve (the same column-group sharding the multi-GPU path uses, at cluster scale):
//   transposed pass   grad[J_c] = A[:, J_c]^T r          local (r is replicated in every CTA)
//   prox epilogue     xn, xnt, vn on the row groups J_c  local (glp_prox.cuh, reference operation order)
//   forward pass      P_c = A[:, J_c] [xn | xnt][J_c]    local partial product, m x 2l
//   exchange          every CTA PULLS the C partials through distributed shared memory and sums them
//                     in rank order -> identical A x, ||r||^2 and scalars in every CTA, so each CTA runs
//                     the control state machine (glp_decide.h) redundantly and reaches the same decision
// Two cluster barriers per trial (scalars + max|xn| ; partial products + count).  No global memory
// traffic inside the loop except f_hist; no launch, no host round trip.
#pragma once
#include <cooperative_groups.h>

#include "glp_decide.h"
#include "glp_device.cuh"
#include "glp_prox.cuh"
#include "glp_types.cuh"

namespace glp {
namespace cg = cooperative_groups;

constexpr int PERSIST_THREADS = 512;
constexpr int PERSIST_WARPS = PERSIST_THREADS / 32;
constexpr int PERSIST_MAX_L = 4;
constexpr int PERSIST_MAX_NC = 128;  // columns of A per CTA

struct PersistArgs {
  const void* A;          // global, row-major [m][lda], dtype T
  int64_t lda;
  const double* b;        // planar fp64 [l][ldm]
  const double* x0;       // planar fp64 [l][ldn]
  double* x_out;          // planar fp64 [l][ldn]
  Ctrl* ctrl_out;
  double* f_hist;
  double* f_hist_best;
  double* s_hist;
  SolveConst sc;
  int32_t m, n, l, nc;    // nc = columns per CTA (multiple of 2)
  int64_t ldm, ldn;
};

// Work split of the two local passes over the resident A slice (host + device).
//   forward : `fparts` threads share a row, each walks a contiguous range of the nc columns
//   transposed: warp -> (column chunk of 32, row group); `tgroups` row groups per chunk
__host__ __device__ inline int persist_mround(int m) { return (m + 31) & ~31; }
__host__ __device__ inline int persist_fparts(int m) {
  const int p = PERSIST_THREADS / persist_mround(m);
  return p < 1 ? 1 : p;
}
__host__ __device__ inline int persist_tgroups(int nc) {
  const int chunks = (nc + 31) / 32;
  return PERSIST_WARPS / chunks;
}

// Shared-memory carve-up, identical on host (capacity check) and device.  Offsets in bytes.
struct PersistLayout {
  int ncp;                // padded row stride of the A slice in elements (odd -> conflict-free row walk)
  size_t o_A, o_b, o_r, o_Av, o_AX, o_P, o_x, o_grad, o_xop, o_scr, o_slots, o_misc, total;
};
__host__ __device__ inline size_t persist_up(size_t x) { return (x + 15) & ~size_t(15); }
__host__ __device__ inline PersistLayout persist_layout(int m, int nc, int l, int esz) {
  PersistLayout L;
  L.ncp = nc | 1;
  size_t o = 0;
  L.o_A = o;     o = persist_up(o + size_t(m) * L.ncp * esz);
  L.o_b = o;     o = persist_up(o + size_t(l) * m * 8);
  L.o_r = o;     o = persist_up(o + size_t(l) * m * 8);
  L.o_Av = o;    o = persist_up(o + size_t(l) * m * 8);
  L.o_AX = o;    o = persist_up(o + size_t(2) * 2 * l * m * 8);        // [2][2l][m]
  L.o_P = o;     o = persist_up(o + size_t(2) * l * m * 8);            // [2l][m], pulled by the peers
  L.o_x = o;     o = persist_up(o + size_t(2) * 3 * l * nc * 8);       // [2][{x,xt,v}][l][nc]
  L.o_grad = o;  o = persist_up(o + size_t(l) * nc * 8);
  L.o_xop = o;   o = persist_up(o + size_t(nc) * 2 * l * 8);           // [nc][2l] interleaved operand
  const size_t fscr = size_t(persist_fparts(m)) * 2 * l * m * 8;       // forward partials [part][2l][m]
  const size_t tscr = size_t(persist_tgroups(nc)) * l * nc * 8;        // transposed partials [group][l][nc]
  L.o_scr = o;   o = persist_up(o + (fscr > tscr ? fscr : tscr));      // CTA-private, the two never overlap in time
  L.o_slots = o; o = persist_up(o + size_t(2) * 8 * 8);                // [2][8] scalars, pulled by the peers
  L.o_misc = o;  o = persist_up(o + 2048);
  L.total = o;
  return L;
}

// max of two non-negative doubles (or NaN) on the integer pipe: for sign-cleared IEEE doubles the bit
// patterns order like the values and every NaN is above +inf, so NaN propagates like in np.max.
__device__ __forceinline__ double pmax_bits(double a, double b) {
  const unsigned long long x = (unsigned long long)__double_as_longlong(a), y = (unsigned long long)__double_as_longlong(b);
  return __longlong_as_double((long long)(x > y ? x : y));
}
__device__ __forceinline__ double warp_pmax(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = pmax_bits(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// Fixed-shape tree over the first 16 lanes (one value per warp of the CTA).
__device__ __forceinline__ double tree16_sum(double v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <typename T, int L>
__global__ void __launch_bounds__(PERSIST_THREADS, 1)
glp_persist_kernel(const PersistArgs a) {
  cg::cluster_group cluster = cg::this_cluster();
  const int C = int(cluster.num_blocks());
  const int rank = int(cluster.block_rank());
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m = a.m, nc = a.nc;
  const int j0 = rank * nc;
  const int nloc = max(0, min(a.n - j0, nc));
  const bool accel = a.sc.algo == ALGO_FPROXGD || a.sc.algo == ALGO_FGD;

  extern __shared__ __align__(16) uint8_t smem[];
  const PersistLayout lay = persist_layout(m, nc, L, int(sizeof(T)));
  const int ncp = lay.ncp;
  T* As = reinterpret_cast<T*>(smem + lay.o_A);
  double* bs = reinterpret_cast<double*>(smem + lay.o_b);
  double* rs = reinterpret_cast<double*>(smem + lay.o_r);
  double* Avs = reinterpret_cast<double*>(smem + lay.o_Av);
  double* AXs = reinterpret_cast<double*>(smem + lay.o_AX);   // [(buf*2L + q)*m + i]
  double* Ps = reinterpret_cast<double*>(smem + lay.o_P);     // [q*m + i]
  double* xs = reinterpret_cast<double*>(smem + lay.o_x);     // [((buf*3 + which)*L + c)*nc + jl]
  double* gs = reinterpret_cast<double*>(smem + lay.o_grad);  // [c*nc + jl]
  double* xop = reinterpret_cast<double*>(smem + lay.o_xop);  // [jl*2L + q]
  double* scr = reinterpret_cast<double*>(smem + lay.o_scr);
  double* slots = reinterpret_cast<double*>(smem + lay.o_slots);  // [par*8 + {mx,gd,dd,reg,cnt}]
  double* s_w = reinterpret_cast<double*>(smem + lay.o_misc);     // [WARPS][6] warp partials
  int* s_cnt = reinterpret_cast<int*>(smem + lay.o_misc + 768);   // [WARPS]
  double* s_one = reinterpret_cast<double*>(smem + lay.o_misc + 896);  // [WARPS] single-value partials
  Ctrl* s_ctrl = reinterpret_cast<Ctrl*>(smem + lay.o_misc + 1024);
  Ctrl* s_new = s_ctrl + 1;
  DecideOut* s_act = reinterpret_cast<DecideOut*>(s_new + 1);
  static_assert(PERSIST_WARPS * 48 <= 768 && PERSIST_WARPS * 8 <= 128 && 1024 + 2 * sizeof(Ctrl) + sizeof(DecideOut) <= 2048, "misc region");

  // ---- load the resident data: A[:, J_rank] (zero padded), b, x0[J_rank]
  const T* Ag = static_cast<const T*>(a.A);
  for (int idx = tid; idx < m * nc; idx += PERSIST_THREADS) {
    const int i = idx / nc, jl = idx - i * nc;
    As[i * ncp + jl] = jl < nloc ? Ag[int64_t(i) * a.lda + j0 + jl] : T(0);
  }
  for (int idx = tid; idx < L * m; idx += PERSIST_THREADS) {
    const int c = idx / m, i = idx - c * m;
    bs[idx] = a.b[int64_t(c) * a.ldm + i];
    rs[idx] = 0.0; Avs[idx] = 0.0;
  }
  for (int idx = tid; idx < 2 * 2 * L * m; idx += PERSIST_THREADS) AXs[idx] = 0.0;
  for (int idx = tid; idx < 2 * 3 * L * nc; idx += PERSIST_THREADS) xs[idx] = 0.0;
  for (int idx = tid; idx < L * nc; idx += PERSIST_THREADS) gs[idx] = 0.0;
  __syncthreads();
  for (int idx = tid; idx < L * nc; idx += PERSIST_THREADS) {  // candidate slot of cur = 0 is buffer 1
    const int c = idx / nc, jl = idx - c * nc;
    xs[(3 * L + c) * nc + jl] = jl < nloc ? a.x0[int64_t(c) * a.ldn + j0 + jl] : 0.0;
  }
  if (tid == 0) {
    ctrl_init(*s_ctrl, a.sc);
    s_ctrl->n_fwd = 1;
  }
  __syncthreads();

  ProxArgs p{};
  p.state[0] = xs;
  p.state[1] = xs + 3 * L * nc;
  p.grad = gs;
  p.xop = nullptr;
  Geom g{};
  g.l = L;
  g.ldn = nc;

  // work split of the local passes
  const int mr = persist_mround(m);
  const int fparts = persist_fparts(m);
  const int f_part = tid / mr, f_row = tid - f_part * mr;           // forward: (column range, row)
  const int f_j0 = (f_part * nloc) / fparts, f_j1 = ((f_part + 1) * nloc) / fparts;
  const int tchunks = (nc + 31) / 32, tgroups = persist_tgroups(nc);
  const int t_chunk = warp % tchunks, t_group = warp / tchunks;     // transposed: (32 columns, row group)

  double r_gy = 0.0;  // thread 0: sum(r^2) at the base point of the running iteration
  int trial_no = 0;
  const int64_t max_slots = int64_t(3) * a.sc.maxit * (int64_t(a.sc.ls_cap) + 1) + 8;
  int init = 1;

  for (int64_t slot = 0; slot <= max_slots; ++slot) {
    // ================= one trial: prox epilogue -> exchange -> forward partial -> exchange =========
    const int cur = s_ctrl->cur;
    const int par = trial_no & 1;
    ++trial_no;
    {
      ProxResolved pr;
      pr.algo = a.sc.algo; pr.cur = cur; pr.init = init;
      pr.theta = s_ctrl->theta; pr.t = s_ctrl->t; pr.mu = s_ctrl->mu; pr.thres = a.sc.thres; pr.delta = a.sc.delta;
      ProxScal ps = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
      for (int jl = tid; jl < nloc; jl += PERSIST_THREADS) prox_group<double, true>(p, pr, g, jl, ps);
      // four interleaved warp reductions (independent chains hide the fp64 latency)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        ps.gd += __shfl_xor_sync(0xffffffffu, ps.gd, o);
        ps.dd += __shfl_xor_sync(0xffffffffu, ps.dd, o);
        ps.reg += __shfl_xor_sync(0xffffffffu, ps.reg, o);
        ps.mx = pmax_bits(ps.mx, __shfl_xor_sync(0xffffffffu, ps.mx, o));
        ps.sx += __shfl_xor_sync(0xffffffffu, ps.sx, o);
        ps.sy += __shfl_xor_sync(0xffffffffu, ps.sy, o);
      }
      if (lane == 0) {
        s_w[warp * 6 + 0] = ps.gd; s_w[warp * 6 + 1] = ps.dd; s_w[warp * 6 + 2] = ps.reg; s_w[warp * 6 + 3] = ps.mx;
        s_w[warp * 6 + 4] = ps.sx; s_w[warp * 6 + 5] = ps.sy;
      }
    }
    const double* cand = xs + (1 - cur) * 3 * L * nc;  // {xn | xnt | vn}
    __syncthreads();  // #1 candidates + warp partials visible
    if (warp == 0) {
      double v0 = 0.0, v1 = 0.0, v2 = 0.0, v3 = 0.0, v4 = 0.0, v5 = 0.0;
      if (lane < PERSIST_WARPS) {
        v0 = s_w[lane * 6 + 0]; v1 = s_w[lane * 6 + 1]; v2 = s_w[lane * 6 + 2]; v3 = s_w[lane * 6 + 3];
        v4 = s_w[lane * 6 + 4]; v5 = s_w[lane * 6 + 5];
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) {
        v0 += __shfl_xor_sync(0xffffffffu, v0, o);
        v1 += __shfl_xor_sync(0xffffffffu, v1, o);
        v2 += __shfl_xor_sync(0xffffffffu, v2, o);
        v3 = pmax_bits(v3, __shfl_xor_sync(0xffffffffu, v3, o));
        v4 += __shfl_xor_sync(0xffffffffu, v4, o);
        v5 += __shfl_xor_sync(0xffffffffu, v5, o);
      }
      if (lane == 0) {
        slots[par * 8 + 0] = v3; slots[par * 8 + 1] = v0; slots[par * 8 + 2] = v1; slots[par * 8 + 3] = v2;
        slots[par * 8 + 5] = v4; slots[par * 8 + 6] = v5;
      }
    }
    for (int idx = tid; idx < nc * 2 * L; idx += PERSIST_THREADS) {
      const int jl = idx / (2 * L), q = idx - jl * 2 * L;
      xop[idx] = cand[q * nc + jl];  // q < L: xn column q ; q >= L: xnt column q-L (planes are adjacent)
    }
    cluster.sync();  // (A) every CTA's scalars are visible; xop visible inside the CTA

    // ---- global max|xn| (integer-pipe max over the C posted values), local sparsity count
    double gmx = 0.0;
    if (lane < C) gmx = cluster.map_shared_rank(slots, lane)[par * 8 + 0];
    gmx = warp_pmax(gmx);
    const double cut = xmul(1e-6, gmx);
    int cnt = 0;
    for (int idx = tid; idx < L * nc; idx += PERSIST_THREADS) {
      const int jl = idx % nc;
      cnt += (jl < nloc && fabs(cand[idx]) > cut) ? 1 : 0;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) s_cnt[warp] = cnt;

    // ---- forward partial product: `fparts` threads per row, two independent chains per operand column
    if (f_part < fparts) {
      for (int i = f_row; i < m; i += PERSIST_THREADS) {
        double acc0[2 * L], acc1[2 * L];
#pragma unroll
        for (int q = 0; q < 2 * L; ++q) { acc0[q] = 0.0; acc1[q] = 0.0; }
        const T* row = As + i * ncp;
        int jl = f_j0;
        for (; jl + 1 < f_j1; jl += 2) {
          const double a0 = double(row[jl]), a1 = double(row[jl + 1]);
#pragma unroll
          for (int q = 0; q < 2 * L; ++q) {
            acc0[q] = fma(a0, xop[jl * 2 * L + q], acc0[q]);
            acc1[q] = fma(a1, xop[(jl + 1) * 2 * L + q], acc1[q]);
          }
        }
        if (jl < f_j1) {
          const double a0 = double(row[jl]);
#pragma unroll
          for (int q = 0; q < 2 * L; ++q) acc0[q] = fma(a0, xop[jl * 2 * L + q], acc0[q]);
        }
#pragma unroll
        for (int q = 0; q < 2 * L; ++q) scr[(f_part * 2 * L + q) * m + i] = acc0[q] + acc1[q];
      }
    }
    __syncthreads();  // #2
    for (int e = tid; e < 2 * L * m; e += PERSIST_THREADS) {
      double sum = 0.0;
      for (int pt = 0; pt < fparts; ++pt) sum += scr[pt * 2 * L * m + e];
      Ps[e] = sum;
    }
    if (warp == 0) {
      int c16 = lane < PERSIST_WARPS ? s_cnt[lane] : 0;
      c16 = __reduce_add_sync(0xffffffffu, c16);
      if (lane == 0) slots[par * 8 + 4] = double(c16);
    }
    cluster.sync();  // (B) every CTA's partial product and count are visible

    // ---- gather: pull the C partials through DSMEM (independent loads), sum in rank order
    double* AXc = AXs + (1 - cur) * 2 * L * m;
    double rr = 0.0;
    for (int e = tid; e < 2 * L * m; e += PERSIST_THREADS) {
      double v[16];
#pragma unroll
      for (int c = 0; c < 16; ++c) v[c] = c < C ? cluster.map_shared_rank(Ps, c)[e] : 0.0;
      double sum = 0.0;
#pragma unroll
      for (int c = 0; c < 16; ++c) sum += v[c];  // + 0.0 beyond C: exact
      AXc[e] = sum;
      if (e < L * m) {
        const double d = sum - bs[e];
        rr = fma(d, d, rr);
      }
    }
    rr = warp_sum(rr);
    if (lane == 0) s_one[warp] = rr;
    init = 0;
    __syncthreads();  // #3

    // ================= decide (every CTA, identical inputs -> identical decision) ===================
    if (warp == 0) {
      const double rr_tot = tree16_sum(lane < PERSIST_WARPS ? s_one[lane] : 0.0);
      double v1 = 0.0, v2 = 0.0, v3 = 0.0, v4 = 0.0, v5 = 0.0, v6 = 0.0;
      if (lane < C) {
        const double* rs_ = cluster.map_shared_rank(slots, lane) + par * 8;
        v1 = rs_[1]; v2 = rs_[2]; v3 = rs_[3]; v4 = rs_[4]; v5 = rs_[5]; v6 = rs_[6];
      }
      double gd = 0.0, dd = 0.0, reg = 0.0, cn = 0.0, sx = 0.0, sy = 0.0;
      for (int c = 0; c < C; ++c) {
        gd += __shfl_sync(0xffffffffu, v1, c); dd += __shfl_sync(0xffffffffu, v2, c);
        reg += __shfl_sync(0xffffffffu, v3, c); cn += __shfl_sync(0xffffffffu, v4, c);
        sx += __shfl_sync(0xffffffffu, v5, c); sy += __shfl_sync(0xffffffffu, v6, c);
      }
      if (lane == 0) {
        Reduced red = {rr_tot, gd, dd, reg, cn, r_gy, sx, sy};
        decide_step(a.sc, *s_ctrl, red, double(a.n) * double(L), *s_new, *s_act, rank == 0 ? a.f_hist : nullptr,
                    rank == 0 ? a.f_hist_best : nullptr, rank == 0 ? a.s_hist : nullptr);
        *s_ctrl = *s_new;
      }
    }
    __syncthreads();  // #4
    const int action = s_act->action;
    if (action != GLP_ACT_ITERATE && action != GLP_ACT_RETRY) break;
    if (action == GLP_ACT_ITERATE) {
      // ---- commit: A v, residual at the next base point, g(base)   (glp_decide.cuh part (c))
      const int ncur = s_act->new_cur;
      const double th_old = s_act->theta_old, th = s_ctrl->theta, omt = xsub(1.0, th);
      const double* Axn = AXs + ncur * 2 * L * m;
      const double* Axnt = Axn + L * m;
      const double* Axt_old = AXs + (1 - ncur) * 2 * L * m + L * m;
      double gy = 0.0;
      for (int i = tid; i < L * m; i += PERSIST_THREADS) {
        double base;
        if (accel) {
          double av;
          if (s_act->reset_v) {
            av = Axn[i];
          } else {
            const double o = Axt_old[i];
            av = xadd(o, xdiv(xsub(Axn[i], o), th_old));
          }
          Avs[i] = av;
          base = xadd(xmul(omt, Axnt[i]), xmul(th, av));
        } else {
          base = Axnt[i];
        }
        const double res = xsub(base, bs[i]);
        rs[i] = res;
        gy = fma(res, res, gy);
      }
      if (accel && s_act->reset_v) {
        double* st = xs + ncur * 3 * L * nc;
        for (int idx = tid; idx < L * nc; idx += PERSIST_THREADS) st[2 * L * nc + idx] = st[idx];
      }
      gy = warp_sum(gy);
      if (lane == 0) s_one[warp] = gy;
      __syncthreads();  // #5 residual visible
      if (warp == 0) r_gy = tree16_sum(lane < PERSIST_WARPS ? s_one[lane] : 0.0);
      // ---- transposed pass: grad[J] = A[:, J]^T r ; warp = (32 columns, row group), 4 independent chains
      if (t_group < tgroups) {
        const int jl = t_chunk * 32 + lane;
        const bool live = jl < nc;
        double acc[4][L];
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int c = 0; c < L; ++c) acc[u][c] = 0.0;
        int i = t_group;
        for (; i + 3 * tgroups < m; i += 4 * tgroups) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int ir = i + u * tgroups;
            const double av = live ? double(As[ir * ncp + jl]) : 0.0;
#pragma unroll
            for (int c = 0; c < L; ++c) acc[u][c] = fma(av, rs[c * m + ir], acc[u][c]);
          }
        }
        for (; i < m; i += tgroups) {
          const double av = live ? double(As[i * ncp + jl]) : 0.0;
#pragma unroll
          for (int c = 0; c < L; ++c) acc[0][c] = fma(av, rs[c * m + i], acc[0][c]);
        }
        if (live) {
#pragma unroll
          for (int c = 0; c < L; ++c) scr[(t_group * L + c) * nc + jl] = (acc[0][c] + acc[1][c]) + (acc[2][c] + acc[3][c]);
        }
      }
      __syncthreads();  // #6
      for (int idx = tid; idx < L * nc; idx += PERSIST_THREADS) {
        double sum = 0.0;
        for (int gq = 0; gq < tgroups; ++gq) sum += scr[gq * L * nc + idx];
        gs[idx] = sum;
      }
      __syncthreads();  // #7
    }
  }

  // ---- publish: x (planar fp64) and the final control block; keep shared memory alive for the peers
  {
    const double* xf = xs + s_ctrl->cur * 3 * L * nc;
    for (int idx = tid; idx < L * nc; idx += PERSIST_THREADS) {
      const int c = idx / nc, jl = idx - c * nc;
      if (jl < nloc) a.x_out[int64_t(c) * a.ldn + j0 + jl] = xf[idx];
    }
    if (rank == 0 && tid == 0) {
      Ctrl cf = *s_ctrl;
      cf.cur = 0;  // x_out is state buffer 0 of the handle
      *a.ctrl_out = cf;
    }
  }
  cluster.sync();
}

}  // namespace glp
