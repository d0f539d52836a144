// glp_fwd.cuh -- K1: forward skinny product  AX[:, q] = A * X[:, q]  (q < RC, RC = 2*l for the solvers)
// with the residual epilogue  rr = sum((AX - b)^2)  fused into the last-arriving CTA of each row block.
//
// Replaces `A @ x` and `0.5*np.sum((A @ x - b)**2)` of the reference
// (gl_ProxGD_primal.py:25,61 ; gl_FProxGD_primal.py:25,63).
//
// Structure (one persistent CTA per SM, 10 warps):
//   warp 8  : TMA producer.  One elected lane streams [64 x 512 B] boxes of A (cp.async.bulk.tensor,
//             L2 evict-first) plus the matching 512 B slices of the RC operand columns into a
//             6-stage shared-memory ring guarded by full/empty mbarriers.
//   warps0-7: consumers.  Warp w owns rows 8w..8w+7 of the tile; each lane reads one 128-bit word
//             per row (conflict-free), keeps 8 x RC accumulators in registers for the whole run of
//             tiles that belong to one row block, and reduces across lanes only once per row block
//             with a transpose butterfly (8*RC-1 shuffles).
//   warp 9  : aux.  Evaluates the reference's sparsity count of the candidate iterate
//             (gl_*_primal.py:22) while A streams.
// The tile sequence (row-block major) is cut into gridDim equal contiguous ranges ("stream-K"), so
// every SM moves the same number of bytes; row blocks that straddle CTAs are completed by the last
// arriver, which sums the partials in CTA order (deterministic).
#pragma once
#include "glp_comm.cuh"
#include "glp_device.cuh"
#include "glp_partition.h"
#include "glp_types.cuh"

namespace glp {

template <int RC> constexpr int fwd_stage_stride() { return TILE_BYTES + RC * TILE_ROW_BYTES; }
template <int RC> constexpr int fwd_smem_bytes() {
  return NUM_STAGES * fwd_stage_stride<RC>() + 2 * NUM_STAGES * 8 + 1024 /*align slack*/;
}
constexpr int FWD_THREADS = CONSUMER_THREADS + 64;
// fp32 streams: the per-lane FMAs run in fp32 and are flushed every FWD_FLUSH_TILES tiles (8 products per lane and
// operand column) into a second level that, like the cross-lane reduction, is kept in fp64.  At C4 the entries of A x are
// ~114 while the residual entries are ~0.15: with an fp32 second level flushed every 8 tiles (round 1; compile with
// -DGLP_FWD_ACC64=0 -DGLP_FWD_FLUSH=8) the noise of ||A x - b||^2 exceeded the margin of ProxGD's sufficient-decrease
// test late in a stage, and the line search flipped coins: 385 forward streams for 150 iterations, against 151 here --
// the count of the fp64 run.  Same 168 registers, same stream rate (profiles/r02_ab_k1_acc64.log).
#ifndef GLP_FWD_ACC64
#define GLP_FWD_ACC64 1
#endif
#ifndef GLP_FWD_FLUSH
#define GLP_FWD_FLUSH 2
#endif
constexpr int FWD_FLUSH_TILES = GLP_FWD_FLUSH;
template <typename T> struct FwdHiAcc { using type = T; };
#if GLP_FWD_ACC64
template <> struct FwdHiAcc<float> { using type = double; };
#endif

__device__ __forceinline__ double nan_max(double a, double b) {
  return (a != a) ? a : ((b != b) ? b : fmax(a, b));
}

// Multi-GPU: called once per CTA by consumer warp 0 and (count launches) once by the aux warp, all 32
// lanes.  The last arrival of the launch publishes this rank's sparsity count and raises flag_ax on
// every peer; every slab store of the launch happens-before it (system fence + counter chain).
__device__ __forceinline__ void fwd_finish(const FwdArgs& a, int lane) {
  unsigned prev = 0;
  if (lane == 0) {
    __threadfence_system();
    prev = atomicAdd(a.fin, 1u);
  }
  prev = __shfl_sync(0xffffffffu, prev, 0);
  const unsigned target = gridDim.x * (a.count_mode ? 2u : 1u);
  if (prev + 1 != target) return;
  __threadfence();
  if (lane == 0) *a.fin = 0u;
  double cnt = 0.0;
  if (a.count_mode) {
    for (int i = lane; i < int(gridDim.x); i += 32) cnt += __ldcg(&a.cnt_partial[i]);
    cnt = warp_sum(cnt);
  }
  if (lane < a.comm.world) {
    if (a.count_mode) mail_sc(a.comm, lane, a.par, a.comm.rank)[4] = cnt;
    __threadfence_system();
    st_release_sys(mail_flag(a.comm, lane, MAIL_FLAG_AX, a.comm.rank), a.seq_f);
  }
}

template <typename T, int RC>
__global__ void __launch_bounds__(FWD_THREADS, 1)
glp_fwd_kernel(const __grid_constant__ CUtensorMap tmapA, const FwdArgs a, const Geom g,
               const int evict_first) {
  using E = Elem<T>;
  using Vec = typename E::Vec;
  constexpr int VEC = E::VEC;
  constexpr int TC = E::TILE_COLS;
  constexpr int STRIDE = fwd_stage_stride<RC>();
  constexpr uint32_t STAGE_TX = TILE_BYTES + RC * TILE_ROW_BYTES;
  constexpr int NV = ROWS_PER_WARP * RC;  // values per lane entering the butterfly
  using Hi = typename FwdHiAcc<T>::type;  // type of the blocked-summation accumulators and of the cross-lane reduction

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + NUM_STAGES * STRIDE);
  uint64_t* empty = full + NUM_STAGES;
  __shared__ int s_last;
  __shared__ double s_red[NUM_CONSUMER_WARPS];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t G = gridDim.x, c = blockIdx.x;
  const int64_t ncb = g.ncb;
  const int64_t N = int64_t(g.nrb) * ncb;
  const int64_t lo = tile_lo(c, G, N), hi = tile_lo(c + 1, G, N);

  if (tid == 0) {
    for (int s = 0; s < NUM_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], NUM_CONSUMER_WARPS);
    }
    fence_mbar_init();
    fence_proxy_async();
    prefetch_tensormap(&tmapA);
  }
  __syncthreads();
  // ---- everything above overlapped the predecessor's tail (programmatic dependent launch)
  pdl_wait();
  pdl_launch();
  if (a.ctrl != nullptr && a.ctrl->done) return;  // grid-uniform
  const int sel = a.ctrl != nullptr ? 1 - a.ctrl->cur : a.sel;

  if (warp == NUM_CONSUMER_WARPS) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && lo < hi) {
      const uint64_t pol = evict_first ? policy_evict_first() : policy_evict_normal();
      const T* X = static_cast<const T*>(a.X[sel]);
      int32_t rb = int32_t(lo / ncb), cb = int32_t(lo % ncb);
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t t = lo; t < hi; ++t) {
        mbar_wait(&empty[stage], phase ^ 1u, a.fault);
        mbar_arrive_expect_tx(&full[stage], STAGE_TX);
        uint8_t* s = smem + stage * STRIDE;
        tma_load_2d(s, &tmapA, &full[stage], cb * TC, rb * TILE_ROWS, pol);
#pragma unroll
        for (int q = 0; q < RC; ++q)
          bulk_load_1d(s + TILE_BYTES + q * TILE_ROW_BYTES, X + int64_t(q) * g.ldn + int64_t(cb) * TC,
                       TILE_ROW_BYTES, &full[stage]);
        if (++cb == ncb) { cb = 0; ++rb; }
        if (++stage == NUM_STAGES) { stage = 0; phase ^= 1u; }
      }
    }
    return;
  }
  if (warp == NUM_CONSUMER_WARPS + 1) {
    // ------------------------------------------------------------------ aux: sparsity count
    if (a.count_mode) {
      double mx = 0.0;
      if (a.comm.world > 1) {
        // global max|xn|: every rank's prox launch pushed its local maximum into this mailbox
        if (lane < a.comm.world) wait_flag(mail_flag(a.comm, a.comm.rank, MAIL_FLAG_SC, lane), a.seq_t, a.fault);
        __syncwarp();
        for (int r = 0; r < a.comm.world; ++r) mx = nan_max(mx, ld_mail(mail_sc(a.comm, a.comm.rank, a.par, r)));
      } else {
        for (int i = lane; i < g.ncb; i += 32) mx = nan_max(mx, __ldcg(&a.scal_partial[SCAL_STRIDE * i + 3]));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = nan_max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      }
      const double cut = xmul(1e-6, mx);
      const double* xc = a.xcand[sel];
      const int64_t j0 = tile_lo(c, G, g.n), j1 = tile_lo(c + 1, G, g.n);
      int cnt = 0;
      for (int col = 0; col < g.l; ++col)
        for (int64_t j = j0 + lane; j < j1; j += 32) cnt += fabs(__ldcg(&xc[col * g.ldn + j])) > cut ? 1 : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
      if (lane == 0) a.cnt_partial[c] = double(cnt);
      if (a.comm.world > 1) fwd_finish(a, lane);
    }
    return;
  }

  // -------------------------------------------------------------------- consumers
  T acc[ROWS_PER_WARP][RC];
  Hi acc_hi[ROWS_PER_WARP][RC];  // fp32: blocked summation (flush every FWD_FLUSH_TILES tiles); unused for fp64
  int stage = 0;
  uint32_t phase = 0;
  int64_t t = lo;
  while (t < hi) {
    const int64_t rb = t / ncb;
    const int64_t t_end = (rb + 1) * ncb < hi ? (rb + 1) * ncb : hi;
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
      for (int q = 0; q < RC; ++q) { acc[i][q] = T(0); acc_hi[i][q] = Hi(0); }
    int run = 0;
    for (; t < t_end; ++t) {
      mbar_wait(&full[stage], phase, a.fault);
      const uint8_t* s = smem + stage * STRIDE;
      Vec xv[RC];
#pragma unroll
      for (int q = 0; q < RC; ++q)
        xv[q] = *reinterpret_cast<const Vec*>(s + TILE_BYTES + q * TILE_ROW_BYTES + lane * 16);
#pragma unroll
      for (int i = 0; i < ROWS_PER_WARP; ++i) {
        const Vec av = *reinterpret_cast<const Vec*>(s + (warp * ROWS_PER_WARP + i) * TILE_ROW_BYTES + lane * 16);
#pragma unroll
        for (int q = 0; q < RC; ++q)
#pragma unroll
          for (int e = 0; e < VEC; ++e) acc[i][q] = fma(vget(av, e), vget(xv[q], e), acc[i][q]);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[stage]);
      if (++stage == NUM_STAGES) { stage = 0; phase ^= 1u; }
      if (sizeof(T) == 4 && ++run == FWD_FLUSH_TILES) {
        run = 0;
#pragma unroll
        for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
          for (int q = 0; q < RC; ++q) { acc_hi[i][q] += Hi(acc[i][q]); acc[i][q] = T(0); }
      }
    }
    // ---- one cross-lane reduction per row block
    Hi v[NV];
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
      for (int q = 0; q < RC; ++q) v[i * RC + q] = Hi(acc[i][q]) + acc_hi[i][q];
    warp_transpose_sum<Hi, NV>(v, lane);
    // lane j < NV holds value j = i*RC + q of row warp*8 + i
    double* slot = a.partial + (c + rb) * int64_t(TILE_ROWS * RC);
    if (lane < NV) slot[warp * NV + lane] = double(v[0]);
    __threadfence();
    consumer_bar();
    const int64_t first = contrib_first(rb, ncb, G, N), last = contrib_last(rb, ncb, G, N);
    if (tid == 0) {
      const unsigned prev = atomicAdd(&a.counter[rb], 1u);
      s_last = (prev == unsigned(last - first)) ? 1 : 0;
    }
    consumer_bar();
    if (s_last) {
      // ---- fused epilogue: complete the row block, r = AX - b, ||r||^2
      __threadfence();
      double sq = 0.0;
      if (tid < TILE_ROWS * RC) {
        const int q = tid / TILE_ROWS, row = tid % TILE_ROWS;
        double sum = 0.0;
        for (int64_t cc = first; cc <= last; ++cc)
          sum += __ldcg(&a.partial[((cc + rb) * TILE_ROWS + row) * RC + q]);
        const int64_t grow = rb * TILE_ROWS + row;
        if (a.comm.world > 1) {
          // push this rank's finished row block into everybody's slab (NVLink stores, 512 B runs)
          // The stores become visible to the peers through the ONE system fence of fwd_finish (lane 0 of warp 0,
          // after the consumer barrier below): no fence per row block.  Peers are visited starting at rank + 1 so
          // that the W ranks do not all hit the same NVLink port at the same time.
          if (grow < g.m) {
            const int64_t off = int64_t(a.col0 + q) * g.ldm + grow;
            int r = a.comm.rank;
            for (int k = 0; k < a.comm.world; ++k) {
              r = r + 1 == a.comm.world ? 0 : r + 1;
              mail_slab(a.comm, r, a.par, a.comm.rank)[off] = sum;
            }
          }
          if (a.push_fence_per_block) __threadfence_system();   // A/B switch (GLPROX_PUSH_FENCE=block)
        } else if (grow < g.m) {
          a.AX[sel][int64_t(q) * g.ldm + grow] = sum;
          if (q < a.nb_cols) {
            const double d = sum - a.b[int64_t(a.bcol0 + q) * g.ldm + grow];
            sq = d * d;
          }
        }
      }
      sq = warp_sum(sq);
      if (lane == 0) s_red[warp] = sq;
      consumer_bar();
      if (tid == 0) {
        double tot = 0.0;
#pragma unroll
        for (int w = 0; w < NUM_CONSUMER_WARPS; ++w) tot += s_red[w];
        a.rr_partial[rb] = tot;
        a.counter[rb] = 0u;
      }
    }
  }
  if (a.comm.world > 1 && warp == 0) fwd_finish(a, lane);
}

}  // namespace glp
