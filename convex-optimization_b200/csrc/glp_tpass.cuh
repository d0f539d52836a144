// glp_tpass.cuh -- K2: transposed stream  grad[:, c] = A^T * R[:, c]  (c < LC, LC = l for the solvers)
// whose last-arriving CTA per column block runs the fused prox epilogue (glp_prox.cuh).
//
// Replaces `A.T @ (A @ y - b)` (gl_ProxGD_primal.py:129 ; gl_FProxGD_primal.py:65-66) -- the residual
// R itself is produced by the decide kernel from carried products -- and, fused, prox_th / Gt /
// extrapolation / the line-search inner products.
//
// Same persistent TMA pipeline as the forward kernel (one CTA per SM, 6 x 32 KiB stages), but the
// tile sequence is column-block major: a CTA keeps VEC x LC register accumulators per lane for a
// 512-byte wide column block of A while it sweeps down the rows, so no cross-lane reduction is
// needed at all; the eight consumer warps are combined through shared memory once per column block.
#pragma once
#include "glp_device.cuh"
#include "glp_partition.h"
#include "glp_prox.cuh"
#include "glp_types.cuh"

namespace glp {

template <typename T, int LC> constexpr int tp_r_bytes() { return LC * TILE_ROWS * int(sizeof(T)); }
template <typename T, int LC> constexpr int tp_stage_stride() {
  return TILE_BYTES + ((tp_r_bytes<T, LC>() + 127) / 128) * 128;
}
template <typename T, int LC> constexpr int tp_smem_bytes() {
  return NUM_STAGES * tp_stage_stride<T, LC>() + 2 * NUM_STAGES * 8 +
         NUM_CONSUMER_WARPS * LC * TILE_ROW_BYTES + 1024;
}
constexpr int TP_THREADS = CONSUMER_THREADS + 32;

template <typename T, int LC>
__global__ void __launch_bounds__(TP_THREADS, 1)
glp_tpass_kernel(const __grid_constant__ CUtensorMap tmapA, const TpassArgs a, const ProxArgs p,
                 const Geom g, const int evict_first) {
  using E = Elem<T>;
  using Vec = typename E::Vec;
  constexpr int VEC = E::VEC;
  constexpr int TC = E::TILE_COLS;
  constexpr int STRIDE = tp_stage_stride<T, LC>();
  constexpr uint32_t R_SLICE = TILE_ROWS * sizeof(T);  // bytes of one residual column per tile
  constexpr uint32_t STAGE_TX = TILE_BYTES + LC * R_SLICE;
  constexpr int RV = ROWS_PER_WARP / VEC;  // Vec loads covering a warp's 8 residual rows

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + NUM_STAGES * STRIDE);
  uint64_t* empty = full + NUM_STAGES;
  T* red = reinterpret_cast<T*>(empty + NUM_STAGES);  // [8 warps][LC][TC]
  __shared__ int s_last;
  __shared__ double s_red[NUM_CONSUMER_WARPS][6];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

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

  if (a.ctrl != nullptr && a.ctrl->retry) {
    // A rejected line-search trial: the gradient is unchanged, only the epilogue is re-evaluated
    // with the shrunken step.  No stream of A.
    if (a.fuse && warp < NUM_CONSUMER_WARPS) {
      const ProxResolved r = resolve(p);
      for (int64_t cb = blockIdx.x; cb < g.ncb; cb += gridDim.x) prox_block<T>(p, r, g, cb, tid, s_red);
      if (warp == 0) prox_post_arrive(p.post, p.scal_partial, g.ncb, lane);
    }
    return;
  }

  const int64_t G = gridDim.x, c = blockIdx.x;
  const int64_t nrb = g.nrb;
  const int64_t N = nrb * int64_t(g.ncb);
  const int64_t lo = tile_lo(c, G, N), hi = tile_lo(c + 1, G, N);

  if (warp == NUM_CONSUMER_WARPS) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && lo < hi) {
      const uint64_t pol = evict_first ? policy_evict_first() : policy_evict_normal();
      const T* R = static_cast<const T*>(a.R);
      int32_t cb = int32_t(lo / nrb), rb = int32_t(lo % nrb);
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t t = lo; t < hi; ++t) {
        mbar_wait(&empty[stage], phase ^ 1u, a.fault);
        mbar_arrive_expect_tx(&full[stage], STAGE_TX);
        uint8_t* s = smem + stage * STRIDE;
        tma_load_2d(s, &tmapA, &full[stage], cb * TC, rb * TILE_ROWS, pol);
#pragma unroll
        for (int q = 0; q < LC; ++q)
          bulk_load_1d(s + TILE_BYTES + q * R_SLICE, R + int64_t(q) * g.ldm + int64_t(rb) * TILE_ROWS, R_SLICE,
                       &full[stage]);
        if (++rb == nrb) { rb = 0; ++cb; }
        if (++stage == NUM_STAGES) { stage = 0; phase ^= 1u; }
      }
    }
    return;
  }

  // -------------------------------------------------------------------- consumers
  const ProxResolved pr = resolve(p);
  T acc[LC][VEC];
  T acc_hi[LC][VEC];
  int stage = 0;
  uint32_t phase = 0;
  int64_t t = lo;
  while (t < hi) {
    const int64_t cb = t / nrb;
    const int64_t t_end = (cb + 1) * nrb < hi ? (cb + 1) * nrb : hi;
#pragma unroll
    for (int q = 0; q < LC; ++q)
#pragma unroll
      for (int e = 0; e < VEC; ++e) { acc[q][e] = T(0); acc_hi[q][e] = T(0); }
    int run = 0;
    for (; t < t_end; ++t) {
      mbar_wait(&full[stage], phase, a.fault);
      const uint8_t* s = smem + stage * STRIDE;
      Vec rv[LC][RV];
#pragma unroll
      for (int q = 0; q < LC; ++q)
#pragma unroll
        for (int u = 0; u < RV; ++u)
          rv[q][u] = *reinterpret_cast<const Vec*>(s + TILE_BYTES + q * R_SLICE +
                                                   (warp * ROWS_PER_WARP + u * VEC) * sizeof(T));
#pragma unroll
      for (int i = 0; i < ROWS_PER_WARP; ++i) {
        const Vec av = *reinterpret_cast<const Vec*>(s + (warp * ROWS_PER_WARP + i) * TILE_ROW_BYTES + lane * 16);
#pragma unroll
        for (int q = 0; q < LC; ++q) {
          const T rq = vget(rv[q][i / VEC], i % VEC);
#pragma unroll
          for (int e = 0; e < VEC; ++e) acc[q][e] = fma(vget(av, e), rq, acc[q][e]);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[stage]);
      if (++stage == NUM_STAGES) { stage = 0; phase ^= 1u; }
      if (sizeof(T) == 4 && ++run == 8) {
        run = 0;
#pragma unroll
        for (int q = 0; q < LC; ++q)
#pragma unroll
          for (int e = 0; e < VEC; ++e) { acc_hi[q][e] += acc[q][e]; acc[q][e] = T(0); }
      }
    }
    // ---- combine the 8 warps (fixed order), publish this CTA's partial for column block cb
#pragma unroll
    for (int q = 0; q < LC; ++q)
#pragma unroll
      for (int e = 0; e < VEC; ++e) red[(warp * LC + q) * TC + lane * VEC + e] = acc[q][e] + acc_hi[q][e];
    consumer_bar();
    double* slot = a.partial + (c + cb) * int64_t(LC * TC);
    for (int o = tid; o < LC * TC; o += CONSUMER_THREADS) {
      double sum = 0.0;
#pragma unroll
      for (int w = 0; w < NUM_CONSUMER_WARPS; ++w) sum += double(red[w * LC * TC + o]);
      slot[o] = sum;
    }
    __threadfence();
    consumer_bar();
    const int64_t first = contrib_first(cb, nrb, G, N), last = contrib_last(cb, nrb, G, N);
    if (tid == 0) {
      const unsigned prev = atomicAdd(&a.counter[cb], 1u);
      s_last = (prev == unsigned(last - first)) ? 1 : 0;
    }
    consumer_bar();
    if (s_last) {
      __threadfence();
      for (int o = tid; o < LC * TC; o += CONSUMER_THREADS) {
        const int q = o / TC, col = o % TC;
        double sum = 0.0;
        for (int64_t cc = first; cc <= last; ++cc) sum += __ldcg(&a.partial[(cc + cb) * int64_t(LC * TC) + o]);
        const int64_t j = cb * TC + col;
        if (j < g.n) a.grad[int64_t(a.gcol0 + q) * g.ldn + j] = sum;
      }
      if (tid == 0) a.counter[cb] = 0u;
      if (a.fuse) {
        __threadfence();
        consumer_bar();  // the block's gradient is complete and visible to the CTA
        prox_block<T>(p, pr, g, cb, tid, s_red);
      }
    }
  }
  if (a.fuse && warp == 0) prox_post_arrive(p.post, p.scal_partial, g.ncb, lane);
}

}  // namespace glp
