// glp_device.cuh -- sm_100a device-side building blocks shared by the streaming kernels:
// mbarrier / TMA (cp.async.bulk[.tensor]) wrappers, deterministic reductions, rounding-exact
// scalar helpers.  Hand-written PTX; no CUTLASS.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace glp {

// ------------------------------------------------------------------------------------------------
// Tile geometry.  A is streamed in [TILE_ROWS x 512-byte] boxes by TMA (32 KiB per stage); one warp
// reads a tile row with one conflict-free 128-bit shared load per lane.
// ------------------------------------------------------------------------------------------------
constexpr int TILE_ROWS = 64;
constexpr int TILE_ROW_BYTES = 512;
constexpr int TILE_BYTES = TILE_ROWS * TILE_ROW_BYTES;  // 32768
constexpr int NUM_CONSUMER_WARPS = 8;
constexpr int ROWS_PER_WARP = TILE_ROWS / NUM_CONSUMER_WARPS;  // 8
constexpr int NUM_STAGES = 6;
constexpr int CONSUMER_THREADS = NUM_CONSUMER_WARPS * 32;  // 256

template <typename T> struct Elem;
template <> struct Elem<float> {
  using Vec = float4;
  static constexpr int VEC = 4;
  static constexpr int TILE_COLS = TILE_ROW_BYTES / 4;  // 128
};
template <> struct Elem<double> {
  using Vec = double2;
  static constexpr int VEC = 2;
  static constexpr int TILE_COLS = TILE_ROW_BYTES / 8;  // 64
};

__device__ __forceinline__ float vget(const float4& v, int i) {
  return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w));
}
__device__ __forceinline__ double vget(const double2& v, int i) { return i == 0 ? v.x : v.y; }

// ------------------------------------------------------------------------------------------------
// Rounding-exact scalar arithmetic (no FMA contraction) for the places where the order of
// operations of the reference NumPy expressions is kept.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double xmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double xadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double xsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double xdiv(double a, double b) { return __ddiv_rn(a, b); }

// ------------------------------------------------------------------------------------------------
// mbarrier + TMA
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug must not hang the GPU.  On timeout the fault word is set and the
// CTA traps; the host reports GLP_ERR_DEVICE_FAULT.
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int* fault) {
  if (mbar_try_wait(bar, parity)) return;
  const uint64_t t0 = global_timer_ns();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0x3ffu) == 0 && global_timer_ns() - t0 > 4000000000ull) {  // 4 s
      if (fault) atomicExch(fault, 1);
      __threadfence_system();
      __trap();
    }
  }
}
// 2D tiled TMA load: box (c0 = innermost/column coordinate, c1 = row coordinate).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar,
                                            int32_t c0, int32_t c1, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
// The same load multicast to every CTA of the cluster whose bit is set in cta_mask: the box lands at the same
// shared-memory offset in each of them and completes bytes on the mbarrier at the same offset in each.
__device__ __forceinline__ void tma_load_2d_mc(void* smem_dst, const CUtensorMap* map, uint64_t* bar,
                                               int32_t c0, int32_t c1, uint16_t cta_mask, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster.L2::cache_hint "
      "[%0], [%1, {%3, %4}], [%2], %5, %6;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(cta_mask), "l"(policy)
      : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// 1D bulk copy global -> shared (bytes multiple of 16, both addresses 16-byte aligned).
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_normal() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void prefetch_tensormap(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
// Programmatic dependent launch (griddepcontrol): a kernel launched with the programmatic-stream-serialization
// attribute may become resident while its predecessor in the stream is still running; everything before pdl_wait()
// (shared-memory carve-up, mbarrier init, tensor-map prefetch) overlaps the predecessor's tail.  pdl_wait() returns
// once the predecessor grid has completed and its memory is visible; NO global memory written by an earlier kernel of
// the stream may be read before it.  pdl_launch() lets the successor's CTAs be scheduled as this kernel's CTAs retire.
// Both are no-ops for a kernel launched without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Named barrier among the consumer warps only (the producer / aux warps keep running).
__device__ __forceinline__ void consumer_bar() {
  asm volatile("bar.sync 1, %0;" ::"n"(CONSUMER_THREADS) : "memory");
}

// ------------------------------------------------------------------------------------------------
// Deterministic reductions (fixed tree => identical bits on every CTA / rank / run)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// "Transpose" butterfly: every lane holds N values (N = 32, 16 or 8); on return lane j (j < N)
// holds in v[0] the sum over all 32 lanes of value j.  N-1 shuffles instead of 5N.
template <typename A, int N>
__device__ __forceinline__ void warp_transpose_sum(A (&v)[N], int lane) {
  static_assert(N == 32 || N == 16 || N == 8, "N");
  constexpr int TOP = N / 2;  // first exchange distance in lanes == number of values kept
#pragma unroll
  for (int cnt = TOP; cnt >= 1; cnt >>= 1) {
    const bool upper = (lane & cnt) != 0;
#pragma unroll
    for (int i = 0; i < cnt; ++i) {
      A send = upper ? v[i] : v[i + cnt];
      A keep = upper ? v[i + cnt] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, cnt);
    }
  }
  // lanes >= N hold copies of lane (lane % N)'s partial over a subset; finish over the
  // remaining lane bits that were never exchanged.
#pragma unroll
  for (int o = 16; o >= N; o >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], o);
}

}  // namespace glp
