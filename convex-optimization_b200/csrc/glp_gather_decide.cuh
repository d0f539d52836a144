// glp_gather_decide.cuh -- multi-GPU path: the gather of one forward group and the decide step that consumes it as
// ONE launch.  Round 1 ran them as two dependent launches per iteration (glp_gather_kernel, then glp_decide_kernel:
// the gather ends in exactly the scalars decide starts from), i.e. two launch gaps and two grid ramps on a 0.7 ms
// iteration at 8 GPUs.  Here every CTA (1) waits for the W peer flags in its own HBM and sums its slice of the W slabs
// in rank order (gather_phase, glp_comm.cuh), (2) crosses a grid-wide barrier -- all CTAs are co-resident: the host
// clamps the grid to the occupancy of the kernel -- and (3) runs the decide phase (decide_phase, glp_decide.cuh):
// fixed-order reduction of the per-CTA partials, the state machine, and its slice of the commit's vector work.
// Results are bit-identical to the two-launch schedule (same partial layout, same reduction order).
#pragma once
#include "glp_comm.cuh"
#include "glp_decide.cuh"

namespace glp {

static_assert(GATHER_THREADS == DECIDE_THREADS, "the fused kernel uses one CTA shape for both phases");

struct GridBarrier {
  unsigned* count;               // arrivals of the current launch; reset by the last arriver
  unsigned long long* flag;      // generation of the last completed barrier (monotonic over the handle's life)
  unsigned long long gen;        // this launch
};

__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// All threads of all CTAs call this once per launch.  Writes made before it by any CTA are visible after it to
// loads that bypass L1 (__ldcg / ld_mail) in every CTA.
__device__ __forceinline__ void grid_barrier(const GridBarrier& b, int* fault) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned prev = atomicAdd(b.count, 1u);
    if (prev + 1 == gridDim.x) {
      *b.count = 0u;
      __threadfence();
      st_release_gpu(b.flag, b.gen);
    } else if (ld_acquire_gpu(b.flag) < b.gen) {
      const uint64_t t0 = global_timer_ns();
      uint32_t spins = 0;
      while (ld_acquire_gpu(b.flag) < b.gen) {
        __nanosleep(32);
        if ((++spins & 0xfffu) == 0 && global_timer_ns() - t0 > 4000000000ull) {   // 4 s: never hang the GPU
          if (fault) atomicExch(fault, 1);
          __threadfence_system();
          __trap();
        }
      }
    }
  }
  __syncthreads();
}

template <typename T>
__global__ void __launch_bounds__(DECIDE_THREADS)
glp_gather_decide_kernel(const GatherArgs ga, const DecideArgs da, const Geom g, const GridBarrier bar) {
  __shared__ DecideSmem sm;
  __shared__ double s_red[GATHER_THREADS / 32];
  pdl_wait();
  pdl_launch();
  const Ctrl cin = *da.cin;
  if (cin.done) {  // keep both control blocks identical once finished
    if (blockIdx.x == 0 && threadIdx.x == 0) *da.cout = cin;
    return;
  }
  gather_phase(ga, g, 1 - cin.cur, s_red);
  grid_barrier(bar, ga.fault);
  decide_phase<T>(da, g, cin, sm);
}

}  // namespace glp
