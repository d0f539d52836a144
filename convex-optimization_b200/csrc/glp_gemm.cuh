// glp_gemm.cuh -- K5: tensor-core products for the wide multi-RHS mode (l >= 16, fp32 data).
//
// With l = 128 right-hand sides the two products of an iteration are real GEMMs
//   forward     A  [m x n] . [xn | xnt] [n x 2l]      (gl_*_primal.py:25,61/63, batched over 2l columns)
//   transposed  A^T[n x m] . r          [m x l]       (gl_ProxGD_primal.py:129 / gl_FProxGD_primal.py:66)
// and belong on the 5th-generation tensor cores.  Both are computed by ONE kernel
//   D[M x N] = P[M x K] . Q[N x K]^T ,  P row-major (K contiguous), Q planar (K contiguous), D planar fp64
// with P = A (forward) or the resident transposed copy A^T (transposed pass), so both operands are
// K-major and need no MN-major descriptors.
//
// Structure (one persistent CTA per SM, 6 warps):
//   warp 0   TMA producer: 128 x 32 (P) and NT x 32 (Q) fp32 boxes, 128-byte swizzle, into a 3-6 stage ring
//   warp 1   allocates TMEM, one elected lane issues tcgen05.mma.cta_group::1.kind::tf32 (M=128, N=NT<=128, K=8),
//            tcgen05.commit releases the smem stage / publishes a chunk accumulator
//   warps2-9 epilogue: tcgen05.ld 32 lanes x 16 columns at a time into fp64 REGISTER accumulators (64 per
//            thread) that live across the K chunks of a tile; the chunk accumulator is double buffered in
//            TMEM so the drain of chunk i overlaps the MMAs of chunk i+1; coalesced planar fp64 stores
// Work decomposition: items = column groups (NT columns each) x K splits x M tiles, dealt round-robin to
// the persistent CTAs.  148 SMs rarely divide the tile count (128 tiles -> 86 % of the SMs busy), so K
// is split until the item count fills whole waves; partial products are summed in split order afterwards.
// Precision: kind::tf32 keeps 10 mantissa bits of each operand and the tensor core accumulates in fp32
// with truncation (measured: a relative bias of ~(K/8)*2^-24 on the result).  PASSES = 3 runs the
// error-compensated product  P.Q ~= Ph.Qh + (Ph.Ql + Pl.Qh)  on hi/lo fp32 splits resident in HBM; the
// main term is accumulated in TMEM over short K chunks only (k_chunk_blocks*32 elements) and summed in
// fp64 by the epilogue, the two small correction terms in a third TMEM accumulator over the whole tile.
#pragma once
#include <cuda_bf16.h>

#include "glp_device.cuh"
#include <cuda_fp16.h>

#include "glp_types.cuh"

namespace glp {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 32;                 // fp32 elements = 128 bytes = one swizzle row
constexpr int GEMM_EPI_WARPS = 8;
constexpr int GEMM_THREADS = (2 + GEMM_EPI_WARPS) * 32;
constexpr int GEMM_UMMA_K = 8;              // tf32

struct GemmArgs {
  int64_t M, K;
  int32_t N;
  int32_t k_chunk_blocks;                   // K blocks accumulated in fp32 before an fp64 flush (0 = all)
  double* D;                                // planar fp64 [N][ldd]
  int64_t ldd;
  int32_t* fault;
  // solver launches: destination / skip decided on the device from the control block
  const Ctrl* ctrl;                         // nullable
  double* D_alt;                            // role 1: D = ctrl->cur ? D : D_alt  (candidate product buffer 1 - cur)
  int32_t role;                             // 0 plain, 1 forward (skip when done), 2 transposed (skip when done or retry)
  // split-K: work item = (column group, K split, M tile); split s writes its partial product to
  // D + s * split_stride (the caller sums the partials in split order: glp_sum_splits_kernel)
  int32_t splits;                           // >= 1
  int64_t split_stride;                     // doubles between the partial outputs
  // thread-block cluster of ct x cg CTAs working on ct neighbouring M tiles x cg column groups of the same K
  // range: every P tile is loaded once per cluster row (each of the cg CTAs fetches 1/cg of its rows and TMA
  // multicasts them to the others), every Q tile once per cluster column.  1 x 1 = no cluster.
  int32_t ct, cg;
  // device-side gate (glp_thrfix.cuh): when set, the launch does nothing unless *gate > gate_cap
  const int32_t* gate;
  int32_t gate_cap;
  // PASSES = 4 (fp16 pairs): the operands were scaled by powers of two; the epilogue multiplies by *scale
  const double* scale;
};

template <int NT, int PASSES> constexpr int gemm_stage_bytes() {
  return (PASSES >= 2 ? 2 : 1) * (GEMM_BM * 128 + NT * 128);
}
template <int NT, int PASSES> constexpr int gemm_stages() {
  return (200 * 1024) / gemm_stage_bytes<NT, PASSES>() > 8 ? 8 : (200 * 1024) / gemm_stage_bytes<NT, PASSES>();
}
template <int NT, int PASSES> constexpr int gemm_smem_bytes() {
  return gemm_stages<NT, PASSES>() * gemm_stage_bytes<NT, PASSES>() + 1024 /*align*/ + 256 /*barriers*/;
}

// Split operand rings of the fp16-pair product (RINGS): the P tiles stream from HBM, the Q tiles are re-read from
// L2 by every M tile.  With one ring of {P | Q} stages half of the shared memory holds Q, and only 2 x 32 KB of P
// per SM are in flight ahead of the tensor core -- too little to cover the loaded HBM latency (ncu: 5.0 TB/s at 60 %
// tensor activity, profiles/r01_ncu_full_gemm_f16x2_fwd_tpass_C5.csv).  Two rings with their own depths put 224 KB
// to work: a deeper one for P (HBM latency), a shallower one for Q (L2 latency).
template <int NT> constexpr int gemm_ring_p_stages() { return NT == 128 ? 4 : (NT == 64 ? 5 : 6); }
template <int NT> constexpr int gemm_ring_q_stages() { return NT == 128 ? 3 : 4; }
template <int NT> constexpr int gemm_ring_smem_bytes() {
  return gemm_ring_p_stages<NT>() * 2 * GEMM_BM * 128 + gemm_ring_q_stages<NT>() * 2 * NT * 128 + 1024 /*align*/ + 256 /*barriers*/;
}
constexpr int GEMM_RING_THREADS = GEMM_THREADS + 32;   // + the Q producer warp

// ---- tcgen05 wrappers (hand-written PTX) -----------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem desc] . B[smem desc]^T, issued by ONE thread for the whole CTA
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// same with bf16 operands (kind::f16, K = 16 per instruction, twice the tf32 rate)
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// the same arrive delivered to the mbarrier at this offset in every CTA of the cluster named in cta_mask
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(cta_mask) : "memory");
}
// mbarrier arrive when all previously issued MMAs of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile [rows][32 fp32] written by TMA with CU_TENSOR_MAP_SWIZZLE_128B: 8-row groups are
// 1024 bytes apart (SBO), LBO unused, descriptor version 1 (Blackwell), layout type 2 (SWIZZLE_128B).
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= uint64_t((smem_addr >> 4) & 0x3fffu);
  d |= uint64_t(1024u >> 4) << 32;
  d |= uint64_t(1) << 46;
  d |= uint64_t(2) << 61;
  return d;
}
// K-major bf16 tile [rows][32 bf16 = 64 bytes] written by TMA with CU_TENSOR_MAP_SWIZZLE_64B: 8-row groups are
// 512 bytes apart, layout type 4 (SWIZZLE_64B)
__device__ __forceinline__ uint64_t umma_desc_k_sw64(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= uint64_t((smem_addr >> 4) & 0x3fffu);
  d |= uint64_t(512u >> 4) << 32;
  d |= uint64_t(1) << 46;
  d |= uint64_t(4) << 61;
  return d;
}
// kind::f16 with bf16 operands (format 1), fp32 accumulate, both K-major, M = 128
template <int NT> __device__ __forceinline__ constexpr uint32_t umma_idesc_bf16() {
  return (1u << 4) | (1u << 7) | (1u << 10) | (uint32_t(NT >> 3) << 17) | (uint32_t(GEMM_BM >> 4) << 24);
}
// kind::f16 with fp16 operands (format 0), fp32 accumulate, both K-major, M = 128
template <int NT> __device__ __forceinline__ constexpr uint32_t umma_idesc_f16() {
  return (1u << 4) | (uint32_t(NT >> 3) << 17) | (uint32_t(GEMM_BM >> 4) << 24);
}
// kind::tf32, fp32 accumulate, both operands K-major, M = 128
template <int NT> __device__ __forceinline__ constexpr uint32_t umma_idesc_tf32() {
  return (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(NT >> 3) << 17) | (uint32_t(GEMM_BM >> 4) << 24);
}

// hi = a with the 13 low mantissa bits cleared after rounding to nearest (what kind::tf32 keeps), lo = a - hi
__device__ __forceinline__ void split_tf32_one(float a, float& h, float& lo) {
  uint32_t u = __float_as_uint(a);
  u = (u + 0x00000fffu + ((u >> 13) & 1u)) & 0xffffe000u;  // round to nearest even on bit 13
  h = __uint_as_float(u);
  if (!(fabsf(h) <= 3.0e38f)) h = a;                        // inf / nan / overflow of the rounding: keep a
  lo = a - h;
}
// gate (nullable): the launch does nothing unless *gate > gate_cap (glp_thrfix.cuh)
__global__ void glp_split_tf32_kernel(const float* src, float* hi, float* lo,  // hi may alias src (in place)
                                      int64_t rows, int64_t cols, int64_t ld_src, int64_t ld_dst,
                                      const int32_t* gate, int gate_cap) {
  if (gate != nullptr && *gate <= gate_cap) return;
  const int64_t total = rows * cols;
  const int64_t step = int64_t(gridDim.x) * blockDim.x, first = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (ld_src == cols && ld_dst == cols && (total & 3) == 0 &&
      ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(hi) | reinterpret_cast<uintptr_t>(lo)) & 15) == 0) {
    // contiguous operand (the solver's small operands): 128-bit accesses, no index arithmetic
    for (int64_t e = first; e < total / 4; e += step) {
      const float4 a = reinterpret_cast<const float4*>(src)[e];
      float4 h, l;
      split_tf32_one(a.x, h.x, l.x); split_tf32_one(a.y, h.y, l.y);
      split_tf32_one(a.z, h.z, l.z); split_tf32_one(a.w, h.w, l.w);
      reinterpret_cast<float4*>(hi)[e] = h;
      reinterpret_cast<float4*>(lo)[e] = l;
    }
    return;
  }
  for (int64_t e = first; e < total; e += step) {
    const int64_t i = e / cols, j = e - i * cols;
    float h, l;
    split_tf32_one(src[i * ld_src + j], h, l);
    hi[i * ld_dst + j] = h;
    lo[i * ld_dst + j] = l;
  }
}

// dst[j][i] = src[i][j]  (32 x 32 tiles through shared memory), split into hi / lo
__global__ void glp_transpose_split_kernel(const float* __restrict__ src, float* __restrict__ hi, float* __restrict__ lo,
                                           int64_t rows, int64_t cols, int64_t ld_src, int64_t ld_dst) {
  __shared__ float t[32][33];
  const int64_t bi = int64_t(blockIdx.y) * 32, bj = int64_t(blockIdx.x) * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t i = bi + r, j = bj + threadIdx.x;
    t[r][threadIdx.x] = (i < rows && j < cols) ? src[i * ld_src + j] : 0.f;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t j = bj + r, i = bi + threadIdx.x;
    if (j < cols && i < rows) {
      float h, l;
      split_tf32_one(t[threadIdx.x][r], h, l);
      hi[j * ld_dst + i] = h;
      lo[j * ld_dst + i] = l;
    }
  }
}

// PASSES = 2 operands: hi as above (fp32 storage, TF32-exact), plus bf16 copies of hi and of lo = a - hi for
// the two correction products (their result is 2^-11 of the main term, so 8 mantissa bits suffice).
__device__ __forceinline__ void split_bf16_one(float a, float& h, __nv_bfloat16& hb, __nv_bfloat16& lb) {
  uint32_t u = __float_as_uint(a);
  u = (u + 0x00000fffu + ((u >> 13) & 1u)) & 0xffffe000u;
  h = __uint_as_float(u);
  if (!(fabsf(h) <= 3.0e38f)) h = a;
  hb = __float2bfloat16_rn(h);
  lb = __float2bfloat16_rn(a - h);
}
__global__ void glp_split_bf16_kernel(const float* src, float* hi, __nv_bfloat16* hb, __nv_bfloat16* lb,  // hi may alias src
                                      int64_t rows, int64_t cols, int64_t ld_src, int64_t ld_dst,
                                      const int32_t* gate, int gate_cap) {
  if (gate != nullptr && *gate <= gate_cap) return;
  const int64_t total = rows * cols;
  const int64_t step = int64_t(gridDim.x) * blockDim.x, first = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (ld_src == cols && ld_dst == cols && (total & 3) == 0 &&
      ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(hi)) & 15) == 0 &&
      ((reinterpret_cast<uintptr_t>(hb) | reinterpret_cast<uintptr_t>(lb)) & 7) == 0) {
    for (int64_t e = first; e < total / 4; e += step) {
      const float4 a = reinterpret_cast<const float4*>(src)[e];
      float4 h;
      __nv_bfloat16 b0[4], b1[4];
      split_bf16_one(a.x, h.x, b0[0], b1[0]); split_bf16_one(a.y, h.y, b0[1], b1[1]);
      split_bf16_one(a.z, h.z, b0[2], b1[2]); split_bf16_one(a.w, h.w, b0[3], b1[3]);
      reinterpret_cast<float4*>(hi)[e] = h;
      reinterpret_cast<uint2*>(hb)[e] = *reinterpret_cast<const uint2*>(b0);
      reinterpret_cast<uint2*>(lb)[e] = *reinterpret_cast<const uint2*>(b1);
    }
    return;
  }
  for (int64_t e = first; e < total; e += step) {
    const int64_t i = e / cols, j = e - i * cols;
    float h; __nv_bfloat16 b0, b1;
    split_bf16_one(src[i * ld_src + j], h, b0, b1);
    hi[i * ld_dst + j] = h;
    hb[i * ld_dst + j] = b0;
    lb[i * ld_dst + j] = b1;
  }
}
__global__ void glp_transpose_split_bf16_kernel(const float* __restrict__ src, float* __restrict__ hi,
                                                __nv_bfloat16* __restrict__ hb, __nv_bfloat16* __restrict__ lb,
                                                int64_t rows, int64_t cols, int64_t ld_src, int64_t ld_dst) {
  __shared__ float t[32][33];
  const int64_t bi = int64_t(blockIdx.y) * 32, bj = int64_t(blockIdx.x) * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t i = bi + r, j = bj + threadIdx.x;
    t[r][threadIdx.x] = (i < rows && j < cols) ? src[i * ld_src + j] : 0.f;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t j = bj + r, i = bi + threadIdx.x;
    if (j < cols && i < rows) {
      float h; __nv_bfloat16 b0, b1;
      split_bf16_one(t[threadIdx.x][r], h, b0, b1);
      hi[j * ld_dst + i] = h;
      hb[j * ld_dst + i] = b0;
      lb[j * ld_dst + i] = b1;
    }
  }
}

__global__ void glp_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t total) {
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x)
    dst[e] = __float2bfloat16_rn(src[e]);
}

// ---- PASSES = 4 operands: fp16 pairs of the scaled value ---------------------------------------------------
// v = a * s (s a power of two chosen so that max|v| < 2^15): v0 = fp16(v), v1 = fp16(v - v0); v0 + v1 carries 22
// significant bits of v while v1 is a normal number (|v| >= 2^-3), and is within 2^-25 of v below that.
__device__ __forceinline__ void split_f16_one(float a, float s, __half& v0, __half& v1) {
  const float v = a * s;
  v0 = __float2half_rn(v);
  v1 = __float2half_rn(v - __half2float(v0));
}
// power-of-two scale that brings max|.| = mx into [2^14, 2^15); 1 for zero / non-finite input
__device__ __host__ __forceinline__ float f16_scale_for(float mx) {
  if (!(mx > 0.0f) || !(mx < 3.0e38f)) return 1.0f;
  int e = 0;
  frexpf(mx, &e);            // mx = f * 2^e, f in [0.5, 1)  ->  mx * 2^(15 - e) in [2^14, 2^15)
  int k = 15 - e;
  k = k > 126 ? 126 : (k < -126 ? -126 : k);
  return ldexpf(1.0f, k);
}
// max |src| as float bits (non-negative floats order like unsigned integers); *out zeroed by the caller
__global__ void glp_absmax_kernel(const float* __restrict__ src, int64_t rows, int64_t cols, int64_t ld, unsigned* out) {
  unsigned mx = 0u;
  const int64_t total = rows * cols;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = e / cols, j = e - i * cols;
    const unsigned u = __float_as_uint(fabsf(src[i * ld + j]));
    mx = u > mx ? u : mx;
  }
  mx = __reduce_max_sync(0xffffffffu, mx);
  if ((threadIdx.x & 31) == 0 && mx != 0u) atomicMax(out, mx);
}
// Per-column and per-row maxima of |src| as float bits (col_max[cols], row_max[rows], zeroed by the caller): the
// dynamic range of the COLUMN / ROW scales of A.  The fp16-pair operand set keeps ~2^-40 of the matrix maximum in
// absolute terms: entries that are small only because their whole column (or row) is scaled down -- un-normalised
// feature columns -- would lose their significance against a large x entry, while an fp32 product keeps 2^-24 of
// every entry.  Isolated small entries of a normally scaled column do not matter (their products are small too).
__global__ void glp_rowcol_absmax_kernel(const float* __restrict__ src, int64_t rows, int64_t cols, int64_t ld,
                                         unsigned* __restrict__ col_max, unsigned* __restrict__ row_max) {
  const int64_t total = rows * cols;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = e / cols, j = e - i * cols;
    const unsigned u = __float_as_uint(fabsf(src[i * ld + j]));
    if (u > __ldg(&col_max[j])) atomicMax(&col_max[j], u);
    if (u > __ldg(&row_max[i])) atomicMax(&row_max[i], u);
  }
}
// out[0] / out[1] = largest / smallest non-zero entry of v (float bits); out zero-initialised to {0, 0x7f800000}
__global__ void glp_minmax_nonzero_kernel(const unsigned* __restrict__ v, int64_t n, unsigned* out) {
  unsigned mx = 0u, mn = 0x7f800000u;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < n; e += int64_t(gridDim.x) * blockDim.x) {
    const unsigned u = v[e];
    if (u != 0u) { mx = u > mx ? u : mx; mn = u < mn ? u : mn; }
  }
  mx = __reduce_max_sync(0xffffffffu, mx);
  mn = __reduce_min_sync(0xffffffffu, mn);
  if ((threadIdx.x & 31) == 0) { if (mx != 0u) atomicMax(&out[0], mx); atomicMin(&out[1], mn); }
}
// contiguous small operand (x or r) with the scale taken from its own max on the device: *amax_bits (float bits from
// glp_absmax_kernel) or, when scal_partial is given, the max over slot 3 of the prox epilogue's per-block partials
// (max|xn|, ncb blocks) -- every block reduces those few hundred values itself, which saves a pass over the operand.
// Block 0 publishes inv_scale_out = 1 / (s * s_other) for the product's epilogue.  gate as in glp_split_tf32_kernel.
__global__ void __launch_bounds__(256)
glp_split_f16_kernel(const float* __restrict__ src, __half* __restrict__ o0, __half* __restrict__ o1,
                     int64_t total, const unsigned* amax_bits, const double* scal_partial, int ncb, float s_other,
                     const unsigned* other_amax_bits, double* inv_scale_out, const int32_t* gate, int gate_cap) {
  __shared__ float s_mx[8];
  pdl_wait();
  pdl_launch();
  if (gate != nullptr && *gate <= gate_cap) return;
  float mxf;
  if (scal_partial != nullptr) {
    float m = 0.0f;   // NaN-propagating max, as in the prox epilogue
    for (int q = threadIdx.x; q < ncb; q += blockDim.x) {
      const float v = float(__ldcg(&scal_partial[SCAL_STRIDE * q + 3]));
      m = (m != m) ? m : ((v != v) ? v : fmaxf(m, v));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float v = __shfl_xor_sync(0xffffffffu, m, o);
      m = (m != m) ? m : ((v != v) ? v : fmaxf(m, v));
    }
    if ((threadIdx.x & 31) == 0) s_mx[threadIdx.x >> 5] = m;
    __syncthreads();
    m = s_mx[0];
    for (int w = 1; w < int(blockDim.x >> 5); ++w) {
      const float v = s_mx[w];
      m = (m != m) ? m : ((v != v) ? v : fmaxf(m, v));
    }
    mxf = m;
  } else {
    mxf = __uint_as_float(*amax_bits);
  }
  const float s = f16_scale_for(mxf);
  if (inv_scale_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    const float so = other_amax_bits != nullptr ? f16_scale_for(__uint_as_float(*other_amax_bits)) : s_other;
    *inv_scale_out = 1.0 / (double(s) * double(so));
  }
  const int64_t step = int64_t(gridDim.x) * blockDim.x, first = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if ((total & 3) == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0 &&
      ((reinterpret_cast<uintptr_t>(o0) | reinterpret_cast<uintptr_t>(o1)) & 7) == 0) {
    for (int64_t e = first; e < total / 4; e += step) {
      const float4 a = reinterpret_cast<const float4*>(src)[e];
      __half h0[4], h1[4];
      split_f16_one(a.x, s, h0[0], h1[0]); split_f16_one(a.y, s, h0[1], h1[1]);
      split_f16_one(a.z, s, h0[2], h1[2]); split_f16_one(a.w, s, h0[3], h1[3]);
      reinterpret_cast<uint2*>(o0)[e] = *reinterpret_cast<const uint2*>(h0);
      reinterpret_cast<uint2*>(o1)[e] = *reinterpret_cast<const uint2*>(h1);
    }
    return;
  }
  for (int64_t e = first; e < total; e += step) split_f16_one(src[e], s, o0[e], o1[e]);
}
// resident operand (A or its transpose) with a fixed scale
__global__ void glp_split_f16_2d_kernel(const float* __restrict__ src, __half* __restrict__ o0, __half* __restrict__ o1,
                                        int64_t rows, int64_t cols, int64_t ld_src, int64_t ld_dst, float s) {
  const int64_t total = rows * cols;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = e / cols, j = e - i * cols;
    split_f16_one(src[i * ld_src + j], s, o0[i * ld_dst + j], o1[i * ld_dst + j]);
  }
}
__global__ void glp_transpose_split_f16_kernel(const float* __restrict__ src, __half* __restrict__ o0, __half* __restrict__ o1,
                                               int64_t rows, int64_t cols, int64_t ld_src, int64_t ld_dst, float s) {
  __shared__ float t[32][33];
  const int64_t bi = int64_t(blockIdx.y) * 32, bj = int64_t(blockIdx.x) * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t i = bi + r, j = bj + threadIdx.x;
    t[r][threadIdx.x] = (i < rows && j < cols) ? src[i * ld_src + j] : 0.f;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t j = bj + r, i = bi + threadIdx.x;
    if (j < cols && i < rows) split_f16_one(t[threadIdx.x][r], s, o0[j * ld_dst + i], o1[j * ld_dst + i]);
  }
}

// ---- tiled resident operand of the CTA-pair kernel (glp_gemm_pair.cuh) ------------------------------------------
// The streamed operand P [rows x cols] (cols = K) is stored as [rows/128][K/64] tiles of 128 rows x 64 fp16 (128 bytes
// per row, 16 KB per tile, contiguous): the tile a CTA fetches for one K block is ONE contiguous 16 KB run of HBM
// instead of 128 runs of 128 bytes that are a whole matrix row (>= 32 KB) apart -- the access pattern that kept the
// row-major operand at 53 % of the DRAM peak (ncu, profiles/r02_ncu_gemm_pair_rowmajor.txt).
__host__ __device__ __forceinline__ int64_t f16_tiled_off(int64_t r, int64_t k, int64_t nkb) {
  return (((r >> 7) * nkb + (k >> 6)) << 13) + ((r & 127) << 6) + (k & 63);
}
__global__ void glp_split_f16_tiled_kernel(const float* __restrict__ src, __half* __restrict__ o0, __half* __restrict__ o1,
                                           int64_t rows, int64_t cols, int64_t ld_src, int64_t nkb, float s) {
  const int64_t total = rows * cols;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = e / cols, j = e - i * cols;
    const int64_t o = f16_tiled_off(i, j, nkb);
    split_f16_one(src[i * ld_src + j], s, o0[o], o1[o]);
  }
}
// the same with the scale taken on the device from the operand's maximum (test / probe entry point glp_gemm_tf32)
__global__ void glp_split_f16_tiled_dev_scale_kernel(const float* __restrict__ src, __half* __restrict__ o0,
                                                     __half* __restrict__ o1, int64_t rows, int64_t cols, int64_t ld_src,
                                                     int64_t nkb, const unsigned* amax_bits) {
  const float s = f16_scale_for(__uint_as_float(*amax_bits));
  const int64_t total = rows * cols;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = e / cols, j = e - i * cols;
    const int64_t o = f16_tiled_off(i, j, nkb);
    split_f16_one(src[i * ld_src + j], s, o0[o], o1[o]);
  }
}
// the transposed operand: dst(j, i) = src(i, j), tiled over (rows = j < cols of src, K = i < rows of src)
__global__ void glp_transpose_split_f16_tiled_kernel(const float* __restrict__ src, __half* __restrict__ o0,
                                                     __half* __restrict__ o1, int64_t rows, int64_t cols, int64_t ld_src,
                                                     int64_t nkb, float s) {
  __shared__ float t[32][33];
  const int64_t bi = int64_t(blockIdx.y) * 32, bj = int64_t(blockIdx.x) * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t i = bi + r, j = bj + threadIdx.x;
    t[r][threadIdx.x] = (i < rows && j < cols) ? src[i * ld_src + j] : 0.f;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int64_t j = bj + r, i = bi + threadIdx.x;
    if (j < cols && i < rows) {
      const int64_t o = f16_tiled_off(j, i, nkb);
      split_f16_one(t[threadIdx.x][r], s, o0[o], o1[o]);
    }
  }
}

// work item w -> (K split, M tile, column group), column group fastest: the CTAs that take the column
// groups of one P tile run side by side, so P is read from HBM once and from L2 by the others (with
// N = 2l = 256 the product would otherwise stream A twice and turn HBM-bound); neighbouring CTAs also
// share the Q slices of their K range.  Split s covers the chunks [s*nchunks/S, (s+1)*nchunks/S).
struct GemmItem { int64_t group, split, tile, ch0, ch1, kb0, kb1; };
__device__ __forceinline__ GemmItem gemm_item(int64_t w, int64_t tiles, int64_t G, int64_t S, int64_t nchunks, int64_t kc,
                                              int64_t nkb) {
  GemmItem it;
  const int64_t ts = w / G;
  it.group = w - ts * G;
  it.split = ts / tiles;
  it.tile = ts - it.split * tiles;
  it.ch0 = it.split * nchunks / S;
  it.ch1 = (it.split + 1) * nchunks / S;
  it.kb0 = it.ch0 * kc;
  it.kb1 = it.ch1 * kc < nkb ? it.ch1 * kc : nkb;
  return it;
}

// out[i] = sum_s part[s * stride + i] in split order (fixed, deterministic); solver launches pick the
// destination / skip from the control block exactly like the product kernel.
__global__ void glp_sum_splits_kernel(const double* __restrict__ part, int splits, int64_t stride, int64_t M, int N,
                                      int64_t ldd, double* out, double* out_alt, const Ctrl* ctrl, int role,
                                      const int32_t* gate, int gate_cap) {
  if (gate != nullptr && *gate <= gate_cap) return;
  if (ctrl != nullptr) {
    if (ctrl->done || (role == 2 && ctrl->retry)) return;
    if (role == 1 && ctrl->cur == 0) out = out_alt;
  }
  const int64_t total = M * N;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int64_t col = e / M, i = col * ldd + (e - col * M);
    double v = part[i];
    for (int s = 1; s < splits; ++s) v += part[s * stride + i];
    out[i] = v;
  }
}

template <int NT, int PASSES, bool RINGS = false>
__global__ void __launch_bounds__(RINGS ? GEMM_RING_THREADS : GEMM_THREADS, 1)
glp_gemm_tf32_kernel(const __grid_constant__ CUtensorMap tmPh, const __grid_constant__ CUtensorMap tmPl,
                     const __grid_constant__ CUtensorMap tmQh, const __grid_constant__ CUtensorMap tmQl,
                     const __grid_constant__ CUtensorMap tmPx, const __grid_constant__ CUtensorMap tmQx,
                     const GemmArgs a) {
  // PASSES = 1: Ph.Qh (plain TF32).  PASSES = 3: + Ph.Ql + Pl.Qh on TF32 hi/lo splits (tmPl / tmQl = fp32 lo).
  // PASSES = 2: the same two corrections on bf16 operands (kind::f16): tmPx / tmQx = bf16(hi), tmPl / tmQl = bf16(lo);
  //             stage = Ph | Qh | bf16 Ph | bf16 Pl | bf16 Qh | bf16 Ql  (same bytes as PASSES = 3)
  // PASSES = 4: fp16 pairs, P = (P0 + P1) / sP and Q = (Q0 + Q1) / sQ with power-of-two scales: P0.Q0 + P0.Q1 + P1.Q0 on
  //             kind::f16 (same 11-bit significands as TF32, twice the rate, half the bytes); tmPh / tmPl = P0 / P1,
  //             tmQh / tmQl = Q0 / Q1; a K block is 64 elements (128-byte rows, same stage layout as PASSES = 3)
  constexpr bool CORR = PASSES >= 2;
  constexpr bool F16X2 = PASSES == 4;
  constexpr int BK = F16X2 ? 2 * GEMM_BK : GEMM_BK;
  constexpr int PB_BYTES = GEMM_BM * 64, QB_BYTES = NT * 64;   // bf16 tiles
  constexpr int STAGES = gemm_stages<NT, PASSES>();
  constexpr int STAGE_BYTES = gemm_stage_bytes<NT, PASSES>();
  constexpr int P_BYTES = GEMM_BM * 128, Q_BYTES = NT * 128;
  // TMEM: two chunk accumulators (main term, double buffered) + one whole-tile accumulator of the two
  // correction terms (their magnitude is 2^-11 of the main term, so their rounding is harmless)
  constexpr uint32_t TMEM_NEED = (CORR ? 3 : 2) * NT;
  constexpr uint32_t TMEM_COLS = TMEM_NEED <= 32 ? 32 : TMEM_NEED <= 64 ? 64 : TMEM_NEED <= 128 ? 128 : TMEM_NEED <= 256 ? 256 : 512;
  constexpr int HC = NT / 2;  // columns per epilogue thread
  static_assert(NT == 32 || NT == 64 || NT == 128, "NT");
  static_assert(!RINGS || PASSES == 4, "split rings: fp16-pair product only");
  // RINGS: P ring [SP] of {P0 | P1} (barriers full / empty, freed by this CTA's MMAs alone: P is not shared),
  //        Q ring [SQ] of {Q0 | Q1} (barriers fullq / emptyq, freed in every CTA of the cluster); launched with cg = 1
  constexpr int SP = RINGS ? gemm_ring_p_stages<NT>() : STAGES, SQ = RINGS ? gemm_ring_q_stages<NT>() : 0;
  constexpr int RING_BYTES = RINGS ? SP * 2 * P_BYTES + SQ * 2 * Q_BYTES : STAGES * STAGE_BYTES;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_q = smem + SP * 2 * P_BYTES;   // RINGS only
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + RING_BYTES);
  uint64_t* empty = full + SP;
  uint64_t* fullq = empty + SP;       // [SQ]
  uint64_t* emptyq = fullq + SQ;      // [SQ]
  uint64_t* tfull = emptyq + SQ;      // [2] chunk accumulator ready
  uint64_t* tempty = tfull + 2;       // [2] chunk accumulator drained
  uint64_t* cempty = tempty + 2;      // [1] correction accumulator drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(cempty + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  double* Dout = a.D;
  if (a.ctrl != nullptr) {  // grid-uniform
    if (a.ctrl->done || (a.role == 2 && a.ctrl->retry)) return;
    if (a.role == 1 && a.ctrl->cur == 0) Dout = a.D_alt;
  }
  if (a.gate != nullptr && *a.gate <= a.gate_cap) return;   // grid-uniform, before any barrier or cluster traffic
  const int64_t tiles = (a.M + GEMM_BM - 1) / GEMM_BM;
  const int64_t nkb = (a.K + BK - 1) / BK;
  const int64_t kc = a.k_chunk_blocks > 0 ? a.k_chunk_blocks : nkb;
  const int64_t nchunks = (nkb + kc - 1) / kc;
  const int64_t S = a.splits > 1 ? a.splits : 1;
  const int64_t G = (a.N + NT - 1) / NT;
  // cluster geometry: this CTA is (dt, dg) of a ct x cg cluster; clusters are the scheduling unit
  const int CT = a.ct > 1 ? a.ct : 1, CG = a.cg > 1 ? a.cg : 1, CS = CT * CG;
  const int crank = CS > 1 ? int(cluster_ctarank()) : 0;
  const int dt = crank % CT, dg = crank / CT;
  const int64_t cl_id = blockIdx.x / CS, ncl = gridDim.x / CS;
  const int64_t tblocks = (tiles + CT - 1) / CT, gblocks = (G + CG - 1) / CG;
  const int64_t items = gblocks * S * tblocks;
  uint16_t mask_p = 0, mask_q = 0;   // CTAs that share my P tile (same dt) / my Q tile (same dg)
  for (int q = 0; q < CG; ++q) mask_p |= uint16_t(1u << (dt + CT * q));
  for (int q = 0; q < CT; ++q) mask_q |= uint16_t(1u << (q + CT * dg));
  const uint16_t mask_all = uint16_t((1u << CS) - 1u);

  if (tid == 0) {
    for (int s = 0; s < SP; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], RINGS ? 1u : uint32_t(CS)); }
    for (int s = 0; s < SQ; ++s) { mbar_init(&fullq[s], 1); mbar_init(&emptyq[s], uint32_t(CS)); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], GEMM_EPI_WARPS); }
    mbar_init(cempty, GEMM_EPI_WARPS);
    fence_mbar_init();
    fence_proxy_async();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (CS > 1) cluster_sync_all();   // every CTA's barriers exist before a peer multicasts into them

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      prefetch_tensormap(&tmPh); prefetch_tensormap(&tmQh);
      if (PASSES == 2) { prefetch_tensormap(&tmPx); prefetch_tensormap(&tmQx); }
      const uint64_t pol_stream = policy_evict_first(), pol_keep = policy_evict_normal();
      int stage = 0; uint32_t phase = 0;
      // my share of each shared tile: rows [dg*BM/CG, ...) of P, rows [dt*NT/CT, ...) of Q (box sizes of the maps)
      const int pr = dg * (GEMM_BM / CG), qr = dt * (NT / CT);
      auto load_p = [&](uint8_t* tile, const CUtensorMap* m, uint64_t* bar, int32_t k0, int32_t r0, int row_bytes) {
        if (CG > 1) tma_load_2d_mc(tile + pr * row_bytes, m, bar, k0, r0 + pr, mask_p, pol_stream);
        else tma_load_2d(tile, m, bar, k0, r0, pol_stream);
      };
      auto load_q = [&](uint8_t* tile, const CUtensorMap* m, uint64_t* bar, int32_t k0, int32_t q0, int row_bytes) {
        if (CT > 1) tma_load_2d_mc(tile + qr * row_bytes, m, bar, k0, q0 + qr, mask_q, pol_keep);
        else tma_load_2d(tile, m, bar, k0, q0, pol_keep);
      };
      for (int64_t w = cl_id; w < items; w += ncl) {
        const GemmItem it = gemm_item(w, tblocks, gblocks, S, nchunks, kc, nkb);
        const int32_t r0 = int32_t((it.tile * CT + dt) * GEMM_BM), q0 = int32_t((it.group * CG + dg) * NT);
        if (RINGS) {   // P only: my own tile, my own slot (the Q producer is warp 10)
          for (int64_t kb = it.kb0; kb < it.kb1; ++kb) {
            mbar_wait(&empty[stage], phase ^ 1u, a.fault);
            mbar_arrive_expect_tx(&full[stage], 2 * P_BYTES);
            uint8_t* s = smem + stage * 2 * P_BYTES;
            const int32_t k0 = int32_t(kb * BK);
            tma_load_2d(s, &tmPh, &full[stage], k0, r0, pol_stream);
            tma_load_2d(s + P_BYTES, &tmPl, &full[stage], k0, r0, pol_stream);
            if (++stage == SP) { stage = 0; phase ^= 1u; }
          }
        } else
        for (int64_t kb = it.kb0; kb < it.kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1u, a.fault);   // the slot is free in EVERY CTA I multicast into
          mbar_arrive_expect_tx(&full[stage], STAGE_BYTES);
          uint8_t* s = smem + stage * STAGE_BYTES;
          const int32_t k0 = int32_t(kb * BK);
          load_p(s, &tmPh, &full[stage], k0, r0, 128);
          load_q(s + P_BYTES, &tmQh, &full[stage], k0, q0, 128);
          if (PASSES == 3 || F16X2) {
            load_p(s + P_BYTES + Q_BYTES, &tmPl, &full[stage], k0, r0, 128);
            load_q(s + 2 * P_BYTES + Q_BYTES, &tmQl, &full[stage], k0, q0, 128);
          }
          if (PASSES == 2) {
            uint8_t* x = s + P_BYTES + Q_BYTES;
            load_p(x, &tmPx, &full[stage], k0, r0, 64);
            load_p(x + PB_BYTES, &tmPl, &full[stage], k0, r0, 64);
            load_q(x + 2 * PB_BYTES, &tmQx, &full[stage], k0, q0, 64);
            load_q(x + 2 * PB_BYTES + QB_BYTES, &tmQl, &full[stage], k0, q0, 64);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (RINGS && warp == GEMM_THREADS / 32) {
    // ------------------------------------------------------------------ TMA producer of the Q ring (RINGS)
    if (lane == 0) {
      prefetch_tensormap(&tmQh); prefetch_tensormap(&tmQl);
      const uint64_t pol_keep = policy_evict_normal();
      int stage = 0; uint32_t phase = 0;
      const int qr = dt * (NT / CT);   // my share of the rows of the Q tile the cluster's M tiles have in common
      for (int64_t w = cl_id; w < items; w += ncl) {
        const GemmItem it = gemm_item(w, tblocks, gblocks, S, nchunks, kc, nkb);
        const int32_t q0 = int32_t(it.group * NT);
        for (int64_t kb = it.kb0; kb < it.kb1; ++kb) {
          mbar_wait(&emptyq[stage], phase ^ 1u, a.fault);   // the slot is free in EVERY CTA I multicast into
          mbar_arrive_expect_tx(&fullq[stage], 2 * Q_BYTES);
          uint8_t* s = smem_q + stage * 2 * Q_BYTES;
          const int32_t k0 = int32_t(kb * BK);
          if (CT > 1) {
            tma_load_2d_mc(s + qr * 128, &tmQh, &fullq[stage], k0, q0 + qr, mask_q, pol_keep);
            tma_load_2d_mc(s + Q_BYTES + qr * 128, &tmQl, &fullq[stage], k0, q0 + qr, mask_q, pol_keep);
          } else {
            tma_load_2d(s, &tmQh, &fullq[stage], k0, q0, pol_keep);
            tma_load_2d(s + Q_BYTES, &tmQl, &fullq[stage], k0, q0, pol_keep);
          }
          if (++stage == SQ) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_tf32<NT>();
      int stage = 0; uint32_t phase = 0;
      int qstage = 0; uint32_t qphase = 0;   // RINGS
      int acc = 0; uint32_t acc_phase = 0, c_phase = 0;
      const uint32_t c_tmem = tmem_base + uint32_t(2 * NT);
      for (int64_t w = cl_id; w < items; w += ncl) {
        const GemmItem it = gemm_item(w, tblocks, gblocks, S, nchunks, kc, nkb);
        if (CORR) {
          mbar_wait(cempty, c_phase ^ 1u, a.fault);  // epilogue has read the previous item's corrections
          c_phase ^= 1u;
        }
        uint32_t c_accumulate = 0;
        for (int64_t ch = it.ch0; ch < it.ch1; ++ch) {
          mbar_wait(&tempty[acc], acc_phase ^ 1u, a.fault);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + uint32_t(acc * NT);
          const int64_t kb1 = (ch + 1) * kc < nkb ? (ch + 1) * kc : nkb;
          uint32_t accumulate = 0;
          for (int64_t kb = ch * kc; kb < kb1; ++kb) {
            mbar_wait(&full[stage], phase, a.fault);
            if (RINGS) mbar_wait(&fullq[qstage], qphase, a.fault);
            tc_fence_after();
            const uint32_t sbase = smem_u32(smem + stage * (RINGS ? 2 * P_BYTES : STAGE_BYTES));
            const uint32_t qbase = RINGS ? smem_u32(smem_q + qstage * 2 * Q_BYTES) : sbase + P_BYTES;
            const uint64_t dPh = umma_desc_k_sw128(sbase), dQh = umma_desc_k_sw128(qbase);
            if (F16X2) {
              // 4 instructions of K = 16 fp16 (32 bytes) per 128-byte row, like the 4 x 8 fp32 of the TF32 tile
              constexpr uint32_t idesc_h = umma_idesc_f16<NT>();
              const uint64_t dPl = umma_desc_k_sw128(RINGS ? sbase + P_BYTES : sbase + P_BYTES + Q_BYTES);
              const uint64_t dQl = umma_desc_k_sw128(RINGS ? qbase + Q_BYTES : sbase + 2 * P_BYTES + Q_BYTES);
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4) {
                umma_bf16(d_tmem, dPh + uint64_t(2 * k4), dQh + uint64_t(2 * k4), idesc_h, accumulate);
                accumulate = 1;
              }
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4) {
                umma_bf16(c_tmem, dPh + uint64_t(2 * k4), dQl + uint64_t(2 * k4), idesc_h, c_accumulate);
                c_accumulate = 1;
                umma_bf16(c_tmem, dPl + uint64_t(2 * k4), dQh + uint64_t(2 * k4), idesc_h, 1u);
              }
            } else {
#pragma unroll
              for (int k4 = 0; k4 < GEMM_BK / GEMM_UMMA_K; ++k4) {
                umma_tf32(d_tmem, dPh + uint64_t(2 * k4), dQh + uint64_t(2 * k4), idesc, accumulate);
                accumulate = 1;
              }
            }
            if (PASSES == 3) {
              const uint64_t dPl = umma_desc_k_sw128(sbase + P_BYTES + Q_BYTES);
              const uint64_t dQl = umma_desc_k_sw128(sbase + 2 * P_BYTES + Q_BYTES);
#pragma unroll
              for (int k4 = 0; k4 < GEMM_BK / GEMM_UMMA_K; ++k4) {
                umma_tf32(c_tmem, dPh + uint64_t(2 * k4), dQl + uint64_t(2 * k4), idesc, c_accumulate);
                c_accumulate = 1;
                umma_tf32(c_tmem, dPl + uint64_t(2 * k4), dQh + uint64_t(2 * k4), idesc, 1u);
              }
            }
            if (PASSES == 2) {
              constexpr uint32_t idesc_b = umma_idesc_bf16<NT>();
              const uint32_t xb = sbase + P_BYTES + Q_BYTES;
              const uint64_t dPhb = umma_desc_k_sw64(xb), dPlb = umma_desc_k_sw64(xb + PB_BYTES);
              const uint64_t dQhb = umma_desc_k_sw64(xb + 2 * PB_BYTES), dQlb = umma_desc_k_sw64(xb + 2 * PB_BYTES + QB_BYTES);
#pragma unroll
              for (int k2 = 0; k2 < GEMM_BK / 16; ++k2) {
                umma_bf16(c_tmem, dPhb + uint64_t(2 * k2), dQlb + uint64_t(2 * k2), idesc_b, c_accumulate);
                c_accumulate = 1;
                umma_bf16(c_tmem, dPlb + uint64_t(2 * k2), dQhb + uint64_t(2 * k2), idesc_b, 1u);
              }
            }
            // frees the smem stage when these MMAs have read it -- in every CTA of the cluster, whose producers
            // multicast into this slot
            if (RINGS) {
              umma_commit(&empty[stage]);                                   // my P slot: mine alone
              if (CS > 1) umma_commit_mc(&emptyq[qstage], mask_all);        // the Q slot: the whole cluster writes it
              else umma_commit(&emptyq[qstage]);
              if (++qstage == SQ) { qstage = 0; qphase ^= 1u; }
            } else if (CS > 1) umma_commit_mc(&empty[stage], mask_all);
            else umma_commit(&empty[stage]);
            if (++stage == SP) { stage = 0; phase ^= 1u; }
          }
          umma_commit(&tfull[acc]);      // chunk accumulator (and, after the last chunk, the corrections) complete
          if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
        }
      }
    }
  } else if (warp < GEMM_THREADS / 32) {
    // ------------------------------------------------------------------ epilogue (8 warps)
    // warp -> (TMEM lane quarter it may read, half of the NT columns); fp64 accumulators live in registers
    // across the K chunks of a tile, so the tensor core's fp32 accumulation never spans more than one chunk
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    int acc = 0; uint32_t acc_phase = 0;
    for (int64_t w = cl_id; w < items; w += ncl) {
      const GemmItem it = gemm_item(w, tblocks, gblocks, S, nchunks, kc, nkb);
      const int64_t row = (it.tile * CT + dt) * GEMM_BM + quarter * 32 + lane;
      double sum[HC];
#pragma unroll
      for (int j = 0; j < HC; ++j) sum[j] = 0.0;
      const uint32_t lane_off = uint32_t(quarter * 32) << 16;
      for (int64_t ch = it.ch0; ch < it.ch1; ++ch) {
        mbar_wait(&tfull[acc], acc_phase, a.fault);
        tc_fence_after();
        const uint32_t t0 = tmem_base + uint32_t(acc * NT + half * HC) + lane_off;
#pragma unroll
        for (int c0 = 0; c0 < HC; c0 += 16) {
          uint32_t v[16];
          tmem_ld_32x32b_x16(t0 + uint32_t(c0), v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) sum[c0 + j] += double(__uint_as_float(v[j]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
      if (CORR) {
        const uint32_t t0 = tmem_base + uint32_t(2 * NT + half * HC) + lane_off;
#pragma unroll
        for (int c0 = 0; c0 < HC; c0 += 16) {
          uint32_t v[16];
          tmem_ld_32x32b_x16(t0 + uint32_t(c0), v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) sum[c0 + j] += double(__uint_as_float(v[j]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(cempty);
      }
      if (row < a.M) {
        double* dst = Dout + it.split * a.split_stride;
        const double sc = (F16X2 && a.scale != nullptr) ? *a.scale : 1.0;   // a power of two: exact
#pragma unroll
        for (int j = 0; j < HC; ++j) {
          const int col = int(it.group * CG + dg) * NT + half * HC + j;
          if (col < a.N) dst[int64_t(col) * a.ldd + row] = F16X2 ? sum[j] * sc : sum[j];
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CS > 1) cluster_sync_all();   // no CTA leaves while a peer may still multicast into it or arrive on its barriers
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// After a tensor-core forward product: residual norm partials and the sparsity count of the candidate
// (what K1 fuses into its epilogue / aux warp on the CUDA-core path).  gl_*_primal.py:22,24-27.
struct PostFwdArgs {
  const Ctrl* ctrl;
  const double* AX[2];          // [2l][ldm]
  const double* b;              // [l][ldm]
  const double* scal_partial;   // [ncb][SCAL_STRIDE], slot 3 = max|xn| per column block
  const double* xcand[2];       // candidate x [l][ldn]
  double* rr_out;               // [gridDim]
  double* cnt_out;              // [gridDim]
};
__global__ void __launch_bounds__(256)
glp_post_fwd_kernel(const PostFwdArgs a, const Geom g) {
  __shared__ double s_red[8];
  __shared__ double s_mx;
  pdl_wait();
  pdl_launch();
  if (a.ctrl != nullptr && a.ctrl->done) return;
  const int sel = a.ctrl != nullptr ? 1 - a.ctrl->cur : 0;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  auto nmx = [](double x, double y) { return (x != x) ? x : ((y != y) ? y : fmax(x, y)); };
  double mx = 0.0;
  for (int i = tid; i < g.ncb; i += 256) mx = nmx(mx, __ldcg(&a.scal_partial[SCAL_STRIDE * i + 3]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = nmx(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) s_red[warp] = mx;
  __syncthreads();
  if (tid == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t = nmx(t, s_red[w]);
    s_mx = t;
  }
  __syncthreads();
  const double cut = xmul(1e-6, s_mx);
  const int64_t stride = int64_t(gridDim.x) * 256, first = int64_t(blockIdx.x) * 256 + tid;
  double cnt = 0.0, rr = 0.0;
  // flat loops over the planar arrays (padding rows / row groups are zero and contribute nothing)
  const double* xc = a.xcand[sel];
  const int64_t lenn = int64_t(g.l) * g.ldn, lenm = int64_t(g.l) * g.ldm;
  for (int64_t j = first; j < lenn; j += stride) cnt += fabs(xc[j]) > cut ? 1.0 : 0.0;
  const double* ax = a.AX[sel];
  for (int64_t i = first; i < lenm; i += stride) {
    const double d = ax[i] - a.b[i];
    rr += d * d;
  }
  cnt = warp_sum(cnt); rr = warp_sum(rr);
  __syncthreads();
  if (lane == 0) s_red[warp] = rr;
  __syncthreads();
  double rr_tot = 0.0;
  if (tid == 0) for (int w = 0; w < 8; ++w) rr_tot += s_red[w];
  __syncthreads();
  if (lane == 0) s_red[warp] = cnt;
  __syncthreads();
  if (tid == 0) {
    double c_tot = 0.0;
    for (int w = 0; w < 8; ++w) c_tot += s_red[w];
    a.rr_out[blockIdx.x] = rr_tot;
    a.cnt_out[blockIdx.x] = c_tot;
  }
}

}  // namespace glp
