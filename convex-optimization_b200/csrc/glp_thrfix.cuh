// Wide (tensor-core) mode: the forward product of the thresholded iterate from the product of the iterate.
//
// The reference thresholds the iterate element-wise AFTER the objective is recorded (gl_*_primal.py:
// x[abs(x) < thres] = 0), so every iteration needs A.xn (objective) and A.xnt (next gradient base point).
// On the CUDA-core path the two ride in one stream over A at no extra HBM cost.  On the tensor cores the product
// is compute-bound and the second operand doubles its price, although xnt differs from xn only where
// 0 < |xn| < thres: E := xn - xnt has |E_jc| < thres and (thres = 1e-3, entries of order 1) about one non-zero
// entry in a thousand.  So here
//     A.xnt = A.xn - A.E,        column c:  (A.E)(:, c) = sum over the entries (j, c) of E:  A(:, j) E_jc
// The entries are listed per operand column in ascending row order by one block per column (fixed order ->
// run-to-run identical bits) and applied by a CUDA-core kernel (fp32 products and sums, subtracted from the fp64 product).  A(:, j) is read from the resident
// transposed split (row j of At_hi + At_lo: coalesced over the output rows); |E| < thres makes the 2^-22 relative
// gap between hi + lo and A irrelevant next to the rounding of the TF32 product itself.  Cost: entries x m x 8
// bytes of L2/HBM traffic instead of a second product.
// When a column holds more entries than the list has room for, the lists are abandoned and a second,
// device-gated launch of the product kernel computes A.xnt directly, as before (GemmArgs::gate).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "glp_device.cuh"
#include "glp_gemm.cuh"
#include "glp_types.cuh"

namespace glp {

constexpr int THRLIST_THREADS = 1024;
constexpr int THRFIX_THREADS = 128;

// count[0] = entries listed (all columns), count[1] = columns whose list overflowed (> 0: lists unusable),
// count[2] = arrival counter of the list blocks (zero between launches)
struct ThrListArgs {
  const Ctrl* ctrl;              // nullable; done -> nothing to do
  const float* xop;              // [2l][ldn] operand copy {xn | xnt}, BEFORE the hi/lo split overwrites it
  int32_t* rows;                 // [l][cap] out: rows of the entries of column c, ascending
  float* vals;                   // [l][cap] out: xn - xnt there (exact: xnt is xn or 0)
  int32_t* col_count;            // [l] out
  int32_t* count;                // [3], see above
  int32_t cap;                   // list entries per column
};

__global__ void __launch_bounds__(THRLIST_THREADS)
glp_thr_list_kernel(const ThrListArgs a, const Geom g) {
  __shared__ int s_warp[THRLIST_THREADS / 32];
  __shared__ int s_carry;
  __shared__ bool s_last;
  pdl_wait();
  pdl_launch();
  if (a.ctrl != nullptr && a.ctrl->done) return;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c = blockIdx.x, l = g.l;
  const float* x = a.xop + int64_t(c) * g.ldn;
  const float* xz = a.xop + int64_t(l + c) * g.ldn;
  int32_t* rows = a.rows + int64_t(c) * a.cap;
  float* vals = a.vals + int64_t(c) * a.cap;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  // ordered compaction, 4 consecutive rows per thread and 4096 rows per round (padding rows are zero in both)
  for (int64_t j0 = 0; j0 < g.ldn; j0 += 4 * THRLIST_THREADS) {
    const int64_t j = j0 + 4 * tid;
    float4 u = make_float4(0.f, 0.f, 0.f, 0.f), v = u;
    if (j + 3 < g.ldn) {
      u = *reinterpret_cast<const float4*>(x + j);
      v = *reinterpret_cast<const float4*>(xz + j);
    } else {
      float* up = &u.x; float* vp = &v.x;
      for (int q = 0; q < 4; ++q) if (j + q < g.ldn) { up[q] = x[j + q]; vp[q] = xz[j + q]; }
    }
    const float d[4] = {u.x - v.x, u.y - v.y, u.z - v.z, u.w - v.w};
    const int mine = (d[0] != 0.f) + (d[1] != 0.f) + (d[2] != 0.f) + (d[3] != 0.f);
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int base = s_carry;
    for (int w = 0; w < warp; ++w) base += s_warp[w];
    int pos = base + incl - mine;
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (d[q] != 0.f) {
        if (pos < a.cap) { rows[pos] = int32_t(j + q); vals[pos] = d[q]; }
        ++pos;
      }
    __syncthreads();
    if (tid == THRLIST_THREADS - 1) s_carry = base + incl;
    __syncthreads();
  }
  if (tid == 0) {
    const int R = s_carry;
    a.col_count[c] = R;
    __threadfence();
    s_last = atomicAdd(reinterpret_cast<unsigned*>(&a.count[2]), 1u) == unsigned(l - 1);
  }
  __syncthreads();
  if (!s_last) return;
  // the last block to arrive totals the columns (fixed order) and re-arms the counter
  __threadfence();
  if (warp == 0) {
    int tot = 0, over = 0;
    for (int q = lane; q < l; q += 32) {
      const int r = __ldcg(&a.col_count[q]);
      tot += r;
      over += r > a.cap ? 1 : 0;
    }
    tot = __reduce_add_sync(0xffffffffu, tot);
    over = __reduce_add_sync(0xffffffffu, over);
    if (lane == 0) { a.count[0] = tot; a.count[1] = over; a.count[2] = 0; }
  }
}

struct ThrFixArgs {
  const Ctrl* ctrl;
  const int32_t* count;             // count[1] > 0: lists unusable, the gated product has computed A.xnt
  const int32_t* col_count;
  const int32_t* rows;
  const float* vals;
  int32_t cap;
  int32_t pad;
  const float* At_hi;               // [n][ldm] transposed A, TF32 part
  const float* At_lo;               // fp32 remainder (3xTF32 operand set) or nullptr
  const __nv_bfloat16* At_lb;       // bf16 remainder (TF32 + bf16 operand set) or nullptr
  const __half* At_f0;              // fp16-pair operand set: A^T s = f0 + f1 (then At_hi is nullptr)
  const __half* At_f1;
  float inv_s;                      // 1 / s
  float thres;                      // the hard threshold: every |vals| is below it
  const double* inv_x;              // fp16-pair set: 1 / (s s_x) of the current operand, max|xn| s_x in [2^14, 2^15)
  double* AX[2];                    // [2l][ldm]: column l + c := column c - sum_r At[rows_r][:] * vals_r
  int64_t tiled_nkb;                // > 0: At_f0 / At_f1 are stored as 128 x 64 tiles with this many K blocks (glp_gemm.cuh)
};

// grid (ceil(ldm / (128 ROWS)), l): block (ib, c) owns 128 ROWS output rows of column c, ROWS consecutive rows per
// thread.  The kernel is a gather of row segments of A^T from HBM (the entries of different columns sit in different
// rows), so what matters is bytes in flight: ROWS = 8 with 8 entries unrolled keeps 8 x 16 bytes (fp16) or 8 x 32
// bytes (fp32) of independent loads per thread outstanding.  Per output element the sum still runs over the entries
// in ascending row order with one fmaf each, so the result does not depend on ROWS or the unroll.
constexpr int THRFIX_ROWS = 8;
constexpr int THRFIX_UNROLL = 8;
__global__ void __launch_bounds__(THRFIX_THREADS)
glp_thr_fix_kernel(const ThrFixArgs a, const Geom g) {
  constexpr int RW = THRFIX_ROWS;
  __shared__ int32_t s_j[THRFIX_THREADS];
  __shared__ float s_v[THRFIX_THREADS];
  pdl_wait();
  pdl_launch();
  if (a.ctrl != nullptr && a.ctrl->done) return;
  if (a.count[1] > 0) return;
  const bool sel1 = a.ctrl != nullptr ? a.ctrl->cur == 0 : false;
  const int tid = threadIdx.x, c = blockIdx.y, l = g.l;
  const int R = a.col_count[c];
  const int64_t i = (int64_t(blockIdx.x) * THRFIX_THREADS + tid) * RW;
  const bool live = i < g.ldm;   // ldm is a multiple of 64: a live thread owns RW in-range rows
  const int32_t* rows = a.rows + int64_t(c) * a.cap;
  const float* vals = a.vals + int64_t(c) * a.cap;
  // fp32 arithmetic: the correction is at most thres / max|xn| of the product, so its own 2^-24 rounding is far
  // below the product's 2^-22; conversions to fp64 run at 1/8 of the FFMA rate and would dominate the kernel
  float acc[RW];
#pragma unroll
  for (int t = 0; t < RW; ++t) acc[t] = 0.f;
  // fp16 pairs: f1 adds 2^-11 (relative) to terms that are at most thres / max|xn| of the product's own; it is read
  // only when that ratio exceeds 2^-10, i.e. when leaving it out could reach the 2^-22 the product itself is good for
  // (max|xn| >= 2^14 / s_x = 2^14 inv_x s).  Halves the traffic of the correction at the default thres.
  const bool f16 = a.At_f0 != nullptr;
  const bool need_f1 = !(f16 && a.inv_x != nullptr && double(a.thres) * double(a.inv_s) <= 16.0 * __ldg(a.inv_x));
  auto h8 = [](const uint4& u, float (&v)[RW], bool add) {   // 8 fp16 values -> fp32 (set or add: both exact)
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[q]));
      if (add) { v[2 * q] += f.x; v[2 * q + 1] += f.y; } else { v[2 * q] = f.x; v[2 * q + 1] = f.y; }
    }
  };
  auto b8 = [](const uint4& u, float (&v)[RW]) {             // + 8 bf16 values
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[q]));
      v[2 * q] += f.x; v[2 * q + 1] += f.y;
    }
  };
  // RW values of row j of A^T (times s for the fp16 pairs) starting at output row i; f0 + f1 and hi + lo are exact in fp32
  auto at8 = [&](int64_t at, float (&v)[RW]) {
    if (f16) {
      h8(__ldg(reinterpret_cast<const uint4*>(a.At_f0 + at)), v, false);
      if (need_f1) h8(__ldg(reinterpret_cast<const uint4*>(a.At_f1 + at)), v, true);
      return;
    }
    const float4 h0 = __ldg(reinterpret_cast<const float4*>(a.At_hi + at));
    const float4 h1 = __ldg(reinterpret_cast<const float4*>(a.At_hi + at + 4));
    v[0] = h0.x; v[1] = h0.y; v[2] = h0.z; v[3] = h0.w; v[4] = h1.x; v[5] = h1.y; v[6] = h1.z; v[7] = h1.w;
    if (a.At_lo != nullptr) {
      const float4 q0 = __ldg(reinterpret_cast<const float4*>(a.At_lo + at));
      const float4 q1 = __ldg(reinterpret_cast<const float4*>(a.At_lo + at + 4));
      v[0] += q0.x; v[1] += q0.y; v[2] += q0.z; v[3] += q0.w; v[4] += q1.x; v[5] += q1.y; v[6] += q1.z; v[7] += q1.w;
    } else if (a.At_lb != nullptr) {
      b8(__ldg(reinterpret_cast<const uint4*>(a.At_lb + at)), v);
    }
  };
  // offset of the RW values (row j of A^T, output rows i .. i + RW - 1; i is a multiple of RW = 8, so they share a tile row)
  auto at_off = [&](int32_t j) {
    return a.tiled_nkb > 0 ? f16_tiled_off(int64_t(j), i, a.tiled_nkb) : int64_t(j) * g.ldm + i;
  };
  for (int r0 = 0; r0 < R; r0 += THRFIX_THREADS) {
    const int nr = R - r0 < THRFIX_THREADS ? R - r0 : THRFIX_THREADS;
    __syncthreads();
    if (tid < nr) { s_j[tid] = rows[r0 + tid]; s_v[tid] = vals[r0 + tid]; }
    __syncthreads();
    if (!live) continue;
    int r = 0;
    for (; r + THRFIX_UNROLL <= nr; r += THRFIX_UNROLL) {   // independent row segments in flight
      float av[THRFIX_UNROLL][RW];
#pragma unroll
      for (int q = 0; q < THRFIX_UNROLL; ++q) at8(at_off(s_j[r + q]), av[q]);
#pragma unroll
      for (int q = 0; q < THRFIX_UNROLL; ++q) {
        const float d = s_v[r + q];
#pragma unroll
        for (int t = 0; t < RW; ++t) acc[t] = fmaf(av[q][t], d, acc[t]);
      }
    }
    for (; r < nr; ++r) {
      float av[RW];
      at8(at_off(s_j[r]), av);
      const float d = s_v[r];
#pragma unroll
      for (int t = 0; t < RW; ++t) acc[t] = fmaf(av[t], d, acc[t]);
    }
  }
  if (!live) return;
  const double sc = f16 ? double(a.inv_s) : 1.0;   // a power of two
  double* ax = sel1 ? a.AX[1] : a.AX[0];
  const double* src = ax + int64_t(c) * g.ldm + i;
  double* dst = ax + int64_t(l + c) * g.ldm + i;
#pragma unroll
  for (int t = 0; t < RW; t += 2) {
    const double2 s2 = *reinterpret_cast<const double2*>(src + t);
    double2 d2;
    d2.x = s2.x - double(acc[t]) * sc;
    d2.y = s2.y - double(acc[t + 1]) * sc;
    *reinterpret_cast<double2*>(dst + t) = d2;
  }
}

}  // namespace glp
