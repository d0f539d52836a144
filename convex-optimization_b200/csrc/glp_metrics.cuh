// glp_metrics.cuh -- the caller-side metrics of the reference driver evaluated where x lives.
//
// code/main.py:48-50 defines, on host arrays,
//   errfun(x1, x2)  = ||x1 - x2||_F / (1 + ||x1||_F)
//   errfun_exact(x) = ||x - u||_F   / (1 + ||x||_F)
//   sparsity(x)     = #{ |x| > 1e-6 * max|x| } / (n l)
// and code/main.py:30-34 the objective with the caller's mu.  A benchmark sweep that keeps x in HBM
// needs none of x on the host: two small reductions over the n x l iterate (row-major, the layout
// the solvers return) give every term.  Reductions are fixed-order (thread-strided partials, warp
// tree, warps in order, CTAs in order), so the numbers are run-to-run identical.
#pragma once
#include "glp_device.cuh"

namespace glp {

constexpr int MET_THREADS = 256;
constexpr int MET_STRIDE = 8;  // per-CTA record {sumsq_x, sumsq_ref, sumsq_diff, reg, max|x|, count, -, -}

__device__ __forceinline__ double met_nmax(double a, double b) {  // NaN-propagating, like np.max
  return (a != a) ? a : ((b != b) ? b : fmax(a, b));
}

// CTA-wide fixed-order reduction of one value; result valid on thread 0.
__device__ __forceinline__ double met_block_reduce(double v, bool is_max, double* s_w) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double y = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? met_nmax(v, y) : v + y;
  }
  __syncthreads();  // s_w may still be read from the previous call
  if (lane == 0) s_w[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0) {
    t = s_w[0];
    for (int w = 1; w < MET_THREADS / 32; ++w) t = is_max ? met_nmax(t, s_w[w]) : t + s_w[w];
  }
  return t;
}

// pass 1: one thread per row group; part[blockIdx][MET_STRIDE]
template <typename T>
__global__ void __launch_bounds__(MET_THREADS)
glp_metrics_sums_kernel(const T* __restrict__ x, const T* __restrict__ ref, int64_t n, int l,
                        double* __restrict__ part) {
  __shared__ double s_w[MET_THREADS / 32];
  double sx = 0.0, sr = 0.0, sd = 0.0, reg = 0.0, mx = 0.0;
  for (int64_t i = int64_t(blockIdx.x) * MET_THREADS + threadIdx.x; i < n; i += int64_t(gridDim.x) * MET_THREADS) {
    double rn2 = 0.0;
    for (int c = 0; c < l; ++c) {
      const double v = double(x[i * l + c]);
      rn2 += v * v;
      mx = met_nmax(mx, fabs(v));
      if (ref != nullptr) {
        const double u = double(ref[i * l + c]);
        sr += u * u;
        sd += (v - u) * (v - u);
      }
    }
    sx += rn2;
    reg += sqrt(rn2);
  }
  const double a0 = met_block_reduce(sx, false, s_w);
  const double a1 = met_block_reduce(sr, false, s_w);
  const double a2 = met_block_reduce(sd, false, s_w);
  const double a3 = met_block_reduce(reg, false, s_w);
  const double a4 = met_block_reduce(mx, true, s_w);
  if (threadIdx.x == 0) {
    double* p = part + int64_t(blockIdx.x) * MET_STRIDE;
    p[0] = a0; p[1] = a1; p[2] = a2; p[3] = a3; p[4] = a4;
  }
}

// pass 2: every CTA re-derives the global max from the pass-1 records (or takes the caller's, for a
// shard whose max was reduced across ranks) and counts its elements above the cut.
template <typename T>
__global__ void __launch_bounds__(MET_THREADS)
glp_metrics_count_kernel(const T* __restrict__ x, int64_t total, double* __restrict__ part, int nparts,
                         double max_override) {
  __shared__ double s_w[MET_THREADS / 32];
  __shared__ double s_cut;
  double mx = 0.0;
  if (max_override >= 0.0) {
    mx = max_override;
  } else {
    for (int i = threadIdx.x; i < nparts; i += MET_THREADS) mx = met_nmax(mx, part[int64_t(i) * MET_STRIDE + 4]);
    mx = met_block_reduce(mx, true, s_w);
  }
  if (threadIdx.x == 0) s_cut = xmul(1e-6, mx);
  __syncthreads();
  const double cut = s_cut;
  double cnt = 0.0;
  for (int64_t e = int64_t(blockIdx.x) * MET_THREADS + threadIdx.x; e < total; e += int64_t(gridDim.x) * MET_THREADS)
    cnt += (fabs(double(x[e])) > cut) ? 1.0 : 0.0;
  const double c = met_block_reduce(cnt, false, s_w);
  if (threadIdx.x == 0) part[int64_t(blockIdx.x) * MET_STRIDE + 5] = c;
}

// pass 3: one CTA folds the records in CTA order -> out[6] = {sumsq_x, sumsq_ref, sumsq_diff, reg, max, count}
__global__ void __launch_bounds__(MET_THREADS)
glp_metrics_fold_kernel(const double* __restrict__ part, int nparts, double* __restrict__ out) {
  __shared__ double s_w[MET_THREADS / 32];
  for (int f = 0; f < 6; ++f) {
    const bool is_max = f == 4;
    double v = 0.0;
    for (int i = threadIdx.x; i < nparts; i += MET_THREADS) {
      const double y = part[int64_t(i) * MET_STRIDE + f];
      v = is_max ? met_nmax(v, y) : v + y;
    }
    const double t = met_block_reduce(v, is_max, s_w);
    if (threadIdx.x == 0) out[f] = t;
  }
}

}  // namespace glp
