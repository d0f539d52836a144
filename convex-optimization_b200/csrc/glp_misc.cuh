// glp_misc.cuh -- small layout / data kernels around the hot path: row-major <-> column-planar
// conversion of the O((m+n) l) vectors, and the counter-based N(0,1) generator used for synthetic
// problems that are too large to draw with NumPy's MT19937 (code/main.py:37-51 restated at scale).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace glp {

// dst[c*ld + i] = src[i*cols + c]   (row-major [rows][cols] of S  ->  planar [cols][ld] of D)
template <typename S, typename D>
__global__ void to_planar_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t rows, int cols,
                                 int64_t ld) {
  const int64_t total = rows * cols;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = e / cols;
    const int c = int(e % cols);
    dst[c * ld + i] = D(src[e]);
  }
}

// dst[i*cols + c] = src[c*ld + i]
template <typename S, typename D>
__global__ void from_planar_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t rows, int cols,
                                   int64_t ld) {
  const int64_t total = rows * cols;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = e / cols;
    const int c = int(e % cols);
    dst[e] = D(src[c * ld + i]);
  }
}

// ---- counter-based generator: value = f(seed, global_row, global_col); any block of the same
// logical matrix can be produced independently on any rank (column shards of A).
__host__ __device__ inline uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ inline double randn_at(uint64_t seed, uint64_t row, uint64_t col) {
  const uint64_t h = mix64(mix64(seed ^ (row * 0xD1342543DE82EF95ull)) + col);
  const uint64_t h2 = mix64(h ^ 0xA0761D6478BD642Full);
  const double u1 = (double((h >> 11) + 1)) * (1.0 / 9007199254740993.0);  // (0,1]
  const double u2 = double(h2 >> 11) * (1.0 / 9007199254740992.0);           // [0,1)
  return sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
}

template <typename T>
__global__ void randn_kernel(T* __restrict__ dst, int64_t rows, int64_t cols, int64_t ld, uint64_t seed,
                             int64_t row0, int64_t col0) {
  const int64_t total = rows * cols;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = e / cols, j = e % cols;
    dst[i * ld + j] = T(randn_at(seed, uint64_t(row0 + i), uint64_t(col0 + j)));
  }
}

}  // namespace glp
