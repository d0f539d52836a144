// glp_misc.cuh -- small layout / data kernels around the hot path: row-major <-> column-planar
// conversion of the O((m+n) l) vectors, and the counter-based N(0,1) generator used for synthetic
// problems that are too large to draw with NumPy's MT19937 (code/main.py:37-51 restated at scale).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>
#include <math.h>

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
//
// Box-Muller on two 53-bit uniforms, written with IEEE-754 basic operations only (+, *, /, sqrt in a fixed
// order, never contracted to FMA; log and cos(2 pi u) by argument reduction in integer arithmetic and fixed
// Horner polynomials), so that the CPU restatement (oracle/glprox_gen.c, oracle/gen.py) reproduces every
// value bit for bit: the CPU reference arm of bench.py then works on the very same problem without
// loading this library.
__host__ __device__ inline uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ inline double rn_mul(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dmul_rn(a, b);
#else
  volatile double r = a * b; return r;
#endif
}
__host__ __device__ inline double rn_add(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dadd_rn(a, b);
#else
  volatile double r = a + b; return r;
#endif
}
__host__ __device__ inline double rn_div(double a, double b) {
#ifdef __CUDA_ARCH__
  return __ddiv_rn(a, b);
#else
  volatile double r = a / b; return r;
#endif
}
__host__ __device__ inline double rn_sqrt(double a) {
#ifdef __CUDA_ARCH__
  return __dsqrt_rn(a);
#else
  return sqrt(a);
#endif
}
__host__ __device__ inline double bits_to_double(uint64_t b) {
#ifdef __CUDA_ARCH__
  return __longlong_as_double((long long)b);
#else
  double d; memcpy(&d, &b, 8); return d;
#endif
}
__host__ __device__ inline uint64_t double_to_bits(double d) {
#ifdef __CUDA_ARCH__
  return (uint64_t)__double_as_longlong(d);
#else
  uint64_t b; memcpy(&b, &d, 8); return b;
#endif
}
// log(u) for a normal u in (0, 1]: u = mant * 2^e, mant in [sqrt(1/2), sqrt(2)], log(mant) = 2 atanh(s),
// s = (mant - 1) / (mant + 1), |s| <= 0.1716: 11 odd terms (truncation < 1e-18)
__host__ __device__ inline double gen_log(double u) {
  const uint64_t bits = double_to_bits(u);
  int e = int(bits >> 52) - 1023;
  double mant = bits_to_double((bits & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull);
  if (mant > 1.4142135623730951) { mant = rn_mul(mant, 0.5); e += 1; }
  const double s = rn_div(rn_add(mant, -1.0), rn_add(mant, 1.0));
  const double z = rn_mul(s, s);
  double p = 0.047619047619047616;
  p = rn_add(rn_mul(p, z), 0.05263157894736842);
  p = rn_add(rn_mul(p, z), 0.058823529411764705);
  p = rn_add(rn_mul(p, z), 0.06666666666666667);
  p = rn_add(rn_mul(p, z), 0.07692307692307693);
  p = rn_add(rn_mul(p, z), 0.09090909090909091);
  p = rn_add(rn_mul(p, z), 0.1111111111111111);
  p = rn_add(rn_mul(p, z), 0.14285714285714285);
  p = rn_add(rn_mul(p, z), 0.2);
  p = rn_add(rn_mul(p, z), 0.3333333333333333);
  p = rn_add(rn_mul(p, z), 1.0);
  return rn_add(rn_mul(double(e), 0.6931471805599453), rn_mul(rn_mul(2.0, s), p));
}
// cos(2 pi N / 2^53) for an integer N in [0, 2^53): nearest quarter turn and remainder in integers (exact),
// then the Taylor polynomial of cos or sin on [-pi/4, pi/4]
__host__ __device__ inline double gen_cos2pi(uint64_t N) {
  const uint64_t kq = (N + (1ull << 50)) >> 51;                      // 0..4
  const int64_t rem = int64_t(N) - int64_t(kq << 51);                 // |rem| <= 2^50
  const double th = rn_mul(rn_mul(double(rem), 1.1102230246251565e-16), 6.283185307179586);   // 2^-53, 2 pi
  const double t2 = rn_mul(th, th);
  double v;
  if ((kq & 1) == 0) {
    double c = -1.5619206968586225e-16;
    c = rn_add(rn_mul(c, t2), 4.779477332387385e-14);
    c = rn_add(rn_mul(c, t2), -1.1470745597729725e-11);
    c = rn_add(rn_mul(c, t2), 2.08767569878681e-09);
    c = rn_add(rn_mul(c, t2), -2.755731922398589e-07);
    c = rn_add(rn_mul(c, t2), 2.48015873015873e-05);
    c = rn_add(rn_mul(c, t2), -0.001388888888888889);
    c = rn_add(rn_mul(c, t2), 0.041666666666666664);
    c = rn_add(rn_mul(c, t2), -0.5);
    v = rn_add(rn_mul(c, t2), 1.0);
  } else {
    double sp = 2.8114572543455206e-15;
    sp = rn_add(rn_mul(sp, t2), -7.647163731819816e-13);
    sp = rn_add(rn_mul(sp, t2), 1.6059043836821613e-10);
    sp = rn_add(rn_mul(sp, t2), -2.505210838544172e-08);
    sp = rn_add(rn_mul(sp, t2), 2.7557319223985893e-06);
    sp = rn_add(rn_mul(sp, t2), -0.0001984126984126984);
    sp = rn_add(rn_mul(sp, t2), 0.008333333333333333);
    sp = rn_add(rn_mul(sp, t2), -0.16666666666666666);
    sp = rn_add(rn_mul(sp, t2), 1.0);
    v = rn_mul(th, sp);
  }
  // quarter turns: cos, -sin, -cos, sin, cos
  return (kq == 1 || kq == 2) ? -v : v;
}
__host__ __device__ inline double randn_at(uint64_t seed, uint64_t row, uint64_t col) {
  const uint64_t h = mix64(mix64(seed ^ (row * 0xD1342543DE82EF95ull)) + col);
  const uint64_t h2 = mix64(h ^ 0xA0761D6478BD642Full);
  const double u1 = rn_mul(double((h >> 11) + 1), 1.1102230246251565e-16);   // (0, 1], exact
  const double radius = rn_sqrt(rn_mul(-2.0, gen_log(u1)));
  return rn_mul(radius, gen_cos2pi(h2 >> 11));
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

// b = A u of the synthetic recipe (code/main.py:45) with fp64 accumulation over the non-zero row groups of u in
// ascending column order, one rounded multiply and one rounded add per term: the order and the arithmetic of the CPU
// restatement (oracle/glprox_gen.c: glp_oracle_matmul), so that both sides build the same right-hand side bit for bit.
// One thread per row of A; nz[] lists the non-zero rows of u.  Set-up only (the rows are read with a stride).
template <typename T>
__global__ void rhs_exact_kernel(const T* __restrict__ A, int64_t m, int64_t lda, const double* __restrict__ u, int l,
                                 const int64_t* __restrict__ nz, int64_t nnz, double* __restrict__ y) {
  const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (i >= m) return;
  for (int c0 = 0; c0 < l; c0 += 8) {
    double acc[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    const int cw = l - c0 < 8 ? l - c0 : 8;
    for (int64_t q = 0; q < nnz; ++q) {
      const int64_t j = nz[q];
      const double v = double(A[i * lda + j]);
      for (int c = 0; c < cw; ++c) acc[c] = rn_add(acc[c], rn_mul(v, u[j * l + c0 + c]));
    }
    for (int c = 0; c < cw; ++c) y[i * l + c0 + c] = acc[c];
  }
}

}  // namespace glp
