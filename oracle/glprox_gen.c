/* oracle/glprox_gen.c -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement of the synthetic-problem generator used at sizes NumPy's MT19937 cannot draw in
 * reasonable time: the recipe of the reference's gen_data (code/main.py:37-51: A ~ N(0,1), b = A u)
 * with a counter-based N(0,1) stream, value = f(seed, global row, global column).  The product's
 * device generator (convex-optimization_b200/csrc/glp_misc.cuh: randn_kernel) computes the same
 * function with the same IEEE-754 basic operations in the same order, so the two agree bit for bit
 * (tests/test_gpu_kernels.py::test_device_generator_matches_cpu_restatement); this lets the CPU
 * reference arm of bench.py work on the identical problem without loading the product library.
 *
 * Only tests/, __graft_entry__.smoke() and the CPU legs of bench.py may load this.
 * Build: gcc -O2 -fopenmp -ffp-contract=off -fPIC -shared (oracle/build_oracle.py).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

static inline uint64_t mix64(uint64_t z) { /* splitmix64 finaliser */
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

/* log(u), u normal in (0,1]: u = mant 2^e with mant in [sqrt(1/2), sqrt(2)]; log(mant) = 2 atanh(s) */
static inline double gen_log(double u) {
  uint64_t bits;
  memcpy(&bits, &u, 8);
  int e = (int)(bits >> 52) - 1023;
  uint64_t mb = (bits & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull;
  double mant;
  memcpy(&mant, &mb, 8);
  if (mant > 1.4142135623730951) { mant = mant * 0.5; e += 1; }
  const double s = (mant + -1.0) / (mant + 1.0);
  const double z = s * s;
  static const double c[11] = {1.0, 0.3333333333333333, 0.2, 0.14285714285714285, 0.1111111111111111,
                               0.09090909090909091, 0.07692307692307693, 0.06666666666666667,
                               0.058823529411764705, 0.05263157894736842, 0.047619047619047616};
  double p = c[10];
  for (int k = 9; k >= 0; --k) p = p * z + c[k];   /* -ffp-contract=off: one multiply, one add */
  return (double)e * 0.6931471805599453 + (2.0 * s) * p;
}

/* cos(2 pi N / 2^53), N integer in [0, 2^53) */
static inline double gen_cos2pi(uint64_t N) {
  const uint64_t kq = (N + (1ull << 50)) >> 51;
  const int64_t rem = (int64_t)N - (int64_t)(kq << 51);
  const double th = ((double)rem * 1.1102230246251565e-16) * 6.283185307179586;
  const double t2 = th * th;
  double v;
  if ((kq & 1) == 0) {
    static const double c[10] = {1.0, -0.5, 0.041666666666666664, -0.001388888888888889, 2.48015873015873e-05,
                                 -2.755731922398589e-07, 2.08767569878681e-09, -1.1470745597729725e-11,
                                 4.779477332387385e-14, -1.5619206968586225e-16};
    double q = c[9];
    for (int k = 8; k >= 0; --k) q = q * t2 + c[k];
    v = q;
  } else {
    static const double c[9] = {1.0, -0.16666666666666666, 0.008333333333333333, -0.0001984126984126984,
                                2.7557319223985893e-06, -2.505210838544172e-08, 1.6059043836821613e-10,
                                -7.647163731819816e-13, 2.8114572543455206e-15};
    double q = c[8];
    for (int k = 7; k >= 0; --k) q = q * t2 + c[k];
    v = th * q;
  }
  return (kq == 1 || kq == 2) ? -v : v;
}

static inline double randn_at(uint64_t seed, uint64_t row, uint64_t col) {
  const uint64_t h = mix64(mix64(seed ^ (row * 0xD1342543DE82EF95ull)) + col);
  const uint64_t h2 = mix64(h ^ 0xA0761D6478BD642Full);
  const double u1 = (double)((h >> 11) + 1) * 1.1102230246251565e-16;
  const double radius = sqrt(-2.0 * gen_log(u1));
  return radius * gen_cos2pi(h2 >> 11);
}

/* dst[i*ld + j] = N(0,1)(seed, row0 + i, col0 + j); dtype 0 = float32, 1 = float64 */
int glp_oracle_randn(void* dst, int dtype, int64_t rows, int64_t cols, int64_t ld, uint64_t seed, int64_t row0,
                     int64_t col0) {
  if (!dst || rows < 0 || cols < 0 || ld < cols || (dtype != 0 && dtype != 1)) return -1;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < rows; ++i) {
    if (dtype == 0) {
      float* d = (float*)dst + i * ld;
      for (int64_t j = 0; j < cols; ++j) d[j] = (float)randn_at(seed, (uint64_t)(row0 + i), (uint64_t)(col0 + j));
    } else {
      double* d = (double*)dst + i * ld;
      for (int64_t j = 0; j < cols; ++j) d[j] = randn_at(seed, (uint64_t)(row0 + i), (uint64_t)(col0 + j));
    }
  }
  return 0;
}

/* y[m][l] (fp64, row-major) = A[m][lda] (float32 or float64) * x[n][l] (fp64), fp64 accumulation in
 * ascending column order, zero rows of x skipped (u of the recipe has 10 % non-zero row groups):
 * the b = A u of the recipe without a 2x host copy of A. */
#include <stdlib.h>
int glp_oracle_matmul(const void* A, int dtype, int64_t m, int64_t n, int64_t lda, const double* x, int l, double* y) {
  if (!A || !x || !y || l < 1 || l > 256) return -1;
  int64_t* nz = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
  if (!nz) return -2;
  int64_t cnt = 0;
  for (int64_t j = 0; j < n; ++j) {
    int any = 0;
    for (int c = 0; c < l; ++c) any |= (x[j * l + c] != 0.0);
    if (any) nz[cnt++] = j;
  }
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < m; ++i) {
    double acc[256];
    for (int c = 0; c < l; ++c) acc[c] = 0.0;
    for (int64_t q = 0; q < cnt; ++q) {
      const int64_t j = nz[q];
      const double v = dtype == 0 ? (double)((const float*)A)[i * lda + j] : ((const double*)A)[i * lda + j];
      for (int c = 0; c < l; ++c) acc[c] += v * x[j * l + c];
    }
    for (int c = 0; c < l; ++c) y[i * l + c] = acc[c];
  }
  free(nz);
  return 0;
}
