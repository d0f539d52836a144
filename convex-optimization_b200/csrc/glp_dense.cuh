// glp_dense.cuh -- fp64 dense building blocks for gl_Admm_dual at streaming sizes.
//
// The reference factors  G = I + rho A A^T  once with scipy (code/gl_ADMM_dual.py:57) and applies two triangular
// solves per iteration (:61-63).  At m = 8192 the factor is 512 MB and a triangular solve is a chain of m dependent
// steps, so here the set-up ALSO forms  Y = L^-T  (the transposed inverse of the Cholesky factor, upper
// triangular) and every iteration applies  z = L^-T (L^-1 w) = Y (Y^T w)  as one transposed and one forward STREAM
// of Y through K2 / K1 -- the same kernels, the same roofline as the streams of A.  cond(L) = sqrt(cond(G)), so the
// explicit triangular inverse costs ~cond(L) eps ~ 1e-13 of accuracy, far inside the 1e-10 parity bar.
//
// Everything in the set-up is one tiled fp64 product  C = beta C + alpha A B^T  ("NT": both operands are read along
// their rows, so no transposed copy is ever needed) plus two single-CTA kernels on 128 x 128 diagonal blocks:
//   Gram              G   = rho A A^T (+ I)                      fp64 inputs (fp32 inputs take the tensor cores, K5)
//   Cholesky, blocked L_kk = chol(G_kk), W_k = L_kk^-1           dense_potrf_trtri_kernel
//                     L_ik = G_ik W_k^T            (i > k)       NT
//                     G_ij -= L_ik L_jk^T          (i >= j > k)  NT, lower tiles only
//   inverse           R       = -Y[<k, <k] L[k, <k]^T            NT   (rows above block k of (E_k - L_k,<k X_<k,:)^T)
//                     Y[<k, k] = R W_k^T ,  Y[k, k] = W_k^T      NT
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "glp_types.cuh"

namespace glp {

constexpr int DENSE_NB = 128;          // block size of the factorisation
constexpr int DENSE_TILE = 64;         // C tile per CTA
constexpr int DENSE_KT = 16;           // K step
constexpr int DENSE_THREADS = 256;     // 16 x 16 threads, 4 x 4 outputs each

// C[i][j] = beta C[i][j] + alpha sum_k A[i][k] B[j][k] + (add_eye && i == j ? 1 : 0),  i < M, j < N, k < K.
// lower_only: tiles strictly above the diagonal are skipped (M == N, same row offsets).  TA: float or double inputs,
// fp64 arithmetic.  C must not alias A or B.
template <typename TA>
__global__ void __launch_bounds__(DENSE_THREADS)
dense_gemm_nt_kernel(int64_t M, int64_t N, int64_t K, double alpha, const TA* __restrict__ A, int64_t lda,
                     const TA* __restrict__ B, int64_t ldb, double beta, double* __restrict__ C, int64_t ldc,
                     int lower_only, int add_eye) {
  __shared__ double sa[DENSE_KT][DENSE_TILE + 4], sb[DENSE_KT][DENSE_TILE + 4];
  const int64_t ti = blockIdx.y, tj = blockIdx.x;
  if (lower_only && tj > ti) return;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t i0 = ti * DENSE_TILE, j0 = tj * DENSE_TILE;
  double acc[4][4];
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int v = 0; v < 4; ++v) acc[u][v] = 0.0;
  // loader: thread -> (row r = tid / 4 of the 64-row strip, 4 consecutive k at kq = (tid % 4) * 4)
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  for (int64_t k0 = 0; k0 < K; k0 += DENSE_KT) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int64_t k = k0 + lk + q;
      const int64_t ia = i0 + lr, jb = j0 + lr;
      sa[lk + q][lr] = (ia < M && k < K) ? double(A[ia * lda + k]) : 0.0;
      sb[lk + q][lr] = (jb < N && k < K) ? double(B[jb * ldb + k]) : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < DENSE_KT; ++kk) {
      double av[4], bv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) av[u] = sa[kk][ty * 4 + u];
#pragma unroll
      for (int v = 0; v < 4; ++v) bv[v] = sb[kk][tx * 4 + v];
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) acc[u][v] = fma(av[u], bv[v], acc[u][v]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int64_t i = i0 + ty * 4 + u;
    if (i >= M) continue;
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      const int64_t j = j0 + tx * 4 + v;
      if (j >= N) continue;
      double c = alpha * acc[u][v];
      if (beta != 0.0) c += beta * C[i * ldc + j];
      if (add_eye && i == j) c += 1.0;
      C[i * ldc + j] = c;
    }
  }
}

// One CTA, one diagonal block of order nb <= 128 held in shared memory:
//   Gkk (row-major, ld) -> Lkk = chol(Gkk) written to Lout (lower triangle, zeros above), W = Lkk^-1 written to Wout
//   ([nb][128] row-major, lower triangle, zeros above).  A non-positive pivot sets *fault = 3.
__global__ void __launch_bounds__(512)
dense_potrf_trtri_kernel(const double* __restrict__ Gkk, int64_t ldg, int nb, double* __restrict__ Lout, int64_t ldl,
                         double* __restrict__ Wout, int32_t* fault) {
  extern __shared__ double sm[];          // [128][129]: lower triangle = the block / its factor, strictly upper = W^T
  double* S = sm;
  double* wdiag = sm + 128 * 129;         // [128] diagonal of W
  const int tid = threadIdx.x;
  for (int e = tid; e < 128 * 128; e += 512) {
    const int i = e >> 7, j = e & 127;
    S[i * 129 + j] = (i < nb && j <= i) ? Gkk[int64_t(i) * ldg + j] : 0.0;
  }
  __syncthreads();
  // right-looking Cholesky by columns
  for (int j = 0; j < nb; ++j) {
    const double d = S[j * 129 + j];
    if (!(d > 0.0)) {
      if (tid == 0 && fault) atomicExch(fault, 3);
      return;
    }
    const double ljj = sqrt(d);
    __syncthreads();
    for (int i = j + tid; i < nb; i += 512) S[i * 129 + j] = (i == j) ? ljj : S[i * 129 + j] / ljj;
    __syncthreads();
    const int r = nb - j - 1;
    for (int e = tid; e < r * r; e += 512) {
      const int ii = e / r, kk = e - ii * r;
      if (kk <= ii) {
        const int i = j + 1 + ii, k = j + 1 + kk;
        S[i * 129 + k] -= S[i * 129 + j] * S[k * 129 + j];
      }
    }
    __syncthreads();
  }
  // W = L^-1 by forward substitution, one thread per column c (W[:, c] solves L w = e_c; the columns are independent).
  // W[i][c], i > c, is kept transposed in the unused upper triangle (S[c][i]); thread c only writes row c there.
  if (tid < nb) {
    const int c = tid;
    const double wd = 1.0 / S[c * 129 + c];
    wdiag[c] = wd;
    for (int i = c + 1; i < nb; ++i) {
      double s = -S[i * 129 + c] * wd;
      for (int k = c + 1; k < i; ++k) s -= S[i * 129 + k] * S[c * 129 + k];
      S[c * 129 + i] = s / S[i * 129 + i];
    }
  }
  __syncthreads();
  for (int e = tid; e < nb * 128; e += 512) {
    const int i = e >> 7, j = e & 127;
    if (j < nb) Lout[int64_t(i) * ldl + j] = (j <= i) ? S[i * 129 + j] : 0.0;
    Wout[i * 128 + j] = (j < nb && j < i) ? S[j * 129 + i] : ((j == i) ? wdiag[i] : 0.0);
  }
}

// planar fp64 [l][ld] -> planar dtype T [l][ld]  (z of the inner streams -> the operand of the transposed stream of A)
template <typename T>
__global__ void dense_cast_planar_kernel(const double* __restrict__ src, T* __restrict__ dst, int64_t total,
                                         const Ctrl* ctrl) {
  if (ctrl != nullptr && ctrl->done) return;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x)
    dst[e] = T(src[e]);
}

// D (planar fp64 [cols][ldd], the layout K5 writes) -> G[row][col0 + col] = alpha D[col][row] + (row == col0 + col)
__global__ void dense_from_planar_kernel(const double* __restrict__ D, int64_t ldd, int64_t rows, int cols, int64_t col0,
                                         double alpha, double* __restrict__ G, int64_t ldg) {
  const int64_t total = rows * cols;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < total; e += int64_t(gridDim.x) * blockDim.x) {
    const int c = int(e / rows);
    const int64_t r = e - int64_t(c) * rows;
    G[r * ldg + col0 + c] = alpha * D[int64_t(c) * ldd + r] + (r == col0 + c ? 1.0 : 0.0);
  }
}

}  // namespace glp
