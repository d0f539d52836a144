// glp_admm.cuh -- SURVEY.md 8(f) row 4: gl_Admm_dual (code/gl_ADMM_dual.py:10-103) on the two stream kernels.
//
// One iteration of the reference is
//   z = (I + rho A A^T)^-1 (A (x - rho u) - b)      :63   Cholesky factor L computed once (:57)
//   u' = P_mu(x / rho - A^T z)                      :64   row-wise projection onto the mu-ball (:45-47)
//   x' = x - tau rho (u' + A^T z)                   :65
//   r = u' + A^T z ,  s = A (u - u')                :67-68  residuals, judged by their matrix 2-norms (:87-88)
//   f = 0.5 ||A x' - b||^2 + mu sum_i ||x'_i||      :72
// Here: ONE forward stream (K1) over the 2l operand columns [x - rho u | u] gives A(x - rho u) and A u; by
// linearity A x = A(x - rho u) + rho A u and s = A u_prev - A u, so the objective, the sparsity and both
// residual norms of iteration k are evaluated after the forward stream of iteration k+1 (kernel `mid`,
// which also does the two triangular solves with L in one CTA); ONE transposed stream (K2) gives A^T z and
// `update` applies the row-local projection / multiplier step.  A is streamed exactly twice per iteration,
// the control (k, the converge_len counter, histories) lives on the device.
// Scope of THIS file: the kernels of the iteration and the single-CTA factor / solves used for m <= 2048; larger
// systems take the blocked set-up of glp_dense.cuh and apply the factor as two streams of Y = L^-T
// (glp_admm_host.inl).  l <= 8: the residual 2-norms come from the l x l Gram matrices, in closed form for l <= 2,
// by Jacobi rotations above.
#pragma once
#include "glp_decide.h"
#include "glp_device.cuh"
#include "glp_types.cuh"

namespace glp {

constexpr int ADMM_MAX_M = 2048;
constexpr int ADMM_THREADS = 1024;
constexpr int ADMM_MAX_L = 8;
constexpr int ADMM_NPAIR = ADMM_MAX_L * (ADMM_MAX_L + 1) / 2;   // 36 entries of an upper triangle
// per-column-block record of the update kernel: [0, 36) Gram matrix of r = u + A^T z (upper triangle, row-major:
// entry (a, b), a <= b, at a*l - a(a-1)/2 + b - a), [36] sum_i ||x_i||, [37] max|x|
constexpr int ADMM_GSTRIDE = 40;
__host__ __device__ inline int admm_pair(int a, int b, int l) { return a * l - a * (a - 1) / 2 + (b - a); }

struct AdmmConst {
  int32_t maxit, converge_len, hist_cap, pad;
  double thres, tau, rho, mu;
};

// ---- G = I + rho A A^T (lower triangle and diagonal), one 32 x 32 tile per CTA ------------------------------
template <typename T>
__global__ void __launch_bounds__(1024) glp_admm_gram_kernel(const T* __restrict__ A, int64_t lda, int m, int n,
                                                             double rho, double* __restrict__ G) {
  __shared__ double sa[32][33], sb[32][33];
  const int bi = blockIdx.y, bj = blockIdx.x;
  if (bj > bi) return;
  const int ty = threadIdx.y, tx = threadIdx.x;
  const int i = bi * 32 + ty, j = bj * 32 + tx;
  double acc = 0.0;
  for (int k0 = 0; k0 < n; k0 += 32) {
    const int k = k0 + tx;
    sa[ty][tx] = (bi * 32 + ty < m && k < n) ? double(A[int64_t(bi * 32 + ty) * lda + k]) : 0.0;
    sb[ty][tx] = (bj * 32 + ty < m && k < n) ? double(A[int64_t(bj * 32 + ty) * lda + k]) : 0.0;
    __syncthreads();
#pragma unroll 8
    for (int q = 0; q < 32; ++q) acc = fma(sa[ty][q], sb[tx][q], acc);
    __syncthreads();
  }
  if (i < m && j < m && j <= i) G[int64_t(i) * m + j] = rho * acc + (i == j ? 1.0 : 0.0);
}

// ---- in-place lower Cholesky G = L L^T, one CTA, right-looking by columns ---------------------------------
__global__ void __launch_bounds__(ADMM_THREADS) glp_admm_chol_kernel(double* __restrict__ G, int m, int32_t* fault) {
  const int tid = threadIdx.x;
  for (int j = 0; j < m; ++j) {
    const double d = G[int64_t(j) * m + j];
    if (!(d > 0.0)) {  // not positive definite (NaN input): report instead of spinning on garbage
      if (tid == 0 && fault) atomicExch(fault, 3);
      return;
    }
    const double ljj = sqrt(d);
    __syncthreads();
    for (int i = j + tid; i < m; i += ADMM_THREADS) G[int64_t(i) * m + j] = (i == j) ? ljj : G[int64_t(i) * m + j] / ljj;
    __syncthreads();
    // trailing update of the lower triangle: G[i][k] -= L[i][j] L[k][j], j < k <= i
    const int r = m - j - 1;
    for (int64_t e = tid; e < int64_t(r) * r; e += ADMM_THREADS) {
      const int ii = int(e / r), kk = int(e - int64_t(ii) * r);
      if (kk <= ii) {
        const int i = j + 1 + ii, k = j + 1 + kk;
        G[int64_t(i) * m + k] -= G[int64_t(i) * m + j] * G[int64_t(k) * m + j];
      }
    }
    __syncthreads();
  }
}

// ---- row-local update (also the initial pass): u' , x' , operands, partial sums per column block -----------
struct AdmmUpdArgs {
  const Ctrl* ctrl;        // done => no-op
  int32_t init;            // 1: x = x0, u = 0: only operands and partials
  int32_t pad;
  double* x;               // [l][ldn]
  double* u;               // [l][ldn]
  const double* grad;      // [l][ldn] = A^T z
  void* xop;               // [2l][ldn] dtype T: [x - rho u | u]
  double* gram_partial;    // [ncb][ADMM_GSTRIDE]: Gram of r (upper triangle), sum_i ||x_i||, max|x|
  double* scal_partial;    // [ncb][SCAL_STRIDE]: slot 3 = max|x| of the block, where K1's sparsity count reads it
  AdmmConst c;
};

template <typename T>
__global__ void glp_admm_update_kernel(const AdmmUpdArgs a, const Geom g) {
  __shared__ double s_red[8][ADMM_NPAIR + 2];
  if (a.ctrl != nullptr && a.ctrl->done) return;
  constexpr int TC = Elem<T>::TILE_COLS;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t j = int64_t(blockIdx.x) * TC + tid;
  const int l = g.l;
  const int64_t ldn = g.ldn;
  T* xop = static_cast<T*>(a.xop);
  double gram[ADMM_NPAIR], reg = 0.0, mx = 0.0;
#pragma unroll
  for (int q = 0; q < ADMM_NPAIR; ++q) gram[q] = 0.0;
  if (tid < TC && j < g.n) {
    const double taurho = xmul(a.c.tau, a.c.rho);
    double nn = 0.0, xv[ADMM_MAX_L], gv[ADMM_MAX_L], vv[ADMM_MAX_L], rr[ADMM_MAX_L];
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c) { xv[c] = 0.0; gv[c] = 0.0; vv[c] = 0.0; rr[c] = 0.0; }
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c) {
      if (c >= l) break;
      xv[c] = a.x[c * ldn + j];
      if (!a.init) {
        gv[c] = __ldcg(&a.grad[c * ldn + j]);
        vv[c] = xsub(xdiv(xv[c], a.c.rho), gv[c]);   // x_k / rho - A.T @ z                   :64
        nn = xadd(nn, xmul(vv[c], vv[c]));
      }
    }
    double n2 = 0.0;
    if (!a.init) {
      const double rn = __dsqrt_rn(nn);
      const double den = rn < a.c.mu ? a.c.mu : rn;    // np.clip(row_norms, a_min=mu)              :47
#pragma unroll
      for (int c = 0; c < ADMM_MAX_L; ++c) {
        if (c >= l) break;
        const double un = xdiv(xmul(a.c.mu, vv[c]), den);
        rr[c] = xadd(un, gv[c]);                       // r_k = u + A.T @ z                        :67
        const double xn = xsub(xv[c], xmul(taurho, rr[c]));   // x_k - tau * rho * (u + A.T @ z)   :65
        a.x[c * ldn + j] = xn;
        a.u[c * ldn + j] = un;
        xop[c * ldn + j] = T(xsub(xn, xmul(a.c.rho, un)));
        xop[(l + c) * ldn + j] = T(un);
        n2 = xadd(n2, xmul(xn, xn));
        mx = fmax(mx, fabs(xn));
      }
    } else {
#pragma unroll
      for (int c = 0; c < ADMM_MAX_L; ++c) {
        if (c >= l) break;
        a.u[c * ldn + j] = 0.0;
        xop[c * ldn + j] = T(xv[c]);
        xop[(l + c) * ldn + j] = T(0.0);
        n2 = xadd(n2, xmul(xv[c], xv[c]));
        mx = fmax(mx, fabs(xv[c]));
      }
    }
    reg = __dsqrt_rn(n2);
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c)
#pragma unroll
      for (int d = c; d < ADMM_MAX_L; ++d)
        gram[admm_pair(c, d, ADMM_MAX_L)] = rr[c] * rr[d];   // compile-time index (rows / columns >= l are zero)
  }
#pragma unroll
  for (int q = 0; q < ADMM_NPAIR; ++q) {
    const double v = warp_sum(gram[q]);
    if (lane == 0) s_red[warp][q] = v;
  }
  reg = warp_sum(reg);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) { s_red[warp][ADMM_NPAIR] = reg; s_red[warp][ADMM_NPAIR + 1] = mx; }
  __syncthreads();
  const int nw = int(blockDim.x >> 5);
  double* out = a.gram_partial + ADMM_GSTRIDE * int64_t(blockIdx.x);
  if (tid < ADMM_NPAIR) {
    // fixed index (pair of an 8 x 8 triangle) -> (c, d) -> index in the l x l triangle the mid kernel reads
    int c = 0, rem = tid;
    while (rem >= ADMM_MAX_L - c) { rem -= ADMM_MAX_L - c; ++c; }
    const int d = c + rem;
    if (d < l) {
      double t = 0.0;
      for (int w = 0; w < nw; ++w) t += s_red[w][tid];
      out[admm_pair(c, d, l)] = t;
    }
  } else if (tid == ADMM_NPAIR) {
    double t = 0.0;
    for (int w = 0; w < nw; ++w) t += s_red[w][tid];
    out[tid] = t;
  } else if (tid == ADMM_NPAIR + 1) {
    double t = 0.0;
    for (int w = 0; w < nw; ++w) t = fmax(t, s_red[w][tid]);
    out[tid] = t;
    a.scal_partial[SCAL_STRIDE * int64_t(blockIdx.x) + 3] = t;
  }
}

// ---- mid: evaluate the iteration whose products just arrived, advance the control, z = L^-T L^-1 (P1 - b) ----
struct AdmmMidArgs {
  const Ctrl* cin;
  Ctrl* cout;
  AdmmConst c;
  const double* P;            // [2l][ldm] = A [x - rho u | u]
  const double* b;            // [l][ldm]
  double* Au_prev;            // [l][ldm]
  const double* L;            // [m][m] lower Cholesky factor
  const double* gram_partial; int32_t n_scal;   // [n_scal][ADMM_GSTRIDE] from the update kernel
  const double* cnt_partial;  int32_t n_cnt;
  void* r;                    // [l][ldm] dtype T: z, the operand of the transposed stream
  double* f_hist;
  double* f_hist_best;
  double size_total;          // n_total * l
  double* w_out;              // blocked path: w = P1 - b goes to this planar fp64 buffer [l][w_ld] and the kernel ends
  int64_t w_ld;               // there (the solves are two streams of Y = L^-T); nullptr: solve here with L
};

// largest singular value of a matrix with Gram matrix [[a, b], [b, c]] (l = 2) or [[a]] (l = 1)
__device__ __forceinline__ double admm_norm2(double a, double b, double c, int l) {
  if (l == 1) return sqrt(a);
  const double h = 0.5 * (a - c);
  return sqrt(0.5 * (a + c) + sqrt(h * h + b * b));
}
// l > 2: matrix 2-norm (numpy LA.norm(., 2), gl_ADMM_dual.py:87-88) = sqrt of the largest eigenvalue of the l x l Gram
// matrix, by cyclic Jacobi rotations (converges quadratically; 12 sweeps are far more than an 8 x 8 matrix needs).
// gu: upper triangle in admm_pair order.
__device__ inline double admm_norm2_jacobi(const double* gu, int l) {
  double M[ADMM_MAX_L][ADMM_MAX_L];
  for (int a = 0; a < l; ++a)
    for (int b = a; b < l; ++b) M[a][b] = M[b][a] = gu[admm_pair(a, b, l)];
  for (int sweep = 0; sweep < 12; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < l; ++p)
      for (int q = p + 1; q < l; ++q) off += M[p][q] * M[p][q];
    if (off == 0.0) break;
    for (int p = 0; p < l; ++p)
      for (int q = p + 1; q < l; ++q) {
        if (M[p][q] == 0.0) continue;
        const double theta = (M[q][q] - M[p][p]) / (2.0 * M[p][q]);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), sn = t * c;
        for (int k = 0; k < l; ++k) {   // columns p, q
          const double kp = M[k][p], kq = M[k][q];
          M[k][p] = c * kp - sn * kq; M[k][q] = sn * kp + c * kq;
        }
        for (int k = 0; k < l; ++k) {   // rows p, q
          const double pk = M[p][k], qk = M[q][k];
          M[p][k] = c * pk - sn * qk; M[q][k] = sn * pk + c * qk;
        }
      }
  }
  double mx = 0.0;
  for (int a = 0; a < l; ++a) mx = fmax(mx, M[a][a]);
  return sqrt(mx);
}

template <typename T>
__global__ void __launch_bounds__(ADMM_THREADS) glp_admm_mid_kernel(const AdmmMidArgs a, const Geom g) {
  extern __shared__ double s_w[];                 // [l][m] right-hand sides / solution
  constexpr int NV = 2 + 2 * ADMM_NPAIR;            // rr, reg, Gram of s (upper), Gram of r (upper)
  __shared__ double s_red[ADMM_THREADS / 32][NV];
  __shared__ double s_tot[NV];
  __shared__ Ctrl s_out;
  __shared__ double s_z[ADMM_MAX_L][32];
  __shared__ double s_blk[32][33];
  __shared__ double s_cnt[ADMM_THREADS / 32];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const Ctrl cin = *a.cin;
  if (cin.done) {
    if (tid == 0) *a.cout = cin;
    return;
  }
  const int l = g.l, m = g.m, npair = l * (l + 1) / 2;
  const int64_t ldm = g.ldm;
  const double* P1 = a.P;
  const double* P2 = a.P + int64_t(l) * ldm;
  // ---- reductions (thread-strided, warp tree, warps in order): ||A x - b||^2 with A x = P1 + rho P2, the Gram
  // matrix of s = A u_prev - A u, and the per-column-block records of the update kernel
  double rrv = 0.0, regv = 0.0, mx = 0.0, cnt = 0.0;
  double sg[ADMM_NPAIR], rg[ADMM_NPAIR];
#pragma unroll
  for (int q = 0; q < ADMM_NPAIR; ++q) { sg[q] = 0.0; rg[q] = 0.0; }
  for (int i = tid; i < m; i += ADMM_THREADS) {
    double sv[ADMM_MAX_L];
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c) {
      sv[c] = 0.0;
      if (c < l) {
        const double p2 = P2[c * ldm + i];
        const double d = xsub(xadd(P1[c * ldm + i], xmul(a.c.rho, p2)), a.b[c * ldm + i]);
        rrv += d * d;
        sv[c] = xsub(a.Au_prev[c * ldm + i], p2);
      }
    }
#pragma unroll
    for (int c = 0; c < ADMM_MAX_L; ++c)
#pragma unroll
      for (int d = c; d < ADMM_MAX_L; ++d) sg[admm_pair(c, d, ADMM_MAX_L)] += sv[c] * sv[d];
  }
  for (int i = tid; i < a.n_scal; i += ADMM_THREADS) {
    const double* p = a.gram_partial + ADMM_GSTRIDE * int64_t(i);
    for (int q = 0; q < npair; ++q) rg[q] += __ldcg(p + q);          // already in the l x l pair order
    regv += __ldcg(p + ADMM_NPAIR);
    mx = fmax(mx, __ldcg(p + ADMM_NPAIR + 1));
  }
  for (int i = tid; i < a.n_cnt; i += ADMM_THREADS) cnt += __ldcg(&a.cnt_partial[i]);
  rrv = warp_sum(rrv); regv = warp_sum(regv); cnt = warp_sum(cnt);
  if (lane == 0) { s_red[warp][0] = rrv; s_red[warp][1] = regv; s_cnt[warp] = cnt; }
#pragma unroll
  for (int c = 0; c < ADMM_MAX_L; ++c)
#pragma unroll
    for (int d = c; d < ADMM_MAX_L; ++d) {
      const double v = warp_sum(sg[admm_pair(c, d, ADMM_MAX_L)]);
      if (lane == 0 && d < l) s_red[warp][2 + admm_pair(c, d, l)] = v;
    }
#pragma unroll
  for (int q = 0; q < ADMM_NPAIR; ++q) {
    const double v = warp_sum(rg[q]);
    if (lane == 0 && q < npair) s_red[warp][2 + ADMM_NPAIR + q] = v;
  }
  __syncthreads();
  if (tid < NV) {
    const bool live = tid < 2 || (tid < 2 + ADMM_NPAIR ? tid - 2 < npair : tid - 2 - ADMM_NPAIR < npair);
    double t = 0.0;
    if (live) for (int w = 0; w < ADMM_THREADS / 32; ++w) t += s_red[w][tid];
    s_tot[tid] = t;
  }
  __syncthreads();
  if (tid == 0) {
    double ctot = 0.0;
    for (int w = 0; w < ADMM_THREADS / 32; ++w) ctot += s_cnt[w];
    Ctrl out = cin;
    const double f_now = xadd(xmul(0.5, s_tot[0]), xmul(a.c.mu, s_tot[1]));
    out.f_cur = f_now;
    out.s_cur = ctot / a.size_total;
    if (cin.k > 0) {   // iteration cin.k has just been completed by the forward stream: record and test it (:72-99)
      if (a.f_hist != nullptr && cin.k - 1 < a.c.hist_cap) a.f_hist[cin.k - 1] = f_now;
      out.f_best = f_now < cin.f_best ? f_now : cin.f_best;
      if (a.f_hist_best != nullptr && cin.k - 1 < a.c.hist_cap) a.f_hist_best[cin.k - 1] = out.f_best;
      const double* sgt = s_tot + 2;
      const double* rgt = s_tot + 2 + ADMM_NPAIR;
      const double rn = l <= 2 ? admm_norm2(rgt[0], l > 1 ? rgt[1] : 0.0, l > 1 ? rgt[2] : 0.0, l) : admm_norm2_jacobi(rgt, l);
      const double sn = l <= 2 ? admm_norm2(sgt[0], l > 1 ? sgt[1] : 0.0, l > 1 ? sgt[2] : 0.0, l) : admm_norm2_jacobi(sgt, l);
      out.stable = (rn < a.c.thres && sn < a.c.thres) ? cin.stable + 1 : 0;
      if (out.stable >= a.c.converge_len) out.done = 1;
    }
    if (!out.done) {
      if (cin.k < a.c.maxit) { out.k = cin.k + 1; out.n_fwd = cin.n_fwd + 1; out.n_tpass = cin.n_tpass + 1; }
      else out.done = 1;
    }
    s_out = out;
    *a.cout = out;
  }
  __syncthreads();
  if (s_out.done) return;

  if (a.w_out != nullptr) {   // blocked path: right-hand side to global memory, the streams of Y do the rest
    for (int i = tid; i < m; i += ADMM_THREADS)
      for (int c = 0; c < l; ++c) {
        a.w_out[c * a.w_ld + i] = xsub(P1[c * ldm + i], a.b[c * ldm + i]);
        a.Au_prev[c * ldm + i] = P2[c * ldm + i];
      }
    return;
  }
  // ---- w = P1 - b ; A u_prev <- A u ; then the two triangular solves in shared memory
  for (int i = tid; i < m; i += ADMM_THREADS)
    for (int c = 0; c < l; ++c) {
      s_w[c * m + i] = xsub(P1[c * ldm + i], a.b[c * ldm + i]);
      a.Au_prev[c * ldm + i] = P2[c * ldm + i];
    }
  __syncthreads();
  const double* L = a.L;
  // forward substitution  L y = w, 32 columns at a time: warp c solves the diagonal block of RHS c with
  // shuffles, then every thread subtracts the panel's contribution from the rows below
  for (int j0 = 0; j0 < m; j0 += 32) {
    const int pw = m - j0 < 32 ? m - j0 : 32;
    // the 32 x 32 diagonal block goes to shared memory first (one element per thread), so that the 32 dependent
    // steps below are shuffles, shared loads and FMAs only; a global load inside the chain costs ~0.7 us a step
    {
      const int bi = tid >> 5, bj = tid & 31;
      s_blk[bi][bj] = (bi < pw && bj <= bi) ? L[int64_t(j0 + bi) * m + j0 + bj] : 0.0;
    }
    __syncthreads();
    if (warp < l) {
      const int c = warp;
      const double dg = lane < pw ? s_blk[lane][lane] : 1.0;
      double wi = lane < pw ? s_w[c * m + j0 + lane] : 0.0;
      for (int jj = 0; jj < pw; ++jj) {
        const double zj = __shfl_sync(0xffffffffu, wi, jj) / __shfl_sync(0xffffffffu, dg, jj);
        if (lane == jj) wi = zj;
        else if (lane > jj && lane < pw) wi -= s_blk[lane][jj] * zj;
      }
      if (lane < pw) { s_w[c * m + j0 + lane] = wi; s_z[c][lane] = wi; }
    }
    __syncthreads();
    for (int i = j0 + pw + tid; i < m; i += ADMM_THREADS) {
      double acc[ADMM_MAX_L];
#pragma unroll
      for (int c = 0; c < ADMM_MAX_L; ++c) acc[c] = 0.0;
      const double* Li = L + int64_t(i) * m + j0;
      for (int jj = 0; jj < pw; ++jj) {
        const double lij = Li[jj];
#pragma unroll
        for (int c = 0; c < ADMM_MAX_L; ++c)
          if (c < l) acc[c] += lij * s_z[c][jj];
      }
#pragma unroll
      for (int c = 0; c < ADMM_MAX_L; ++c)
        if (c < l) s_w[c * m + i] -= acc[c];
    }
    __syncthreads();
  }
  // back substitution  L^T z = y, panels from the bottom
  for (int j1 = m; j1 > 0; j1 -= 32) {
    const int j0 = j1 - 32 > 0 ? j1 - 32 : 0, pw = j1 - j0;
    {
      const int bi = tid >> 5, bj = tid & 31;   // s_blk[bi][bj] = L[j0+bi][j0+bj], lower triangle
      s_blk[bi][bj] = (bi < pw && bj <= bi) ? L[int64_t(j0 + bi) * m + j0 + bj] : 0.0;
    }
    __syncthreads();
    if (warp < l) {
      const int c = warp;
      const double dg = lane < pw ? s_blk[lane][lane] : 1.0;
      double wi = lane < pw ? s_w[c * m + j0 + lane] : 0.0;
      for (int jj = pw - 1; jj >= 0; --jj) {
        const double zj = __shfl_sync(0xffffffffu, wi, jj) / __shfl_sync(0xffffffffu, dg, jj);
        if (lane == jj) wi = zj;
        else if (lane < jj) wi -= s_blk[jj][lane] * zj;   // (L^T)[lane][jj]
      }
      if (lane < pw) { s_w[c * m + j0 + lane] = wi; s_z[c][lane] = wi; }
    }
    __syncthreads();
    for (int i = tid; i < j0; i += ADMM_THREADS) {
      double acc[ADMM_MAX_L];
#pragma unroll
      for (int c = 0; c < ADMM_MAX_L; ++c) acc[c] = 0.0;
      for (int jj = 0; jj < pw; ++jj) {
        const double lji = L[int64_t(j0 + jj) * m + i];   // coalesced over i
#pragma unroll
        for (int c = 0; c < ADMM_MAX_L; ++c)
          if (c < l) acc[c] += lji * s_z[c][jj];
      }
#pragma unroll
      for (int c = 0; c < ADMM_MAX_L; ++c)
        if (c < l) s_w[c * m + i] -= acc[c];
    }
    __syncthreads();
  }
  T* r = static_cast<T*>(a.r);
  for (int i = tid; i < m; i += ADMM_THREADS)
    for (int c = 0; c < l; ++c) r[c * ldm + i] = T(s_w[c * m + i]);
}

}  // namespace glp
