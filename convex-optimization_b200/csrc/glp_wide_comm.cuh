// glp_wide_comm.cuh -- multi-GPU exchange of the wide multi-RHS mode (K5 products, l >= 16).
//
// With 2l = 256 operand columns the partial product of a rank is m x 2l fp64 = 33.5 MB (C5), too much to
// broadcast to every peer the way the l = 2 path does (7 x 33.5 MB per rank and iteration).  The sum
//   A X = sum_g A[:, J_g] X[J_g]
// is therefore taken as a reduce-scatter + all-gather over row slices, still push-model over the peer
// mailboxes (NVLink stores + 64-bit release flags, no collective library on the data path):
//   push    rank r stores rows I_j of its partial product into slab[r] of rank j, for every j != r, and
//           raises flag AX on all peers (with its share of the sparsity count);
//   reduce  rank j waits for the W flags, sums the W partial slices of ITS rows I_j in rank order, keeps
//           the result and stores it into rows I_j of slab[j] on every peer; raises flag SUM;
//   finish  every rank waits for the W-1 SUM flags, assembles the full product from the owners'
//           slices, and reduces ||A X - b||^2 and the exchanged scalars exactly like the l = 2 gather.
// Every row slice is summed once, by its owner, so all ranks hold bit-identical products and take
// identical line-search / stop decisions.  Traffic per rank and forward product: 2 (W-1)/W of one slab.
#pragma once
#include "glp_comm.cuh"

namespace glp {

constexpr int MAIL_FLAG_SUM = 24;   // between MAIL_FLAG_EP (16..23) and MAIL_SC (32)
constexpr int WIDEX_THREADS = 256;

struct WideXArgs {
  Comm comm;
  const Ctrl* ctrl;               // nullable
  int32_t sel, par;
  int32_t ncols;                  // 2l
  int32_t slice_rows;             // rows per owner slice (multiple of 32); slice j = [j*slice_rows, ...) of ldm
  unsigned long long seq_f, seq_t;
  double* AX[2];                  // [2l][ldm]: the local partial on entry of push, the full sum after finish
  const double* b;                // [l][ldm]
  const double* xcand[2];         // candidate x [l][ldn] (sparsity count)
  double* cnt_partial;            // [gridDim]
  double* rr_partial;             // [gridDim]
  double* scal_red;               // [SCAL_STRIDE]
  double* cnt_red;                // [1]
  unsigned* fin;
  int32_t* fault;
};

// last-CTA election shared by push and reduce: returns true on every lane of warp 0 of the last CTA
__device__ __forceinline__ bool widex_last_cta(unsigned* fin, int lane) {
  unsigned prev = 0;
  if (lane == 0) {
    __threadfence_system();
    prev = atomicAdd(fin, 1u);
  }
  prev = __shfl_sync(0xffffffffu, prev, 0);
  if (prev + 1 != gridDim.x) return false;
  __threadfence();
  if (lane == 0) *fin = 0u;
  return true;
}

__global__ void __launch_bounds__(WIDEX_THREADS)
glp_wide_push_kernel(const WideXArgs a, const Geom g) {
  __shared__ double s_red[WIDEX_THREADS / 32];
  if (a.ctrl != nullptr && a.ctrl->done) return;
  const int sel = a.ctrl != nullptr ? 1 - a.ctrl->cur : a.sel;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int W = a.comm.world, me = a.comm.rank;
  // ---- sparsity count of the local rows of the candidate against the global max|x| (posted by the prox tail)
  if (tid < W) wait_flag(mail_flag(a.comm, me, MAIL_FLAG_SC, tid), a.seq_t, a.fault);
  __syncthreads();
  double mx = 0.0;
  for (int r = 0; r < W; ++r) mx = comm_nmax(mx, ld_mail(mail_sc(a.comm, me, a.par, r)));
  const double cut = xmul(1e-6, mx);
  const int64_t stride = int64_t(gridDim.x) * WIDEX_THREADS, first = int64_t(blockIdx.x) * WIDEX_THREADS + tid;
  double cnt = 0.0;
  const double* xc = a.xcand[sel];
  const int64_t lenn = int64_t(g.l) * g.ldn;   // flat: padding row groups are zero and never counted
  for (int64_t j = first; j < lenn; j += stride) cnt += fabs(xc[j]) > cut ? 1.0 : 0.0;
  cnt = warp_sum(cnt);
  if (lane == 0) s_red[warp] = cnt;
  __syncthreads();
  if (tid == 0) {
    double t = 0.0;
    for (int w = 0; w < WIDEX_THREADS / 32; ++w) t += s_red[w];
    a.cnt_partial[blockIdx.x] = t;
  }
  // ---- push the row slices the peers own
  const double* src = a.AX[sel];
  const int64_t rs = a.slice_rows;
  for (int d = 1; d < W; ++d) {
    const int j = (me + d) % W;   // staggered start: every rank targets a different peer first
    const int64_t r0 = int64_t(j) * rs;
    const int64_t rows = (r0 + rs <= g.ldm ? rs : (g.ldm > r0 ? g.ldm - r0 : 0));
    if (rows <= 0) continue;
    double* dst = mail_slab(a.comm, j, a.par, me);
    const int64_t total = rows * a.ncols;
    for (int64_t e = first; e < total; e += stride) {
      const int64_t c = e / rows, idx = c * g.ldm + r0 + (e - c * rows);
      dst[idx] = src[idx];
    }
  }
  __syncthreads();
  if (warp == 0 && widex_last_cta(a.fin, lane)) {
    double tot = 0.0;
    for (int i = lane; i < int(gridDim.x); i += 32) tot += __ldcg(&a.cnt_partial[i]);
    tot = warp_sum(tot);
    if (lane < W) {
      mail_sc(a.comm, lane, a.par, me)[4] = tot;
      __threadfence_system();
      st_release_sys(mail_flag(a.comm, lane, MAIL_FLAG_AX, me), a.seq_f);
    }
  }
}

__global__ void __launch_bounds__(WIDEX_THREADS)
glp_wide_reduce_kernel(const WideXArgs a, const Geom g) {
  if (a.ctrl != nullptr && a.ctrl->done) return;
  const int sel = a.ctrl != nullptr ? 1 - a.ctrl->cur : a.sel;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int W = a.comm.world, me = a.comm.rank;
  if (tid < W) wait_flag(mail_flag(a.comm, me, MAIL_FLAG_AX, tid), a.seq_f, a.fault);
  __syncthreads();
  const int64_t rs = a.slice_rows, r0 = int64_t(me) * rs;
  const int64_t rows = (r0 + rs <= g.ldm ? rs : (g.ldm > r0 ? g.ldm - r0 : 0));
  double* mine = a.AX[sel];
  const int64_t total = rows * a.ncols;
  const int64_t stride = int64_t(gridDim.x) * WIDEX_THREADS, first = int64_t(blockIdx.x) * WIDEX_THREADS + tid;
  for (int64_t e = first; e < total; e += stride) {
    const int64_t c = e / rows, idx = c * g.ldm + r0 + (e - c * rows);
    double s = 0.0;
    for (int r = 0; r < W; ++r) s += (r == me) ? mine[idx] : ld_mail(mail_slab(a.comm, me, a.par, r) + idx);
    mine[idx] = s;
    for (int d = 1; d < W; ++d) mail_slab(a.comm, (me + d) % W, a.par, me)[idx] = s;
  }
  __syncthreads();
  if (warp == 0 && widex_last_cta(a.fin, lane) && lane < W) {
    __threadfence_system();
    st_release_sys(mail_flag(a.comm, lane, MAIL_FLAG_SUM, me), a.seq_f);
  }
}

__global__ void __launch_bounds__(WIDEX_THREADS)
glp_wide_finish_kernel(const WideXArgs a, const Geom g) {
  __shared__ double s_red[WIDEX_THREADS / 32];
  if (a.ctrl != nullptr && a.ctrl->done) return;
  const int sel = a.ctrl != nullptr ? 1 - a.ctrl->cur : a.sel;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int W = a.comm.world, me = a.comm.rank;
  if (tid < W && tid != me) wait_flag(mail_flag(a.comm, me, MAIL_FLAG_SUM, tid), a.seq_f, a.fault);
  __syncthreads();
  double* out = a.AX[sel];
  const int64_t rs = a.slice_rows;
  const int64_t total = int64_t(a.ncols) * g.ldm, lenb = int64_t(g.l) * g.ldm;
  double rr = 0.0;
  for (int64_t idx = int64_t(blockIdx.x) * WIDEX_THREADS + tid; idx < total; idx += int64_t(gridDim.x) * WIDEX_THREADS) {
    const int64_t i = idx % g.ldm;
    int owner = int(i / rs);
    owner = owner < W ? owner : W - 1;
    double v;
    if (owner == me) {
      v = out[idx];
    } else {
      v = ld_mail(mail_slab(a.comm, me, a.par, owner) + idx);
      out[idx] = v;
    }
    if (idx < lenb) {
      const double d = v - a.b[idx];
      rr += d * d;
    }
  }
  rr = warp_sum(rr);
  if (lane == 0) s_red[warp] = rr;
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < WIDEX_THREADS / 32; ++w) tot += s_red[w];
    a.rr_partial[blockIdx.x] = tot;
    if (blockIdx.x == 0) {
      double mx = 0.0, gd = 0.0, dd = 0.0, reg = 0.0, cnt = 0.0, sx = 0.0, sy = 0.0;
      for (int r = 0; r < W; ++r) {
        const double* sc = mail_sc(a.comm, me, a.par, r);
        mx = comm_nmax(mx, ld_mail(sc + 0));
        gd += ld_mail(sc + 1); dd += ld_mail(sc + 2); reg += ld_mail(sc + 3); cnt += ld_mail(sc + 4);
        sx += ld_mail(sc + 5); sy += ld_mail(sc + 6);
      }
      a.scal_red[0] = gd; a.scal_red[1] = dd; a.scal_red[2] = reg; a.scal_red[3] = mx;
      a.scal_red[4] = sx; a.scal_red[5] = sy;
      a.cnt_red[0] = cnt;
    }
  }
}

}  // namespace glp
