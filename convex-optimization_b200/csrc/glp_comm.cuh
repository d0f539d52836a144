// glp_comm.cuh -- multi-GPU exchange over NVLink 5 / NVSwitch peer memory, written as part of the
// streaming kernels themselves (no collective launch on the data path).
//
// The reference has no distributed code; the exchange below is what column-block sharding of A
//  needs per iteration:  A*X = sum_g A[:, J_g] * X[J_g]  (one m x 2l SUM) plus the
// line-search / objective scalars (SUM) and max|x| (MAX).
//
// Push model, one-shot: the CTA that completes a 64-row block of its rank's partial product stores
// it straight into the `slab[src]` of EVERY rank's mailbox while the rest of A is still streaming,
// so the transfer overlaps the stream block by block.  The last CTA of the launch raises a 64-bit
// flag on every peer (st.release.sys after a system fence).  The gather kernel on each rank waits
// for the W flags in its own HBM and sums the W slabs in rank order: every rank performs the same
// additions in the same order, so all ranks hold bit-identical products / scalars and therefore
// take identical line-search, stop-rule and stage decisions without any broadcast.
#pragma once
#include "glp_device.cuh"
#include "glp_types.cuh"

namespace glp {

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// Data that a peer GPU stored into this GPU's HBM: read at L2 (the coherence point), never from L1.
__device__ __forceinline__ double ld_mail(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
// Bounded spin (20 s): a lost peer must not hang the GPU; the fault word makes the host report it.
__device__ __forceinline__ void wait_flag(const unsigned long long* p, unsigned long long target, int* fault) {
  if (ld_acquire_sys(p) >= target) return;
  const uint64_t t0 = global_timer_ns();
  uint32_t spins = 0;
  while (ld_acquire_sys(p) < target) {
    __nanosleep(64);
    if ((++spins & 0xffu) == 0 && global_timer_ns() - t0 > 20000000000ull) {
      if (fault) atomicExch(fault, 2);
      __threadfence_system();
      __trap();
    }
  }
}

__device__ __forceinline__ unsigned long long* mail_flag(const Comm& c, int owner, int base, int src) {
  return c.mail[owner] + base + src;
}
__device__ __forceinline__ double* mail_sc(const Comm& c, int owner, int par, int src) {
  return reinterpret_cast<double*>(c.mail[owner] + MAIL_SC) + (par * MAX_WORLD + src) * 8;
}
__device__ __forceinline__ double* mail_slab(const Comm& c, int owner, int par, int src) {
  return reinterpret_cast<double*>(c.mail[owner] + MAIL_SLAB) + (int64_t(par) * MAX_WORLD + src) * c.slab_words;
}

// ---- solve-epoch barrier (two launches so that G logical ranks can share one stream in tests) ----
__global__ void glp_epoch_post_kernel(const Comm c, const unsigned long long epoch) {
  const int r = threadIdx.x;
  if (r < c.world) {
    __threadfence_system();
    st_release_sys(mail_flag(c, r, MAIL_FLAG_EP, c.rank), epoch);
  }
}
__global__ void glp_epoch_wait_kernel(const Comm c, const unsigned long long epoch, int* fault) {
  const int r = threadIdx.x;
  if (r < c.world) wait_flag(mail_flag(c, c.rank, MAIL_FLAG_EP, r), epoch, fault);
}

// ---- scalar post: called by one full warp of the last CTA of a prox-producing launch ------------
// scal_partial[ncb][SCAL_STRIDE] = {gd, dd, reg, max, sx, sy}: lane-strided fixed-order sums, then one record
// per peer and the flag.
__device__ __forceinline__ double comm_nmax(double a, double b) {
  return (a != a) ? a : ((b != b) ? b : fmax(a, b));
}
__device__ __forceinline__ void post_prox_scalars(const ProxPost& P, const double* scal_partial, int ncb, int lane) {
  double gd = 0.0, dd = 0.0, reg = 0.0, mx = 0.0, sx = 0.0, sy = 0.0;
  for (int i = lane; i < ncb; i += 32) {
    gd += __ldcg(&scal_partial[SCAL_STRIDE * i + 0]);
    dd += __ldcg(&scal_partial[SCAL_STRIDE * i + 1]);
    reg += __ldcg(&scal_partial[SCAL_STRIDE * i + 2]);
    mx = comm_nmax(mx, __ldcg(&scal_partial[SCAL_STRIDE * i + 3]));
    sx += __ldcg(&scal_partial[SCAL_STRIDE * i + 4]);
    sy += __ldcg(&scal_partial[SCAL_STRIDE * i + 5]);
  }
  gd = warp_sum(gd); dd = warp_sum(dd); reg = warp_sum(reg); sx = warp_sum(sx); sy = warp_sum(sy);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = comm_nmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane < P.comm.world) {
    double* dst = mail_sc(P.comm, lane, P.par, P.comm.rank);
    dst[0] = mx; dst[1] = gd; dst[2] = dd; dst[3] = reg; dst[5] = sx; dst[6] = sy;
    __threadfence_system();
    st_release_sys(mail_flag(P.comm, lane, MAIL_FLAG_SC, P.comm.rank), P.seq_t);
  }
}

// Every CTA of a prox-producing launch calls this once, with all threads of warp 0, after its last
// prox_block: the last CTA to arrive publishes.
__device__ __forceinline__ void prox_post_arrive(const ProxPost& P, const double* scal_partial, int ncb, int lane) {
  if (P.comm.world <= 1) return;
  unsigned prev = 0;
  if (lane == 0) {
    __threadfence();
    prev = atomicAdd(P.fin, 1u);
  }
  prev = __shfl_sync(0xffffffffu, prev, 0);
  if (prev + 1 == gridDim.x) {
    __threadfence();
    if (lane == 0) *P.fin = 0u;
    post_prox_scalars(P, scal_partial, ncb, lane);
  }
}

// ---- gather: sum the W slabs (rank order), residual norm, scalars -------------------------------
struct GatherArgs {
  Comm comm;
  const Ctrl* ctrl;            // nullable
  int32_t sel;
  int32_t par;
  unsigned long long seq_f, seq_t;
  double* AX[2];               // [2l][ldm]
  const double* b;             // [l][ldm]
  double* rr_partial;          // [gridDim]
  double* scal_red;            // [SCAL_STRIDE] = {gd, dd, reg, max, sx, sy}
  double* cnt_red;             // [1]
  int32_t* fault;
};
constexpr int GATHER_THREADS = 256;

// The gather proper: executed by every CTA of a grid of GATHER_THREADS-wide CTAs (the stand-alone gather kernel, or the
// first half of the fused gather + decide kernel).
__device__ __forceinline__ void gather_phase(const GatherArgs& a, const Geom& g, int sel, double* s_red) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int W = a.comm.world, me = a.comm.rank;
  if (tid < W) wait_flag(mail_flag(a.comm, me, MAIL_FLAG_AX, tid), a.seq_f, a.fault);
  else if (tid >= 32 && tid < 32 + W) wait_flag(mail_flag(a.comm, me, MAIL_FLAG_SC, tid - 32), a.seq_t, a.fault);
  __syncthreads();
  const int64_t len = a.comm.slab_words, lenb = int64_t(g.l) * g.ldm;
  const double* slab0 = mail_slab(a.comm, me, a.par, 0);
  double* out = a.AX[sel];
  double rr = 0.0;
  for (int64_t i = int64_t(blockIdx.x) * GATHER_THREADS + tid; i < len; i += int64_t(gridDim.x) * GATHER_THREADS) {
    double s = 0.0;
    for (int r = 0; r < W; ++r) s += ld_mail(slab0 + r * len + i);
    out[i] = s;
    if (i < lenb) {
      const double d = s - a.b[i];
      rr += d * d;
    }
  }
  rr = warp_sum(rr);
  if (lane == 0) s_red[warp] = rr;
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < GATHER_THREADS / 32; ++w) tot += s_red[w];
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

__global__ void __launch_bounds__(GATHER_THREADS)
glp_gather_kernel(const GatherArgs a, const Geom g) {
  __shared__ double s_red[GATHER_THREADS / 32];
  pdl_wait();
  pdl_launch();
  if (a.ctrl != nullptr && a.ctrl->done) return;
  const int sel = a.ctrl != nullptr ? 1 - a.ctrl->cur : a.sel;
  gather_phase(a, g, sel, s_red);
}

}  // namespace glp

// ---- NCCL baseline exchange (A/B measurement only; the product path is the peer-mailbox push above) ----------
// The same sums taken with two ncclAllReduce calls per forward stream on the solver's stream: MAX of max|xn|
// (one double, after the prox) and SUM of this rank's partial product with the scalar partials appended
// (2 l ldm + 8 doubles, after the forward stream).  The kernels below only pack / unpack the staging buffer;
// the destination buffer of the product is chosen on the device (1 - ctrl->cur), which a host-side
// collective cannot see, hence the staging copy.
namespace glp {

struct NcclXArgs {
  const Ctrl* ctrl;
  double* stage;               // [2l*ldm] partial product, then tail[8] = {gd, dd, reg, sx, sy, count, -, -}
  double* stage_m;             // [1] max|xn| (MAX-reduced)
  int64_t slab_words;          // 2l*ldm
  const double* scal_partial;  // [ncb][SCAL_STRIDE]
  int32_t ncb, pad;
  const double* xcand[2];
  double* AX[2];
  const double* b;
  double* rr_partial;          // [gridDim of the post kernel]
  double* scal_red;
  double* cnt_red;
};

__global__ void __launch_bounds__(256) glp_nccl_fold_kernel(const NcclXArgs a) {
  __shared__ double s[8][6];
  if (a.ctrl != nullptr && a.ctrl->done) return;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  double gd = 0.0, dd = 0.0, reg = 0.0, mx = 0.0, sx = 0.0, sy = 0.0;
  for (int i = tid; i < a.ncb; i += 256) {
    const double* p = a.scal_partial + SCAL_STRIDE * i;
    gd += __ldcg(p + 0); dd += __ldcg(p + 1); reg += __ldcg(p + 2); mx = comm_nmax(mx, __ldcg(p + 3));
    sx += __ldcg(p + 4); sy += __ldcg(p + 5);
  }
  gd = warp_sum(gd); dd = warp_sum(dd); reg = warp_sum(reg); sx = warp_sum(sx); sy = warp_sum(sy);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = comm_nmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) { s[warp][0] = gd; s[warp][1] = dd; s[warp][2] = reg; s[warp][3] = mx; s[warp][4] = sx; s[warp][5] = sy; }
  __syncthreads();
  if (tid == 0) {
    double t[6] = {0, 0, 0, 0, 0, 0};
    for (int w = 0; w < 8; ++w) {
      t[0] += s[w][0]; t[1] += s[w][1]; t[2] += s[w][2]; t[3] = comm_nmax(t[3], s[w][3]); t[4] += s[w][4]; t[5] += s[w][5];
    }
    double* tail = a.stage + a.slab_words;
    tail[0] = t[0]; tail[1] = t[1]; tail[2] = t[2]; tail[3] = t[4]; tail[4] = t[5]; tail[5] = 0.0; tail[6] = 0.0; tail[7] = 0.0;
    a.stage_m[0] = t[3];
  }
}

// local sparsity count against the MAX-reduced max|xn| (counts are integers: the atomic sum is exact)
__global__ void __launch_bounds__(256) glp_nccl_count_kernel(const NcclXArgs a, const Geom g) {
  if (a.ctrl != nullptr && a.ctrl->done) return;
  const int sel = a.ctrl != nullptr ? 1 - a.ctrl->cur : 0;
  const double cut = xmul(1e-6, a.stage_m[0]);
  const double* xc = a.xcand[sel];
  const int64_t lenn = int64_t(g.l) * g.ldn;
  double cnt = 0.0;
  for (int64_t j = int64_t(blockIdx.x) * 256 + threadIdx.x; j < lenn; j += int64_t(gridDim.x) * 256)
    cnt += fabs(xc[j]) > cut ? 1.0 : 0.0;
  cnt = warp_sum(cnt);
  if ((threadIdx.x & 31) == 0 && cnt != 0.0) atomicAdd(a.stage + a.slab_words + 5, cnt);
}

__global__ void __launch_bounds__(GATHER_THREADS) glp_nccl_post_kernel(const NcclXArgs a, const Geom g) {
  __shared__ double s_red[GATHER_THREADS / 32];
  if (a.ctrl != nullptr && a.ctrl->done) return;
  const int sel = a.ctrl != nullptr ? 1 - a.ctrl->cur : 0;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  double* out = a.AX[sel];
  const int64_t lenb = int64_t(g.l) * g.ldm;
  double rr = 0.0;
  for (int64_t i = int64_t(blockIdx.x) * GATHER_THREADS + tid; i < a.slab_words; i += int64_t(gridDim.x) * GATHER_THREADS) {
    const double v = a.stage[i];
    out[i] = v;
    if (i < lenb) {
      const double d = v - a.b[i];
      rr += d * d;
    }
  }
  rr = warp_sum(rr);
  if (lane == 0) s_red[warp] = rr;
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
    for (int w = 0; w < GATHER_THREADS / 32; ++w) tot += s_red[w];
    a.rr_partial[blockIdx.x] = tot;
    if (blockIdx.x == 0) {
      const double* tail = a.stage + a.slab_words;
      a.scal_red[0] = tail[0]; a.scal_red[1] = tail[1]; a.scal_red[2] = tail[2]; a.scal_red[3] = a.stage_m[0];
      a.scal_red[4] = tail[3]; a.scal_red[5] = tail[4];
      a.cnt_red[0] = tail[5];
    }
  }
}

}  // namespace glp
