// glp_gemm_pair.cuh -- K5, fp16-pair product on CTA PAIRS: tcgen05.mma.cta_group::2 (M = 256 across two SMs of one TPC)
// with a stream-K split of the K chunks over all SM pairs.
//
// Same product as glp_gemm_tf32_kernel<NT, 4> (glp_gemm.cuh):  D[M x N] = P[M x K] . Q[N x K]^T  with
// P s_P = P0 + P1, Q s_Q = Q0 + Q1 in fp16 and  P0.Q0 + (P0.Q1 + P1.Q0)  on kind::f16, the main term flushed to fp64
// every k_chunk_blocks * 64 elements of K, batched forms of `A @ x` / `A.T @ r` (gl_*_primal.py:25,61,129 / :63,66).
// What changes, and why (round-1 ncu of the one-CTA kernel at C5, profiles/r01_ncu_full_gemm_f16x2_fwd_tpass_C5.csv:
// tensor pipe 60 % active, smem->tensor datapath 51 %, 128 / 132 of 148 CTAs launched):
//   * M = 128, N = 128 instructions read (128 + 128) x 32 B of shared memory each, more than an SM's shared-memory
//     port delivers per MMA slot: the tensor pipe cannot be fed.  A CTA pair runs ONE M = 256 instruction over both
//     SMs; each CTA stages its own 128 rows of P and only HALF of the Q tile (N/2 rows), so the shared-memory
//     traffic per SM and MMA drops from 8 KB to 6 KB and the Q bytes through L2 halve.
//   * Work unit = (256-row tile pair, column group, K chunk).  The unit sequence (tile-major) is cut into equal
//     contiguous ranges, one per CTA pair (stream-K): all 148 SMs move the same bytes whatever the tile count.
//     A tile whose chunks are shared by several pairs is finished by the last arriver, which sums the partial
//     tiles in pair order = ascending K order (fixed order: run-to-run identical bits).
// Roles per CTA (19 warps): warp 0 TMA producer (both CTAs; every load signals the LEADER's full barrier),
// warp 1 allocates TMEM (both CTAs) and, in the leader CTA only, one lane issues the MMAs for the pair;
// warps 2-17 epilogue (both CTAs, each on its own 128 rows: 4 TMEM lane quarters x 4 column groups, tcgen05.ld into
// fp64 register accumulators; three chunk accumulators in flight so that the drain never stalls the MMA issuer);
// warp 18: producer of the separate Q ring when PAIR_RINGS is set.
#pragma once
#include "glp_gemm.cuh"
#include "glp_partition.h"

namespace glp {

// 16 epilogue warps: 4 TMEM lane quarters x up to 4 column groups of PAIR_HC columns each (the drain of a chunk
// accumulator -- tcgen05.ld, 64-bit conversion, fp64 add -- sits between two MMA batches: the shorter, the better)
constexpr int PAIR_EPI_WARPS = 16;
constexpr int PAIR_THREADS = (3 + PAIR_EPI_WARPS) * 32;   // + P producer, MMA issuer, Q producer
constexpr int PAIR_NACC = 3;                // chunk accumulators in flight (TMEM: (3 + 1) * NT <= 512 columns)
template <int NT> constexpr int pair_hc() { return NT >= 64 ? NT / 4 : 16; }
template <int NT> constexpr int pair_ncg() { return NT / pair_hc<NT>(); }
// Two operand rings with their own depths: the P tiles stream from HBM (what has to be in flight to cover its loaded
// latency: ~3 us x 44 GB/s per SM = 130 KB), the Q tiles are re-read from L2 by every tile pair.  One ring of {P | Q}
// stages (first version: 4 x 48 KB) kept 96 KB of P in flight and measured 4.9 TB/s; 5 P stages keep 128 KB.
// Measured at C5 on one B200 (profiles/r02_ab_gemm_pair_c5.log): one ring 406 it/s (fwd window 1.087 ms / tpass
// 0.880 ms), two rings 398.5 it/s (1.114 / 0.910), one-CTA kernel 394 (1.116 / 0.960).  More P bytes in flight do
// not help: after the epilogue fix the MMA issuer never waits for data or for a free accumulator (ncu source view),
// and the tensor pipe sits at 59 % -- the ceiling of shared-memory bandwidth for this operand mix: per K = 16 step a
// CTA's three MMAs read 3 x (4 KB of P + 2 KB of Q) and TMA writes 4 KB, 22 KB per 192 tensor cycles against
// 128 B/clk (the one-CTA kernel reads 3 x 8 KB + 5.3 KB written: the 0.9 ms of round 1).  Kept switchable.
constexpr bool PAIR_RINGS = false;
template <int NT> constexpr int pair_p_stages() { return PAIR_RINGS ? (NT == 128 ? 5 : 6) : (NT == 128 ? 4 : 5); }
template <int NT> constexpr int pair_q_stages() { return PAIR_RINGS ? 4 : pair_p_stages<NT>(); }
constexpr int PAIR_P_STAGE = 2 * GEMM_BM * 128;                              // {P0 | P1}: 32 KB
template <int NT> constexpr int pair_q_stage() { return 2 * (NT / 2) * 128; }   // {Q0 half | Q1 half}
template <int NT> constexpr int pair_smem_bytes() {
  return pair_p_stages<NT>() * PAIR_P_STAGE + pair_q_stages<NT>() * pair_q_stage<NT>() + 1024 + 256;
}
constexpr int PAIR_COUNTER_WORDS = 65536;   // one arrival counter per (128-row tile, column group)
// workspace: [PAIR_COUNTER_WORDS] unsigned counters (zero between launches), then 2 partial-tile slots per CTA
template <int NT> constexpr size_t pair_slot_doubles() { return size_t(NT) * GEMM_BM; }
inline size_t pair_ws_bytes(int ctas, int NT) { return size_t(PAIR_COUNTER_WORDS) * 4 + size_t(2) * ctas * NT * GEMM_BM * 8; }

// ---- cluster / cta_group::2 PTX -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t mapa_shared(uint32_t local_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta_rank));
  return r;
}
// Arrive on an mbarrier of the leader CTA.  Default semantics (release at CTA scope): what is released are this
// thread's TMEM reads, already ordered by tcgen05.fence::before_thread_sync; a cluster-scope release costs a GPU-wide
// memory barrier per arrive (MEMBAR.ALL.GPU + ERRBAR: 20 % of the kernel's stall samples in the first version).
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load of a CTA pair: lands in THIS CTA's shared memory, completes bytes on the mbarrier at `bar_cluster_addr`,
// which may live in the peer (leader) CTA
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, uint32_t bar_cluster_addr,
                                                 int32_t c0, int32_t c1, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the mbarrier at this offset in both CTAs of the pair when the pair's MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(uint16_t(3)) : "memory");
}
// kind::f16 with fp16 operands, fp32 accumulate, both K-major, M = 256 over the pair
template <int NT> __device__ __forceinline__ constexpr uint32_t umma_idesc_f16_m256() {
  return (1u << 4) | (uint32_t(NT >> 3) << 17) | (uint32_t(256 >> 4) << 24);
}

struct PairArgs {
  int64_t M, K;
  int32_t N;
  int32_t k_chunk_blocks;        // K blocks (64 fp16 each) accumulated in fp32 before an fp64 flush (0 = all)
  double* D;                     // planar fp64 [N][ldd]
  int64_t ldd;
  int32_t* fault;
  const Ctrl* ctrl;              // nullable; role / D_alt / gate as in GemmArgs
  double* D_alt;
  int32_t role;
  const int32_t* gate;
  int32_t gate_cap;
  const double* scale;           // the epilogue multiplies by *scale (power of two: exact)
  int32_t p_tiled;               // 1: P0 / P1 are stored as 128 x 64 tiles (f16_tiled_off), maps over [tiles*nkb*128][64]
  unsigned* counters;            // [tiles128 * groups] zero between launches
  double* slots;                 // [2 * gridDim.x][NT][128] partial tiles
};

// one contiguous piece of a pair's unit range inside one (tile pair, group)
struct PairSeg { int64_t tp, group, ch0, ch1, first, last; bool whole; };
__device__ __forceinline__ PairSeg pair_seg(int64_t u, int64_t hi, int64_t nchunks, int64_t G, int64_t NP, int64_t U) {
  PairSeg s;
  const int64_t tg = u / nchunks;
  s.tp = tg / G;
  s.group = tg - s.tp * G;
  s.ch0 = u - tg * nchunks;
  s.ch1 = s.ch0 + (hi - u) < nchunks ? s.ch0 + (hi - u) : nchunks;
  s.whole = s.ch0 == 0 && s.ch1 == nchunks;
  s.first = tile_owner(tg * nchunks, NP, U);
  s.last = tile_owner((tg + 1) * nchunks - 1, NP, U);
  return s;
}

template <int NT>
__global__ void __launch_bounds__(PAIR_THREADS, 1)
glp_gemm_pair_kernel(const __grid_constant__ CUtensorMap tmP0, const __grid_constant__ CUtensorMap tmP1,
                     const __grid_constant__ CUtensorMap tmQ0, const __grid_constant__ CUtensorMap tmQ1,
                     const PairArgs a) {
  constexpr int BK = 2 * GEMM_BK;                       // 64 fp16 = one 128-byte swizzle row
  constexpr int QH = NT / 2;                            // rows of the Q tile staged by each CTA
  constexpr int P_BYTES = GEMM_BM * 128, Q_BYTES = QH * 128;
  constexpr int SP = pair_p_stages<NT>(), SQ = pair_q_stages<NT>();
  constexpr int PS_BYTES = PAIR_P_STAGE, QS_BYTES = pair_q_stage<NT>();
  constexpr int NACC = PAIR_NACC;
  constexpr uint32_t TMEM_NEED = (NACC + 1) * NT;       // chunk accumulators + the correction accumulator
  constexpr uint32_t TMEM_COLS = TMEM_NEED <= 128 ? 128 : TMEM_NEED <= 256 ? 256 : 512;
  constexpr int HC = pair_hc<NT>();                     // columns per epilogue thread
  constexpr int NCG = pair_ncg<NT>();                   // active column groups
  constexpr int EPI_ACTIVE = 4 * NCG;                   // epilogue warps with work, per CTA
  static_assert(TMEM_NEED <= 512, "TMEM");
  static_assert(NT == 32 || NT == 64 || NT == 128, "NT");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_q = smem + SP * PS_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_q + SQ * QS_BYTES);   // [SP] P ring, used in the leader CTA
  uint64_t* empty = full + SP;          // [SP] in each CTA: the pair's MMAs have read this CTA's P stage
  uint64_t* fullq = empty + SP;         // [SQ] Q ring, used in the leader CTA
  uint64_t* emptyq = fullq + SQ;        // [SQ] in each CTA
  uint64_t* tfull = emptyq + SQ;        // [NACC] in each CTA: chunk accumulator ready
  uint64_t* tempty = tfull + NACC;      // [NACC] leader: chunk accumulator drained by the epilogues of BOTH CTAs
  uint64_t* cempty = tempty + NACC;     // [1] leader: correction accumulator drained by both
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(cempty + 1);
  __shared__ int s_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();              // 0 = leader
  const bool leader = rank == 0;

  if (tid == 0) {
    for (int s = 0; s < SP; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int s = 0; s < SQ; ++s) { mbar_init(&fullq[s], 1); mbar_init(&emptyq[s], 1); }
    for (int s = 0; s < NACC; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 2 * EPI_ACTIVE); }
    mbar_init(cempty, 2 * EPI_ACTIVE);
    fence_mbar_init();
    fence_proxy_async();
    prefetch_tensormap(&tmP0); prefetch_tensormap(&tmP1); prefetch_tensormap(&tmQ0); prefetch_tensormap(&tmQ1);
  }
  if (warp == 1) tmem_alloc_pair(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  cluster_sync_all();   // both CTAs' barriers exist and both hold their TMEM before any cross-CTA signal
  pdl_wait();
  pdl_launch();

  // grid-uniform early exits (after the collective set-up so that the teardown below stays symmetric)
  bool skip = false;
  double* Dout = a.D;
  if (a.ctrl != nullptr) {
    if (a.ctrl->done || (a.role == 2 && a.ctrl->retry)) skip = true;
    if (a.role == 1 && a.ctrl->cur == 0) Dout = a.D_alt;
  }
  if (a.gate != nullptr && *a.gate <= a.gate_cap) skip = true;

  const int64_t tps = (a.M + 2 * GEMM_BM - 1) / (2 * GEMM_BM);
  const int64_t nkb = (a.K + BK - 1) / BK;
  const int64_t kc = a.k_chunk_blocks > 0 ? a.k_chunk_blocks : nkb;
  const int64_t nchunks = (nkb + kc - 1) / kc;
  const int64_t G = (a.N + NT - 1) / NT;
  const int64_t U = tps * G * nchunks;
  // active pairs: never more than units, so that every contributor range of a shared tile is non-empty (the last
  // arriver of a tile counts last - first + 1 arrivals)
  const int64_t NP = (gridDim.x / 2) < U ? (gridDim.x / 2) : U, pair = blockIdx.x / 2;
  if (pair >= NP) skip = true;
  const int64_t lo = skip ? 0 : tile_lo(pair, NP, U), hi = skip ? 0 : tile_lo(pair + 1, NP, U);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer of the P ring (both CTAs)
    if (lane == 0) {
      const uint64_t pol_stream = policy_evict_first();
      int stage = 0; uint32_t phase = 0;
      for (int64_t u = lo; u < hi;) {
        const PairSeg sg = pair_seg(u, hi, nchunks, G, NP, U);
        const int32_t r0 = int32_t((sg.tp * 2 + rank) * GEMM_BM);
        const int64_t kb0 = sg.ch0 * kc, kb1 = sg.ch1 * kc < nkb ? sg.ch1 * kc : nkb;
        for (int64_t kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1u, a.fault);
          // the leader's barrier counts the bytes of both CTAs (a peer load that lands first only drives the
          // transaction count negative until the leader's expect_tx: the phase cannot complete before that arrive)
          if (leader) mbar_arrive_expect_tx(&full[stage], PAIR_RINGS ? 2 * PS_BYTES : 2 * (PS_BYTES + QS_BYTES));
          const uint32_t bar = mapa_shared(smem_u32(&full[stage]), 0);
          uint8_t* s = smem + stage * PS_BYTES;
          const int32_t k0 = int32_t(kb * BK);
          // tiled operand: tile (row block, K block) is 128 consecutive rows of the [tiles * nkb * 128][64] view
          const int32_t pc0 = a.p_tiled ? 0 : k0;
          const int32_t pc1 = a.p_tiled ? int32_t(((sg.tp * 2 + rank) * nkb + kb) * GEMM_BM) : r0;
          tma_load_2d_pair(s, &tmP0, bar, pc0, pc1, pol_stream);
          tma_load_2d_pair(s + P_BYTES, &tmP1, bar, pc0, pc1, pol_stream);
          if (!PAIR_RINGS) {   // one ring: the Q stage of the same index rides on the same barrier
            const int32_t q0 = int32_t(sg.group * NT + rank * QH);
            uint8_t* sq = smem_q + stage * QS_BYTES;
            const uint64_t pol_keep = policy_evict_normal();
            tma_load_2d_pair(sq, &tmQ0, bar, k0, q0, pol_keep);
            tma_load_2d_pair(sq + Q_BYTES, &tmQ1, bar, k0, q0, pol_keep);
          }
          if (++stage == SP) { stage = 0; phase ^= 1u; }
        }
        u += sg.ch1 - sg.ch0;
      }
    }
  } else if (warp == 2 + PAIR_EPI_WARPS) {
    // ------------------------------------------------------------------ TMA producer of the Q ring (both CTAs)
    if (PAIR_RINGS && lane == 0) {
      const uint64_t pol_keep = policy_evict_normal();
      int stage = 0; uint32_t phase = 0;
      for (int64_t u = lo; u < hi;) {
        const PairSeg sg = pair_seg(u, hi, nchunks, G, NP, U);
        const int32_t q0 = int32_t(sg.group * NT + rank * QH);
        const int64_t kb0 = sg.ch0 * kc, kb1 = sg.ch1 * kc < nkb ? sg.ch1 * kc : nkb;
        for (int64_t kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&emptyq[stage], phase ^ 1u, a.fault);
          if (leader) mbar_arrive_expect_tx(&fullq[stage], 2 * QS_BYTES);
          const uint32_t bar = mapa_shared(smem_u32(&fullq[stage]), 0);
          uint8_t* s = smem_q + stage * QS_BYTES;
          const int32_t k0 = int32_t(kb * BK);
          tma_load_2d_pair(s, &tmQ0, bar, k0, q0, pol_keep);
          tma_load_2d_pair(s + Q_BYTES, &tmQ1, bar, k0, q0, pol_keep);
          if (++stage == SQ) { stage = 0; phase ^= 1u; }
        }
        u += sg.ch1 - sg.ch0;
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA, one thread)
    if (leader && lane == 0) {
      constexpr uint32_t idesc = umma_idesc_f16_m256<NT>();
      int stage = 0; uint32_t phase = 0;
      int qstage = 0; uint32_t qphase = 0;
      int acc = 0; uint32_t acc_phase = 0, c_phase = 0;
      const uint32_t c_tmem = tmem_base + uint32_t(NACC * NT);
      for (int64_t u = lo; u < hi;) {
        const PairSeg sg = pair_seg(u, hi, nchunks, G, NP, U);
        mbar_wait(cempty, c_phase ^ 1u, a.fault);   // both epilogues have read the previous segment's corrections
        c_phase ^= 1u;
        uint32_t c_accumulate = 0;
        for (int64_t ch = sg.ch0; ch < sg.ch1; ++ch) {
          mbar_wait(&tempty[acc], acc_phase ^ 1u, a.fault);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + uint32_t(acc * NT);
          const int64_t kb1 = (ch + 1) * kc < nkb ? (ch + 1) * kc : nkb;
          uint32_t accumulate = 0;
          for (int64_t kb = ch * kc; kb < kb1; ++kb) {
            mbar_wait(&full[stage], phase, a.fault);
            if (PAIR_RINGS) mbar_wait(&fullq[qstage], qphase, a.fault);
            tc_fence_after();
            const uint32_t sbase = smem_u32(smem + stage * PS_BYTES), qbase = smem_u32(smem_q + qstage * QS_BYTES);
            const uint64_t dP0 = umma_desc_k_sw128(sbase), dP1 = umma_desc_k_sw128(sbase + P_BYTES);
            const uint64_t dQ0 = umma_desc_k_sw128(qbase), dQ1 = umma_desc_k_sw128(qbase + Q_BYTES);
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {   // 4 x K = 16 fp16 per 128-byte row
              umma_f16_pair(d_tmem, dP0 + uint64_t(2 * k4), dQ0 + uint64_t(2 * k4), idesc, accumulate);
              accumulate = 1;
            }
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
              umma_f16_pair(c_tmem, dP0 + uint64_t(2 * k4), dQ1 + uint64_t(2 * k4), idesc, c_accumulate);
              c_accumulate = 1;
              umma_f16_pair(c_tmem, dP1 + uint64_t(2 * k4), dQ0 + uint64_t(2 * k4), idesc, 1u);
            }
            umma_commit_pair(&empty[stage]);    // frees the P stage in both CTAs once these MMAs have read it
            if (PAIR_RINGS) umma_commit_pair(&emptyq[qstage]);  // and the Q stage
            if (++stage == SP) { stage = 0; phase ^= 1u; }
            if (++qstage == SQ) { qstage = 0; qphase ^= 1u; }
          }
          umma_commit_pair(&tfull[acc]);       // chunk accumulator (and, after the last chunk, the corrections) complete
          if (++acc == NACC) { acc = 0; acc_phase ^= 1u; }
        }
        u += sg.ch1 - sg.ch0;
      }
    }
  } else if (warp < 2 + PAIR_EPI_WARPS) {
    // ------------------------------------------------------------------ epilogue (16 warps in each CTA)
    const int quarter = warp & 3, cgp = (warp - 2) >> 2;        // TMEM lane quarter (fixed by warp % 4), column group
    if (cgp < NCG) {
    const int etid = (cgp * 4 + quarter) * 32 + lane;           // 0 .. 32 * EPI_ACTIVE - 1 among the active epilogue threads
    uint32_t tempty_l[NACC];
#pragma unroll
    for (int q = 0; q < NACC; ++q) tempty_l[q] = mapa_shared(smem_u32(&tempty[q]), 0);
    const uint32_t cempty_l = mapa_shared(smem_u32(cempty), 0);
    const uint32_t lane_off = uint32_t(quarter * 32) << 16;
    const int lrow = quarter * 32 + lane;                       // row inside this CTA's 128-row tile
    auto epi_bar = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_ACTIVE) : "memory"); };
    // HC columns of one accumulator: all tcgen05.ld in flight before the single wait
    auto drain = [&](uint32_t t0, double (&sum)[HC]) {
      uint32_t v[HC / 16][16];
#pragma unroll
      for (int c0 = 0; c0 < HC / 16; ++c0) tmem_ld_32x32b_x16(t0 + uint32_t(16 * c0), v[c0]);
      tmem_ld_wait();
#pragma unroll
      for (int c0 = 0; c0 < HC / 16; ++c0)
#pragma unroll
        for (int j = 0; j < 16; ++j) sum[16 * c0 + j] += double(__uint_as_float(v[c0][j]));
    };
    int acc = 0; uint32_t acc_phase = 0;
    for (int64_t u = lo; u < hi;) {
      const PairSeg sg = pair_seg(u, hi, nchunks, G, NP, U);
      const int64_t row = (sg.tp * 2 + rank) * GEMM_BM + lrow;
      double sum[HC];
#pragma unroll
      for (int j = 0; j < HC; ++j) sum[j] = 0.0;
      for (int64_t ch = sg.ch0; ch < sg.ch1; ++ch) {
        mbar_wait(&tfull[acc], acc_phase, a.fault);
        tc_fence_after();
        drain(tmem_base + uint32_t(acc * NT + cgp * HC) + lane_off, sum);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(tempty_l[acc]);
        if (++acc == NACC) { acc = 0; acc_phase ^= 1u; }
      }
      {
        drain(tmem_base + uint32_t(NACC * NT + cgp * HC) + lane_off, sum);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(cempty_l);
      }
      bool store = true;
      if (!sg.whole) {
        // ---- a tile shared with other pairs: publish my partial, the last arriver sums all of them in pair order
        double* mine = a.slots + (size_t(2) * blockIdx.x + (sg.ch0 == 0 ? 1 : 0)) * pair_slot_doubles<NT>();
#pragma unroll
        for (int j = 0; j < HC; ++j) mine[size_t(cgp * HC + j) * GEMM_BM + lrow] = sum[j];
        __threadfence();
        epi_bar();
        unsigned* cnt = a.counters + ((sg.tp * 2 + rank) * G + sg.group);
        if (etid == 0) {
          const unsigned prev = atomicAdd(cnt, 1u);
          s_last = prev == unsigned(sg.last - sg.first) ? 1 : 0;
        }
        epi_bar();
        store = s_last != 0;
        if (store) {
          __threadfence();
          if (etid == 0) *cnt = 0u;
          // every partial -- mine included -- is read back from the slots, so that the sum runs over the
          // contributors in pair order (= ascending K) without a second register array
#pragma unroll
          for (int j = 0; j < HC; ++j) sum[j] = 0.0;
          for (int64_t c = sg.first; c <= sg.last; ++c) {
            // contributor c: its range starts inside this tile unless it is the first contributor
            const double* part = a.slots + (size_t(2) * (2 * c + rank) + (c == sg.first ? 1 : 0)) * pair_slot_doubles<NT>();
#pragma unroll
            for (int j = 0; j < HC; ++j) sum[j] += __ldcg(&part[size_t(cgp * HC + j) * GEMM_BM + lrow]);
          }
        }
        epi_bar();   // s_last is reused by the next segment
      }
      if (store && row < a.M) {
        const double sc = a.scale != nullptr ? *a.scale : 1.0;
#pragma unroll
        for (int j = 0; j < HC; ++j) {
          const int col = int(sg.group) * NT + cgp * HC + j;
          if (col < a.N) Dout[int64_t(col) * a.ldd + row] = sum[j] * sc;
        }
      }
      u += sg.ch1 - sg.ch0;
    }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // no CTA leaves while its peer may still signal its barriers or read its shared memory
  if (warp == 1) tmem_dealloc_pair(tmem_base, TMEM_COLS);
}

}  // namespace glp
