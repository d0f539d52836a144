// glp_wide_host.inl -- host side of the wide multi-RHS mode (l >= 16, fp32), included by glprox.cu: the resident
// tensor-core operand sets of A and A^T, and the forward / transposed products of a solver slot on K5.
// ---- wide multi-RHS mode: both products of an iteration run on the tensor cores (K5) ---------------
static bool wide_active(const glp_handle* h) {
  return h->dtype == GLP_F32 && h->g.l >= 16 && 2 * h->g.l <= 256 && h->tensor_mode != 1 && !h->xch_nccl;
}
static int wide_passes(const glp_handle* h) {   // the PASSES of glp_gemm_tf32_kernel
  switch (h->tensor_mode) {
    case 2: return 1;
    case 3: return 2;
    case 4: return 4;
    case 5: return 3;
    default: return h->f16_unusable ? 3 : 4;   // auto: fp16 pairs unless A cannot be scaled (non-finite entries)
  }
}
static int gemm_splits(int64_t M, int64_t K, int N, int k_chunk, int num_sms, int bk = GEMM_BK);
constexpr int WIDE_K_CHUNK = 8;  // the tensor core's fp32 accumulator spans 8 x 32 = 256 elements of K, then fp64

struct GemmGate { const int32_t* count; int32_t cap; };            // run only when *count > cap (nullptr: always)
static int split_f32(cudaStream_t st, const float* src, int64_t rows, int64_t cols, int64_t ld_src, float* hi, float* lo,
                     int64_t ld_dst, bool transpose, GemmGate gate = GemmGate{nullptr, 0}) {
  if (transpose) {
    dim3 grid(unsigned((cols + 31) / 32), unsigned((rows + 31) / 32));
    if (grid.y > 65535u) return fail(GLP_ERR_ARG, "too many rows for the transposing split");
    glp_transpose_split_kernel<<<grid, dim3(32, 8), 0, st>>>(src, hi, lo, rows, cols, ld_src, ld_dst);
  } else {
    const int grid = int(std::min<int64_t>((rows * cols + 255) / 256, 148 * 16));
    glp_split_tf32_kernel<<<grid, 256, 0, st>>>(src, hi, lo, rows, cols, ld_src, ld_dst, gate.count, gate.cap);
  }
  CU_TRY(cudaGetLastError());
  return GLP_OK;
}

static int split_bf16(cudaStream_t st, const float* src, int64_t rows, int64_t cols, int64_t ld_src, float* hi,
                      __nv_bfloat16* hb, __nv_bfloat16* lb, int64_t ld_dst, bool transpose,
                      GemmGate gate = GemmGate{nullptr, 0}) {
  if (transpose) {
    dim3 grid(unsigned((cols + 31) / 32), unsigned((rows + 31) / 32));
    if (grid.y > 65535u) return fail(GLP_ERR_ARG, "too many rows for the transposing split");
    glp_transpose_split_bf16_kernel<<<grid, dim3(32, 8), 0, st>>>(src, hi, hb, lb, rows, cols, ld_src, ld_dst);
  } else {
    const int grid = int(std::min<int64_t>((rows * cols + 255) / 256, 148 * 16));
    glp_split_bf16_kernel<<<grid, 256, 0, st>>>(src, hi, hb, lb, rows, cols, ld_src, ld_dst, gate.count, gate.cap);
  }
  CU_TRY(cudaGetLastError());
  return GLP_OK;
}

template <typename Q>
static int alloc_zero(Q** p, size_t bytes, cudaStream_t st) {
  CU_TRY(cudaMalloc(reinterpret_cast<void**>(p), bytes));
  CU_TRY(cudaMemsetAsync(*p, 0, bytes, st));
  return GLP_OK;
}

// One-time set-up per operand set, resident in HBM afterwards ("memory laid out for 180 GB"), plus the matching
// halves of the two small operands:
//   fp16 pairs (default):  A s = A_f0 + A_f1 and A^T s = At_f0 + At_f1 in fp16, s from max|A|: 2 x |A| in all
//   3xTF32:                A = A_hi + A_lo and A^T = At_hi + At_lo in fp32: 4 x |A|
//   TF32 + bf16:           fp32 hi plus bf16 copies of hi and lo: 4 x |A|
static int ensure_wide(glp_handle* h) {
  const Geom& g = h->g;
  const size_t a_el = size_t(g.m) * g.ldn, at_el = size_t(g.n) * g.ldm;
  const size_t x_el = size_t(2) * g.l * g.ldn, r_el = size_t(g.l) * g.ldm;
  const float* A = static_cast<const float*>(h->A);
  if (h->cnt_g == nullptr) {   // shared by all operand sets
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->cnt_g), size_t(h->gather_grid) * 8));
    // A.xnt by sparse correction: on unless GLPROX_WIDE_THRFIX=0.  Room for n/32 entries per operand column
    // (>= 64): ~30 x the expected load at thres = 1e-3, and at that load the correction still costs well under
    // half of the product it replaces.  GLPROX_WIDE_THRFIX_CAP overrides (tests).
    const char* e_fix = getenv("GLPROX_WIDE_THRFIX");
    if (e_fix == nullptr || atoi(e_fix) != 0) {
      int64_t cap = std::max<int64_t>(64, g.n / 32);
      if (const char* e_cap = getenv("GLPROX_WIDE_THRFIX_CAP")) cap = std::max(0, atoi(e_cap));
      const size_t el = size_t(std::max<int64_t>(cap, 1)) * g.l;
      GLP_TRY(alloc_zero(&h->thr_rows, el * 4, h->stream));
      GLP_TRY(alloc_zero(&h->thr_vals, el * 4, h->stream));
      GLP_TRY(alloc_zero(&h->thr_col_count, size_t(g.l) * 4, h->stream));
      GLP_TRY(alloc_zero(&h->thr_count, 4 * 4, h->stream));
      h->thr_cap = int32_t(cap);
    }
    // split-K workspace for the larger of the products
    size_t fwd = size_t(gemm_splits(g.m, g.n, 2 * g.l, WIDE_K_CHUNK, h->num_sms)) * 2 * g.l * g.ldm * 8;
    fwd = std::max(fwd, size_t(gemm_splits(g.m, g.n, g.l, WIDE_K_CHUNK, h->num_sms)) * g.l * g.ldm * 8);
    const size_t tps = size_t(gemm_splits(g.n, g.m, g.l, WIDE_K_CHUNK, h->num_sms)) * g.l * g.ldn * 8;
    h->gemm_ws_bytes = std::max(fwd, tps);
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->gemm_ws), h->gemm_ws_bytes));
    h->pair_ws_bytes = pair_ws_bytes(h->num_sms, 128);
    GLP_TRY(alloc_zero(&h->pair_ws, h->pair_ws_bytes, h->stream));
  }
  if (wide_passes(h) == 4) {
    if (h->A_f0 != nullptr) return GLP_OK;
    // scale of the resident operand from max|A| (one synchronous read-back at set-up)
    if (h->f16_amax == nullptr) {
      GLP_TRY(alloc_zero(&h->f16_amax, 4 * sizeof(unsigned), h->stream));
      GLP_TRY(alloc_zero(&h->f16_inv, 2 * sizeof(double), h->stream));
    }
    CU_TRY(cudaMemsetAsync(h->f16_amax + 2, 0, sizeof(unsigned), h->stream));
    const int ag = int(std::min<int64_t>((int64_t(g.m) * g.n + 255) / 256, int64_t(h->num_sms) * 16));
    glp_absmax_kernel<<<ag, 256, 0, h->stream>>>(A, g.m, g.n, h->lda, h->f16_amax + 2);
    CU_TRY(cudaGetLastError());
    unsigned bits = 0;
    CU_TRY(cudaMemcpyAsync(&bits, h->f16_amax + 2, sizeof(bits), cudaMemcpyDeviceToHost, h->stream));
    CU_TRY(cudaStreamSynchronize(h->stream));
    float amax;
    memcpy(&amax, &bits, sizeof(amax));
    if (!(amax < 3.0e38f)) {
      if (h->tensor_mode != 0) return fail(GLP_ERR_ARG, "A holds non-finite values: the fp16-pair operand set cannot be scaled");
      h->f16_unusable = true;   // auto: propagate them through the 3xTF32 operand set like any fp32 product would
      return ensure_wide(h);
    }
    // dynamic range of the column / row scales of A: beyond 2^18 the fp16 pairs (2^-40 of the matrix maximum) fall
    // below fp32 accuracy on the small columns -> auto takes the 3xTF32 operand set, an explicit 'f16x2' is refused
    {
      unsigned* rc = nullptr;
      CU_TRY(cudaMalloc(reinterpret_cast<void**>(&rc), size_t(g.m + g.n + 4) * sizeof(unsigned)));
      CU_TRY(cudaMemsetAsync(rc, 0, size_t(g.m + g.n + 4) * sizeof(unsigned), h->stream));
      const unsigned init[4] = {0u, 0x7f800000u, 0u, 0x7f800000u};
      CU_TRY(cudaMemcpyAsync(rc + g.m + g.n, init, sizeof(init), cudaMemcpyHostToDevice, h->stream));
      glp_rowcol_absmax_kernel<<<ag, 256, 0, h->stream>>>(A, g.m, g.n, h->lda, rc, rc + g.n);
      glp_minmax_nonzero_kernel<<<64, 256, 0, h->stream>>>(rc, g.n, rc + g.m + g.n);
      glp_minmax_nonzero_kernel<<<64, 256, 0, h->stream>>>(rc + g.n, g.m, rc + g.m + g.n + 2);
      unsigned mm[4] = {0, 0, 0, 0};
      cudaError_t e1 = cudaMemcpyAsync(mm, rc + g.m + g.n, sizeof(mm), cudaMemcpyDeviceToHost, h->stream);
      if (e1 == cudaSuccess) e1 = cudaStreamSynchronize(h->stream);
      cudaFree(rc);
      if (e1 != cudaSuccess) return fail(GLP_ERR_CUDA, "dynamic-range probe failed: %s", cudaGetErrorString(e1));
      float fv[4];
      memcpy(fv, mm, sizeof(fv));
      const float range = std::max(fv[1] > 0.f ? fv[0] / fv[1] : 1.f, fv[3] > 0.f ? fv[2] / fv[3] : 1.f);
      if (range > 262144.0f) {
        if (h->tensor_mode != 0)
          return fail(GLP_ERR_ARG, "the column / row scales of A span a factor %.3g > 2^18: the fp16-pair operand set would lose "
                                   "the small ones; use tensor_cores 'auto' or '3xtf32'", double(range));
        h->f16_unusable = true;
        return ensure_wide(h);
      }
    }
    h->f16_sA = f16_scale_for(amax);
    // the CTA-pair kernel streams a tiled operand (GLPROX_GEMM_PAIR=0 keeps the row-major set of the one-CTA kernel)
    const char* e_pair = getenv("GLPROX_GEMM_PAIR");
    h->f16_tiled = e_pair == nullptr || atoi(e_pair) != 0;
    // K blocks of the tiled sets: what the product kernel derives from its K (n for A, m for A^T)
    const int64_t nkb_a = (g.n + 63) / 64, nkb_at = (g.m + 63) / 64;
    const size_t a_tiled = size_t((g.m + 127) / 128) * 128 * size_t(nkb_a) * 64;
    const size_t at_tiled = size_t((g.n + 127) / 128) * 128 * size_t(nkb_at) * 64;
    GLP_TRY(alloc_zero(&h->A_f0, (h->f16_tiled ? a_tiled : a_el) * 2, h->stream));
    GLP_TRY(alloc_zero(&h->A_f1, (h->f16_tiled ? a_tiled : a_el) * 2, h->stream));
    GLP_TRY(alloc_zero(&h->At_f0, (h->f16_tiled ? at_tiled : at_el) * 2, h->stream));
    GLP_TRY(alloc_zero(&h->At_f1, (h->f16_tiled ? at_tiled : at_el) * 2, h->stream));
    GLP_TRY(alloc_zero(&h->xop_f0, x_el * 2, h->stream));
    GLP_TRY(alloc_zero(&h->xop_f1, x_el * 2, h->stream));
    GLP_TRY(alloc_zero(&h->r_f0, r_el * 2, h->stream));
    GLP_TRY(alloc_zero(&h->r_f1, r_el * 2, h->stream));
    dim3 tgrid(unsigned((g.n + 31) / 32), unsigned((g.m + 31) / 32));
    if (tgrid.y > 65535u) return fail(GLP_ERR_ARG, "too many rows for the transposing split");
    if (h->f16_tiled) {
      glp_split_f16_tiled_kernel<<<ag, 256, 0, h->stream>>>(A, h->A_f0, h->A_f1, g.m, g.n, h->lda, nkb_a, h->f16_sA);
      glp_transpose_split_f16_tiled_kernel<<<tgrid, dim3(32, 8), 0, h->stream>>>(A, h->At_f0, h->At_f1, g.m, g.n, h->lda,
                                                                                 nkb_at, h->f16_sA);
    } else {
      glp_split_f16_2d_kernel<<<ag, 256, 0, h->stream>>>(A, h->A_f0, h->A_f1, g.m, g.n, h->lda, g.ldn, h->f16_sA);
      glp_transpose_split_f16_kernel<<<tgrid, dim3(32, 8), 0, h->stream>>>(A, h->At_f0, h->At_f1, g.m, g.n, h->lda, g.ldm, h->f16_sA);
    }
    CU_TRY(cudaGetLastError());
    return GLP_OK;
  }
  if (h->A_hi == nullptr) {
    GLP_TRY(alloc_zero(&h->A_hi, a_el * 4, h->stream));
    GLP_TRY(alloc_zero(&h->At_hi, at_el * 4, h->stream));
  }
  if (wide_passes(h) == 2) {
    if (h->A_hb == nullptr) {
      GLP_TRY(alloc_zero(&h->A_hb, a_el * 2, h->stream));
      GLP_TRY(alloc_zero(&h->A_lb, a_el * 2, h->stream));
      GLP_TRY(alloc_zero(&h->At_hb, at_el * 2, h->stream));
      GLP_TRY(alloc_zero(&h->At_lb, at_el * 2, h->stream));
      GLP_TRY(alloc_zero(&h->xop_hb, x_el * 2, h->stream));
      GLP_TRY(alloc_zero(&h->xop_lb, x_el * 2, h->stream));
      GLP_TRY(alloc_zero(&h->r_hb, r_el * 2, h->stream));
      GLP_TRY(alloc_zero(&h->r_lb, r_el * 2, h->stream));
      GLP_TRY(split_bf16(h->stream, A, g.m, g.n, h->lda, h->A_hi, h->A_hb, h->A_lb, g.ldn, false));
      GLP_TRY(split_bf16(h->stream, A, g.m, g.n, h->lda, h->At_hi, h->At_hb, h->At_lb, g.ldm, true));
    }
  } else if (h->A_lo == nullptr) {
    GLP_TRY(alloc_zero(&h->A_lo, a_el * 4, h->stream));
    GLP_TRY(alloc_zero(&h->At_lo, at_el * 4, h->stream));
    GLP_TRY(alloc_zero(&h->xop_lo, x_el * 4, h->stream));
    GLP_TRY(alloc_zero(&h->r_lo, r_el * 4, h->stream));
    GLP_TRY(split_f32(h->stream, A, g.m, g.n, h->lda, h->A_hi, h->A_lo, g.ldn, false));
    GLP_TRY(split_f32(h->stream, A, g.m, g.n, h->lda, h->At_hi, h->At_lo, g.ldm, true));
  }
  return GLP_OK;
}

// split-K partial products (buf == nullptr: no splitting) and the workspace of the CTA-pair kernel (pair == nullptr:
// the one-CTA kernel runs instead)
struct GemmWs { double* buf; size_t bytes; void* pair = nullptr; size_t pair_bytes = 0; bool p_tiled = false; bool pdl = false; };
struct GemmBf16 { const __nv_bfloat16 *Phb, *Plb, *Qhb, *Qlb; };  // passes = 2: bf16 copies of hi / lo, same pitches
static int gemm_splits(int64_t M, int64_t K, int N, int k_chunk, int num_sms, int bk);
static int gemm_enqueue(cudaStream_t st, int num_sms, int passes, int64_t M, int64_t K, int N, const float* Ph,
                        const float* Pl, int64_t ldp, const float* Qh, const float* Ql, int64_t ldq, double* D,
                        int64_t ldd, int k_chunk, int32_t* fault, const Ctrl* ctrl = nullptr, double* D_alt = nullptr,
                        int role = 0, GemmWs ws = GemmWs{nullptr, 0},
                        GemmBf16 bf = GemmBf16{nullptr, nullptr, nullptr, nullptr}, GemmGate gate = GemmGate{nullptr, 0},
                        const double* scale = nullptr);

static uint64_t seq_base(const glp_handle* h);

// Multi-GPU wide mode (glp_wide_comm.cuh): stage 0 push (right after the local product), 1 reduce, 2 finish.
// One device-wide stage per phase, so that a multi-rank job on ONE device can enqueue each stage for all
// of its logical ranks before the next.
static int enqueue_wide_exchange(glp_handle* h, const Ctrl* ctrl, int stage) {
  const Geom& g = h->g;
  WideXArgs a{};
  a.comm = h->comm;
  a.ctrl = ctrl;
  a.sel = 0;
  a.par = h->group & 1;
  a.ncols = 2 * g.l;
  a.slice_rows = int32_t(((g.ldm + h->comm.world - 1) / h->comm.world + 31) / 32 * 32);
  if (stage == 0) ++h->idx_f;
  a.seq_f = seq_base(h) | h->idx_f;
  a.seq_t = h->seq_t_cur;
  a.AX[0] = h->AX[0]; a.AX[1] = h->AX[1];
  a.b = h->b;
  a.xcand[0] = h->state[0]; a.xcand[1] = h->state[1];
  a.cnt_partial = h->cnt_g;
  a.rr_partial = h->rr_g;
  a.scal_red = h->scal_red;
  a.cnt_red = h->cnt_red;
  a.fin = h->fin_f;
  a.fault = h->fault;
  if (stage == 0) glp_wide_push_kernel<<<h->gather_grid, WIDEX_THREADS, 0, h->stream>>>(a, g);
  else if (stage == 1) glp_wide_reduce_kernel<<<h->gather_grid, WIDEX_THREADS, 0, h->stream>>>(a, g);
  else glp_wide_finish_kernel<<<h->gather_grid, WIDEX_THREADS, 0, h->stream>>>(a, g);
  CU_TRY(cudaGetLastError());
  h->n_launches++;
  return GLP_OK;
}

static cudaError_t launch_split_f16(const glp_handle* h, int grid, const float* src, __half* o0, __half* o1, int64_t total,
                                    const unsigned* amax_bits, const double* scal_partial, int ncb, float s_other,
                                    const unsigned* other_amax_bits, double* inv_scale_out, const int32_t* gate, int gate_cap) {
  return launch_k(h, glp_split_f16_kernel, dim3(grid), dim3(256), 0, src, o0, o1, total, amax_bits, scal_partial, ncb, s_other,
                  other_amax_bits, inv_scale_out, gate, gate_cap);
}

// forward product of the solver on the tensor cores: AX[1-cur] = A . [xn | xnt], then ||r||^2 and the count
static int enqueue_wide_fwd(glp_handle* h, const Ctrl* ctrl) {
  const Geom& g = h->g;
  float* xop = static_cast<float*>(h->xop);
  const bool bf = wide_passes(h) == 2;
  if (h->thr_cap >= 0) {   // list the rows the hard threshold changed, before the split overwrites the operand
    ThrListArgs la{};
    la.ctrl = ctrl;
    la.xop = xop;
    la.rows = h->thr_rows; la.vals = h->thr_vals; la.col_count = h->thr_col_count; la.count = h->thr_count;
    la.cap = h->thr_cap;
    CU_TRY(launch_k(h, glp_thr_list_kernel, dim3(g.l), dim3(THRLIST_THREADS), 0, la, g));
    h->n_launches++;
  }
  const bool f16 = wide_passes(h) == 4;
  const int64_t q2 = int64_t(g.l) * g.ldn;
  const GemmGate xnt_gate{h->thr_count + 1, 0};
  if (f16) {
    // fp16 pairs of xn s_x; s_x from max|xn|, which the prox epilogue left per column block in scal_partial
    // (max|xnt| <= max|xn|: one scale for both halves)
    const int sg = int(std::min<int64_t>((q2 + 1023) / 1024, int64_t(h->num_sms) * 16));
    if (h->thr_cap < 0) {
      CU_TRY(launch_split_f16(h, sg, xop, h->xop_f0, h->xop_f1, 2 * q2, nullptr, h->scal_partial, g.ncb, h->f16_sA, nullptr,
                              h->f16_inv, nullptr, 0));
    } else {
      CU_TRY(launch_split_f16(h, sg, xop, h->xop_f0, h->xop_f1, q2, nullptr, h->scal_partial, g.ncb, h->f16_sA, nullptr,
                              h->f16_inv, nullptr, 0));
      CU_TRY(launch_split_f16(h, sg, xop + q2, h->xop_f0 + q2, h->xop_f1 + q2, q2, nullptr, h->scal_partial, g.ncb,
                              h->f16_sA, nullptr, nullptr, xnt_gate.count, xnt_gate.cap));
      h->n_launches++;
    }
    h->n_launches++;
  } else if (h->thr_cap < 0) {
    if (bf) GLP_TRY(split_bf16(h->stream, xop, 2 * g.l, g.ldn, g.ldn, xop, h->xop_hb, h->xop_lb, g.ldn, false));
    else GLP_TRY(split_f32(h->stream, xop, 2 * g.l, g.ldn, g.ldn, xop, h->xop_lo, g.ldn, false));  // hi in place
  } else {   // the xnt half is only needed by the gated fallback product
    if (bf) {
      GLP_TRY(split_bf16(h->stream, xop, g.l, g.ldn, g.ldn, xop, h->xop_hb, h->xop_lb, g.ldn, false));
      GLP_TRY(split_bf16(h->stream, xop + q2, g.l, g.ldn, g.ldn, xop + q2, h->xop_hb + q2, h->xop_lb + q2, g.ldn, false, xnt_gate));
    } else {
      GLP_TRY(split_f32(h->stream, xop, g.l, g.ldn, g.ldn, xop, h->xop_lo, g.ldn, false));
      GLP_TRY(split_f32(h->stream, xop + q2, g.l, g.ldn, g.ldn, xop + q2, h->xop_lo + q2, g.ldn, false, xnt_gate));
    }
    h->n_launches++;
  }
  prof_mark(h, 0, 0);
  const GemmWs ws{h->gemm_ws, h->gemm_ws_bytes, h->pair_ws, h->pair_ws_bytes, f16 && h->f16_tiled, h->pdl};
  // operand pointers of the selected set (fp16 pairs travel as float* through gemm_enqueue, pitches in elements)
  const float* Ph = f16 ? reinterpret_cast<const float*>(h->A_f0) : h->A_hi;
  const float* Pl = f16 ? reinterpret_cast<const float*>(h->A_f1) : h->A_lo;
  auto qh = [&](int64_t off) { return f16 ? reinterpret_cast<const float*>(h->xop_f0 + off) : xop + off; };
  auto ql = [&](int64_t off) {
    return f16 ? reinterpret_cast<const float*>(h->xop_f1 + off) : (bf ? static_cast<const float*>(nullptr) : h->xop_lo + off);
  };
  auto qb = [&](int64_t off) {
    return GemmBf16{h->A_hb, h->A_lb, bf ? h->xop_hb + off : nullptr, bf ? h->xop_lb + off : nullptr};
  };
  const double* sc = f16 ? h->f16_inv : nullptr;
  const int kch = f16 ? WIDE_K_CHUNK / 2 : WIDE_K_CHUNK;   // 256 elements of K per fp32 accumulation either way
  if (h->thr_cap < 0) {   // one product over both operand halves
    GLP_TRY(gemm_enqueue(h->stream, h->num_sms, wide_passes(h), g.m, g.n, 2 * g.l, Ph, Pl, g.ldn, qh(0), ql(0),
                         g.ldn, h->AX[0], g.ldm, kch, h->fault, ctrl, h->AX[1], 1, ws, qb(0), GemmGate{nullptr, 0}, sc));
  } else {
    // A.xn on the tensor cores; A.xnt = A.xn - A.(xn - xnt) on the CUDA cores from the listed entries, or, when
    // a list overflowed, by a second product launch that the device gates on the overflow count
    const int64_t d2 = int64_t(g.l) * g.ldm;
    GLP_TRY(gemm_enqueue(h->stream, h->num_sms, wide_passes(h), g.m, g.n, g.l, Ph, Pl, g.ldn, qh(0), ql(0),
                         g.ldn, h->AX[0], g.ldm, kch, h->fault, ctrl, h->AX[1], 1, ws, qb(0), GemmGate{nullptr, 0}, sc));
    GLP_TRY(gemm_enqueue(h->stream, h->num_sms, wide_passes(h), g.m, g.n, g.l, Ph, Pl, g.ldn, qh(q2), ql(q2),
                         g.ldn, h->AX[0] + d2, g.ldm, kch, h->fault, ctrl, h->AX[1] + d2, 1, ws, qb(q2), xnt_gate, sc));
    ThrFixArgs fa{};
    fa.ctrl = ctrl;
    fa.count = h->thr_count; fa.col_count = h->thr_col_count;
    fa.rows = h->thr_rows; fa.vals = h->thr_vals; fa.cap = h->thr_cap;
    fa.At_hi = h->At_hi; fa.At_lo = bf ? nullptr : h->At_lo; fa.At_lb = bf ? h->At_lb : nullptr;
    if (f16) {
      fa.At_hi = nullptr; fa.At_lo = nullptr; fa.At_lb = nullptr;
      fa.At_f0 = h->At_f0; fa.At_f1 = h->At_f1; fa.inv_s = 1.0f / h->f16_sA;
      fa.thres = float(h->sc.thres); fa.inv_x = h->f16_inv;
      fa.tiled_nkb = h->f16_tiled ? (g.m + 63) / 64 : 0;   // == ldm / 64: the padded output rows stay inside the last tile
    }
    fa.AX[0] = h->AX[0]; fa.AX[1] = h->AX[1];
    const int64_t frows = int64_t(THRFIX_THREADS) * THRFIX_ROWS;
    const dim3 fgrid(unsigned((g.ldm + frows - 1) / frows), unsigned(g.l));
    CU_TRY(launch_k(h, glp_thr_fix_kernel, fgrid, dim3(THRFIX_THREADS), 0, fa, g));
    h->n_launches += 2;
  }
  prof_mark(h, 0, 1);
  h->n_launches += 2;
  if (multi(h)) return enqueue_wide_exchange(h, ctrl, 0);
  PostFwdArgs pa{};
  pa.ctrl = ctrl;
  pa.AX[0] = h->AX[0]; pa.AX[1] = h->AX[1];
  pa.b = h->b;
  pa.scal_partial = h->scal_partial;
  pa.xcand[0] = h->state[0]; pa.xcand[1] = h->state[1];
  pa.rr_out = h->rr_g;
  pa.cnt_out = h->cnt_g;
  CU_TRY(launch_k(h, glp_post_fwd_kernel, dim3(h->gather_grid), dim3(256), 0, pa, g));
  h->n_launches += 2;
  h->n_gemm++;
  return GLP_OK;
}

// transposed product of the solver on the tensor cores: grad = A^T . r  (skipped by the kernel on a retry)
static int enqueue_wide_tpass(glp_handle* h, const Ctrl* ctrl) {
  const Geom& g = h->g;
  float* r = static_cast<float*>(h->r);
  const GemmWs ws{h->gemm_ws, h->gemm_ws_bytes, h->pair_ws, h->pair_ws_bytes, wide_passes(h) == 4 && h->f16_tiled, h->pdl};
  if (wide_passes(h) == 4) {
    const int64_t tot = int64_t(g.l) * g.ldm;
    // max|r| was left in f16_amax[1] by the decide kernel that wrote r
    const int sg = int(std::min<int64_t>((tot + 1023) / 1024, int64_t(h->num_sms) * 16));
    CU_TRY(launch_split_f16(h, sg, r, h->r_f0, h->r_f1, tot, h->f16_amax + 1, nullptr, 0, h->f16_sA, nullptr, h->f16_inv + 1,
                            nullptr, 0));
    h->n_launches++;
    prof_mark(h, 1, 0);
    GLP_TRY(gemm_enqueue(h->stream, h->num_sms, 4, g.n, g.m, g.l, reinterpret_cast<const float*>(h->At_f0),
                         reinterpret_cast<const float*>(h->At_f1), g.ldm, reinterpret_cast<const float*>(h->r_f0),
                         reinterpret_cast<const float*>(h->r_f1), g.ldm, h->grad, g.ldn, WIDE_K_CHUNK / 2, h->fault, ctrl,
                         nullptr, 2, ws, GemmBf16{nullptr, nullptr, nullptr, nullptr}, GemmGate{nullptr, 0}, h->f16_inv + 1));
  } else {
    if (wide_passes(h) == 2) GLP_TRY(split_bf16(h->stream, r, g.l, g.ldm, g.ldm, r, h->r_hb, h->r_lb, g.ldm, false));
    else GLP_TRY(split_f32(h->stream, r, g.l, g.ldm, g.ldm, r, h->r_lo, g.ldm, false));
    prof_mark(h, 1, 0);
    GLP_TRY(gemm_enqueue(h->stream, h->num_sms, wide_passes(h), g.n, g.m, g.l, h->At_hi, h->At_lo, g.ldm, r, h->r_lo,
                         g.ldm, h->grad, g.ldn, WIDE_K_CHUNK, h->fault, ctrl, nullptr, 2, ws,
                         GemmBf16{h->At_hb, h->At_lb, h->r_hb, h->r_lb}));
  }
  prof_mark(h, 1, 1);
  h->n_launches += 2;
  h->n_gemm++;
  return GLP_OK;
}
