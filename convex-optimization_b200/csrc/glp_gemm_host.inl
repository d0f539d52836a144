// glp_gemm_host.inl -- host side of K5, included by glprox.cu: tensor maps, launch of the one-CTA and the CTA-pair
// kernels, split-K bookkeeping, and the step-level entry points glp_gemm_tf32 / glp_tf32_split.
// ------------------------------------------------------------------------------------------------
// K5: tensor-core products (glp_gemm.cuh)
// ------------------------------------------------------------------------------------------------
static int make_map_f32_k128(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t stride_bytes,
                             uint32_t box_outer) {
  PFN_encodeTiled fn = encode_tiled_fn();
  if (fn == nullptr) return fail(GLP_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  const cuuint64_t dims[2] = {inner, outer};
  const cuuint64_t strides[1] = {stride_bytes};
  const cuuint32_t box[2] = {cuuint32_t(GEMM_BK), box_outer};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims,
                                                     strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(GLP_ERR_CUDA, "cuTensorMapEncodeTiled (gemm operand) failed with CUresult %d", int(r));
  return GLP_OK;
}

// fp16 operand tile of the fp16-pair mode: 64 elements = 128 bytes of K per row, 128-byte swizzle
static int make_map_f16_k128(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t stride_bytes,
                             uint32_t box_outer) {
  PFN_encodeTiled fn = encode_tiled_fn();
  if (fn == nullptr) return fail(GLP_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  const cuuint64_t dims[2] = {inner, outer};
  const cuuint64_t strides[1] = {stride_bytes};
  const cuuint32_t box[2] = {cuuint32_t(2 * GEMM_BK), box_outer};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(GLP_ERR_CUDA, "cuTensorMapEncodeTiled (fp16 operand) failed with CUresult %d", int(r));
  return GLP_OK;
}

// bf16 operand tile: 32 elements = 64 bytes of K per row, 64-byte swizzle
static int make_map_bf16_k64(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t stride_bytes,
                             uint32_t box_outer) {
  PFN_encodeTiled fn = encode_tiled_fn();
  if (fn == nullptr) return fail(GLP_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  const cuuint64_t dims[2] = {inner, outer};
  const cuuint64_t strides[1] = {stride_bytes};
  const cuuint32_t box[2] = {cuuint32_t(GEMM_BK), box_outer};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(GLP_ERR_CUDA, "cuTensorMapEncodeTiled (bf16 operand) failed with CUresult %d", int(r));
  return GLP_OK;
}

template <int NT, int PASSES, bool RINGS = false>
static cudaError_t launch_gemm_t(cudaStream_t st, int num_sms, int64_t cluster_items, const CUtensorMap& ph,
                                 const CUtensorMap& pl, const CUtensorMap& qh, const CUtensorMap& ql,
                                 const CUtensorMap& px, const CUtensorMap& qx, const GemmArgs& a) {
  auto kfn = glp_gemm_tf32_kernel<NT, PASSES, RINGS>;
  constexpr int smem = RINGS ? gemm_ring_smem_bytes<NT>() : gemm_smem_bytes<NT, PASSES>();
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  const int cs = std::max(1, a.ct) * std::max(1, a.cg);
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(RINGS ? GEMM_RING_THREADS : GEMM_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = unsigned(cs);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = cs > 1 ? 1 : 0;
  int64_t ncl = num_sms / cs;
  if (cs > 1) {  // clusters must fit inside a GPC: ask how many can be resident at once
    cfg.gridDim = dim3(unsigned(ncl * cs), 1, 1);
    int max_cl = 0;
    e = cudaOccupancyMaxActiveClusters(&max_cl, kfn, &cfg);
    if (e != cudaSuccess) return e;
    if (max_cl < 1) return cudaErrorInvalidConfiguration;
    ncl = std::min<int64_t>(ncl, max_cl);
  }
  ncl = std::max<int64_t>(1, std::min<int64_t>(ncl, cluster_items));
  cfg.gridDim = dim3(unsigned(ncl * cs), 1, 1);
  return cudaLaunchKernelEx(&cfg, kfn, ph, pl, qh, ql, px, qx, a);
}

// K5 on CTA pairs (glp_gemm_pair.cuh): clusters of 2, one pair per TPC, stream-K ranges over the pairs.
template <int NT>
static cudaError_t launch_gemm_pair_t(cudaStream_t st, int num_sms, const CUtensorMap& p0, const CUtensorMap& p1,
                                      const CUtensorMap& q0, const CUtensorMap& q1, const PairArgs& a, bool pdl) {
  auto kfn = glp_gemm_pair_kernel<NT>;
  constexpr int smem = pair_smem_bytes<NT>();
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(PAIR_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int pairs = num_sms / 2;
  cfg.gridDim = dim3(unsigned(2 * pairs), 1, 1);
  int max_cl = 0;
  e = cudaOccupancyMaxActiveClusters(&max_cl, kfn, &cfg);
  if (e != cudaSuccess) return e;
  if (max_cl < 1) return cudaErrorInvalidConfiguration;
  pairs = std::min(pairs, max_cl);
  cfg.gridDim = dim3(unsigned(2 * pairs), 1, 1);
  cfg.numAttrs = pdl ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kfn, p0, p1, q0, q1, a);
}

// Number of K splits.  Measured on B200 (tests/bench_gemm_probe.py split): the kernel runs under the
// board's power cap, so filling the SMs left idle by the tail wave buys nothing at 128+ items (3.15 ms
// unsplit vs 3.56 ms with 4 splits at M=16384 K=65536 N=256) and fine splits cost epilogue + workspace
// traffic.  K is therefore split only when the items cover less than half of the SMs (column shards of a
// multi-GPU job), and then just enough to occupy them, keeping >= 8 accumulation chunks per item.
static int gemm_splits(int64_t M, int64_t K, int N, int k_chunk, int num_sms, int bk) {
  const int NT = N <= 32 ? 32 : (N <= 64 ? 64 : 128);
  const int64_t base = ((M + GEMM_BM - 1) / GEMM_BM) * ((N + NT - 1) / NT);
  const int64_t nkb = (K + bk - 1) / bk;
  const int64_t kc = k_chunk > 0 ? k_chunk : nkb;
  const int64_t nchunks = (nkb + kc - 1) / kc;
  if (const char* e = getenv("GLPROX_GEMM_SPLITS")) {   // measurement override
    const int64_t S = atoi(e);
    return int(std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(S, 16), nchunks)));
  }
  if (base * 2 > num_sms) return 1;
  int64_t S = num_sms / base;
  S = std::min<int64_t>(S, std::max<int64_t>(1, nchunks / 8));
  return int(std::max<int64_t>(1, std::min<int64_t>(S, 16)));
}

// D[N][ldd] = P[M x K] . Q[N x K]^T ; passes 1 (TF32), 3 (hi/lo compensated), 2 (bf16 corrections) or 4 (fp16 pairs:
// Ph / Pl / Qh / Ql then point at __half arrays v0 / v1 with pitches in fp16 elements, and the product is multiplied
// by *scale); N <= 256 (groups of <= 128 columns)
static int gemm_enqueue(cudaStream_t st, int num_sms, int passes, int64_t M, int64_t K, int N, const float* Ph,
                        const float* Pl, int64_t ldp, const float* Qh, const float* Ql, int64_t ldq, double* D,
                        int64_t ldd, int k_chunk, int32_t* fault, const Ctrl* ctrl, double* D_alt, int role, GemmWs ws,
                        GemmBf16 bf, GemmGate gate, const double* scale) {
  if (N < 1 || N > 256) return fail(GLP_ERR_ARG, "gemm: N must be 1..256 per launch (got %d)", N);
  if (passes < 1 || passes > 4) return fail(GLP_ERR_ARG, "gemm: passes must be 1, 2, 3 or 4");
  const int es = passes == 4 ? 2 : 4, bk = passes == 4 ? 2 * GEMM_BK : GEMM_BK;
  if ((ldp * es) % 16 || (ldq * es) % 16 || (reinterpret_cast<uintptr_t>(Ph) % 16) || (reinterpret_cast<uintptr_t>(Qh) % 16))
    return fail(GLP_ERR_ARG, "gemm: operands must be 16-byte aligned with 16-byte aligned pitches");
  const int NT = N <= 32 ? 32 : (N <= 64 ? 64 : 128);
  const int64_t tiles = (M + GEMM_BM - 1) / GEMM_BM, groups = (N + NT - 1) / NT;
  // cluster geometry (TMA multicast of the shared operand tiles): neighbouring M tiles share every Q tile,
  // the 2 column groups of N = 256 share every P tile.  Measured in the solver at C5 (it/s, fwd / tpass ms):
  // 1x1 171.6 (3.45 / 1.78), 2x2 176.0 (3.34 / 1.76), 4 tiles 154.7 (4.28 with 8-CTA clusters / 1.61): 2 x 2 for
  // the two-group forward product, 4 x 1 for the one-group transposed product.
  // GLPROX_GEMM_CLUSTER=ct,cg overrides (1,1 = off).
  int cg = groups == 2 ? 2 : 1;
  int ct = cg == 2 ? (tiles >= 2 ? 2 : 1) : (tiles >= 4 ? 4 : (tiles >= 2 ? 2 : 1));
  if (const char* e = getenv("GLPROX_GEMM_CLUSTER")) {
    int a1 = 0, a2 = 0;
    if (sscanf(e, "%d,%d", &a1, &a2) == 2 && (a1 == 1 || a1 == 2 || a1 == 4) && (a2 == 1 || a2 == 2)) {
      ct = std::min<int64_t>(a1, tiles >= a1 ? a1 : 1);
      cg = groups == 2 ? a2 : 1;
    }
  }
  if (NT / ct < 8) ct = 1;
  const uint32_t pbox = uint32_t(GEMM_BM / cg), qbox = uint32_t(NT / ct);
  CUtensorMap mph, mpl, mqh, mql, mpx, mqx;
  if (passes == 4) {
    if (Pl == nullptr || Ql == nullptr) return fail(GLP_ERR_ARG, "gemm: passes = 4 needs both halves of the fp16 pairs");
    GLP_TRY(make_map_f16_k128(&mph, Ph, uint64_t(K), uint64_t(M), uint64_t(ldp) * 2, pbox));
    GLP_TRY(make_map_f16_k128(&mqh, Qh, uint64_t(K), uint64_t(N), uint64_t(ldq) * 2, qbox));
    GLP_TRY(make_map_f16_k128(&mpl, Pl, uint64_t(K), uint64_t(M), uint64_t(ldp) * 2, pbox));
    GLP_TRY(make_map_f16_k128(&mql, Ql, uint64_t(K), uint64_t(N), uint64_t(ldq) * 2, qbox));
  } else {
    GLP_TRY(make_map_f32_k128(&mph, Ph, uint64_t(K), uint64_t(M), uint64_t(ldp) * 4, pbox));
    GLP_TRY(make_map_f32_k128(&mqh, Qh, uint64_t(K), uint64_t(N), uint64_t(ldq) * 4, qbox));
  }
  mpx = mph; mqx = mqh;
  // fp16 pairs: the CTA-pair kernel (cta_group::2 + stream-K) unless GLPROX_GEMM_PAIR=0 or no workspace was given
  const char* e_pair = getenv("GLPROX_GEMM_PAIR");   // read per call: tests switch it inside one process
  if (passes == 4 && ws.p_tiled && (ws.pair == nullptr || (e_pair != nullptr && atoi(e_pair) == 0)))
    return fail(GLP_ERR_STATE, "gemm: the resident fp16 operands are tiled for the CTA-pair kernel, which is switched off");
  if (passes == 4 && ws.pair != nullptr && (e_pair == nullptr || atoi(e_pair) != 0)) {
    const int ctas = 2 * (num_sms / 2);
    if (ws.pair_bytes < pair_ws_bytes(ctas, NT)) return fail(GLP_ERR_ARG, "gemm: pair workspace too small");
    if (tiles * groups > PAIR_COUNTER_WORDS) return fail(GLP_ERR_ARG, "gemm: too many output tiles for the pair kernel");
    CUtensorMap p0, p1, q0, q1;
    if (ws.p_tiled) {   // [tiles * nkb * 128 rows][64 fp16]: every 128-row box is one contiguous 16 KB tile
      const uint64_t trows = uint64_t((M + 127) / 128) * uint64_t((K + 63) / 64) * 128;
      GLP_TRY(make_map_f16_k128(&p0, Ph, 64, trows, 128, GEMM_BM));
      GLP_TRY(make_map_f16_k128(&p1, Pl, 64, trows, 128, GEMM_BM));
    } else {
      GLP_TRY(make_map_f16_k128(&p0, Ph, uint64_t(K), uint64_t(M), uint64_t(ldp) * 2, GEMM_BM));
      GLP_TRY(make_map_f16_k128(&p1, Pl, uint64_t(K), uint64_t(M), uint64_t(ldp) * 2, GEMM_BM));
    }
    GLP_TRY(make_map_f16_k128(&q0, Qh, uint64_t(K), uint64_t(N), uint64_t(ldq) * 2, uint32_t(NT / 2)));
    GLP_TRY(make_map_f16_k128(&q1, Ql, uint64_t(K), uint64_t(N), uint64_t(ldq) * 2, uint32_t(NT / 2)));
    PairArgs pa{};
    pa.M = M; pa.K = K; pa.N = N; pa.k_chunk_blocks = k_chunk; pa.D = D; pa.ldd = ldd; pa.fault = fault;
    pa.ctrl = ctrl; pa.D_alt = D_alt; pa.role = role;
    pa.gate = gate.count; pa.gate_cap = gate.cap;
    pa.scale = scale;
    pa.p_tiled = ws.p_tiled ? 1 : 0;
    pa.counters = static_cast<unsigned*>(ws.pair);
    pa.slots = reinterpret_cast<double*>(static_cast<char*>(ws.pair) + size_t(PAIR_COUNTER_WORDS) * 4);
    const cudaError_t ep = NT == 32 ? launch_gemm_pair_t<32>(st, num_sms, p0, p1, q0, q1, pa, ws.pdl)
                         : NT == 64 ? launch_gemm_pair_t<64>(st, num_sms, p0, p1, q0, q1, pa, ws.pdl)
                                    : launch_gemm_pair_t<128>(st, num_sms, p0, p1, q0, q1, pa, ws.pdl);
    if (ep != cudaSuccess) return fail(GLP_ERR_CUDA, "gemm (CTA pairs) launch failed: %s", cudaGetErrorString(ep));
    return GLP_OK;
  }
  if (passes == 4) {
  } else if (passes == 2) {
    if (bf.Phb == nullptr || bf.Plb == nullptr || bf.Qhb == nullptr || bf.Qlb == nullptr)
      return fail(GLP_ERR_ARG, "gemm: passes = 2 needs the bf16 operands");
    GLP_TRY(make_map_bf16_k64(&mpx, bf.Phb, uint64_t(K), uint64_t(M), uint64_t(ldp) * 2, pbox));
    GLP_TRY(make_map_bf16_k64(&mpl, bf.Plb, uint64_t(K), uint64_t(M), uint64_t(ldp) * 2, pbox));
    GLP_TRY(make_map_bf16_k64(&mqx, bf.Qhb, uint64_t(K), uint64_t(N), uint64_t(ldq) * 2, qbox));
    GLP_TRY(make_map_bf16_k64(&mql, bf.Qlb, uint64_t(K), uint64_t(N), uint64_t(ldq) * 2, qbox));
  } else if (passes == 3) {
    if (Pl == nullptr || Ql == nullptr) return fail(GLP_ERR_ARG, "gemm: passes = 3 needs the lo operands");
    GLP_TRY(make_map_f32_k128(&mpl, Pl, uint64_t(K), uint64_t(M), uint64_t(ldp) * 4, pbox));
    GLP_TRY(make_map_f32_k128(&mql, Ql, uint64_t(K), uint64_t(N), uint64_t(ldq) * 4, qbox));
  } else {
    mpl = mph; mql = mqh;
  }
  int S = ws.buf != nullptr ? gemm_splits(M, K, N, k_chunk, num_sms, bk) : 1;
  const int64_t stride = int64_t(N) * ldd;
  while (S > 1 && size_t(S) * stride * 8 > ws.bytes) --S;
  GemmArgs a{};
  a.M = M; a.K = K; a.N = N; a.k_chunk_blocks = k_chunk; a.D = D; a.ldd = ldd; a.fault = fault;
  a.ctrl = ctrl; a.D_alt = D_alt; a.role = role;
  a.splits = S; a.split_stride = stride;
  if (S > 1) { a.D = ws.buf; a.D_alt = ws.buf; }   // partials first; the sum below picks the real destination
  a.ct = ct; a.cg = cg;
  a.gate = gate.count; a.gate_cap = gate.cap;
  a.scale = scale;
  const int64_t items = ((tiles + ct - 1) / ct) * ((groups + cg - 1) / cg) * S;   // cluster items
  // fp16-pair product with one column group (P not shared inside the cluster): separate P / Q operand rings
  // (glp_gemm.cuh, RINGS) when GLPROX_GEMM_RINGS=1.  Off by default: measured at C5 on one B200 in the same run
  // (profiles/r01_ab_gemm_rings_c5.log) 372.8 it/s with the rings (fwd / tpass windows 1.174 / 1.023 ms) against
  // 389.1 it/s with the single ring (1.116 / 0.966 ms), identical objective bits and all GPU tests green with
  // either -- more P bytes in flight do not help, the kernel is not HBM-latency-bound.
  const char* e_rings = getenv("GLPROX_GEMM_RINGS");   // read per call: tests switch it inside one process
  const bool rings = passes == 4 && cg == 1 && e_rings != nullptr && atoi(e_rings) != 0;
  cudaError_t e;
#define GLP_GEMM_LAUNCH(NTV, PV) launch_gemm_t<NTV, PV>(st, num_sms, items, mph, mpl, mqh, mql, mpx, mqx, a)
  if (passes == 1) e = NT == 32 ? GLP_GEMM_LAUNCH(32, 1) : NT == 64 ? GLP_GEMM_LAUNCH(64, 1) : GLP_GEMM_LAUNCH(128, 1);
  else if (passes == 2) e = NT == 32 ? GLP_GEMM_LAUNCH(32, 2) : NT == 64 ? GLP_GEMM_LAUNCH(64, 2) : GLP_GEMM_LAUNCH(128, 2);
  else if (passes == 4 && rings) {
#define GLP_GEMM_LAUNCH_R(NTV) launch_gemm_t<NTV, 4, true>(st, num_sms, items, mph, mpl, mqh, mql, mpx, mqx, a)
    e = NT == 32 ? GLP_GEMM_LAUNCH_R(32) : NT == 64 ? GLP_GEMM_LAUNCH_R(64) : GLP_GEMM_LAUNCH_R(128);
#undef GLP_GEMM_LAUNCH_R
  }
  else if (passes == 4) e = NT == 32 ? GLP_GEMM_LAUNCH(32, 4) : NT == 64 ? GLP_GEMM_LAUNCH(64, 4) : GLP_GEMM_LAUNCH(128, 4);
  else e = NT == 32 ? GLP_GEMM_LAUNCH(32, 3) : NT == 64 ? GLP_GEMM_LAUNCH(64, 3) : GLP_GEMM_LAUNCH(128, 3);
#undef GLP_GEMM_LAUNCH
  if (e != cudaSuccess) return fail(GLP_ERR_CUDA, "gemm launch failed: %s", cudaGetErrorString(e));
  if (S > 1) {
    const int sgrid = int(std::min<int64_t>((M * N + 255) / 256, int64_t(num_sms) * 8));
    glp_sum_splits_kernel<<<sgrid, 256, 0, st>>>(ws.buf, S, stride, M, N, ldd, D, D_alt, ctrl, role, gate.count, gate.cap);
    CU_TRY(cudaGetLastError());
  }
  return GLP_OK;
}

extern "C" int glp_gemm_tf32(int device, void* stream, int passes, int64_t M, int64_t K, int N, const void* P_hi,
                             const void* P_lo, int64_t ldp, const void* Q_hi, const void* Q_lo, int64_t ldq,
                             double* D_dev, int64_t ldd, int k_chunk_blocks) {
  if (P_hi == nullptr || Q_hi == nullptr || D_dev == nullptr || M < 1 || K < 1) return fail(GLP_ERR_ARG, "bad argument");
  int sms = 0;
  GLP_TRY(check_device(device, &sms));
  CU_TRY(cudaSetDevice(device));
  if (N < 1 || N > 256) return fail(GLP_ERR_ARG, "gemm: N must be 1..256 (got %d)", N);
  // k_chunk_blocks < 0: also exercise the split-K path (|k_chunk_blocks| blocks per chunk, partials in a
  // stream-ordered temporary)
  const int kc = k_chunk_blocks < 0 ? -k_chunk_blocks : k_chunk_blocks;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GemmWs ws{nullptr, 0};
  if (k_chunk_blocks < 0) {
    ws.bytes = size_t(gemm_splits(M, K, N, kc, sms, passes == 4 ? 2 * GEMM_BK : GEMM_BK)) * N * ldd * 8;
    CU_TRY(cudaMallocAsync(reinterpret_cast<void**>(&ws.buf), ws.bytes, st));
  }
  GemmBf16 bf{nullptr, nullptr, nullptr, nullptr};
  __nv_bfloat16* tmp = nullptr;
  if (passes == 2) {  // bf16 copies of the caller's hi / lo arrays (stream-ordered temporaries; test / probe entry point)
    if (P_lo == nullptr || Q_lo == nullptr) return fail(GLP_ERR_ARG, "gemm: passes = 2 needs the lo operands");
    const size_t pe = size_t(M) * ldp, qe = size_t(N) * ldq;
    CU_TRY(cudaMallocAsync(reinterpret_cast<void**>(&tmp), (2 * pe + 2 * qe) * 2, st));
    __nv_bfloat16 *phb = tmp, *plb = tmp + pe, *qhb = tmp + 2 * pe, *qlb = tmp + 2 * pe + qe;
    const int g1 = int(std::min<size_t>((pe + 255) / 256, 148 * 16)), g2 = int(std::min<size_t>((qe + 255) / 256, 148 * 16));
    glp_to_bf16_kernel<<<g1, 256, 0, st>>>(static_cast<const float*>(P_hi), phb, int64_t(pe));
    glp_to_bf16_kernel<<<g1, 256, 0, st>>>(static_cast<const float*>(P_lo), plb, int64_t(pe));
    glp_to_bf16_kernel<<<g2, 256, 0, st>>>(static_cast<const float*>(Q_hi), qhb, int64_t(qe));
    glp_to_bf16_kernel<<<g2, 256, 0, st>>>(static_cast<const float*>(Q_lo), qlb, int64_t(qe));
    bf = GemmBf16{phb, plb, qhb, qlb};
  }
  int rc;
  if (passes == 4) {
    // fp16 pairs of P_hi and Q_hi themselves (the lo arguments are not used): scales from the operands' maxima,
    // all on the device in stream-ordered temporaries (test / probe entry point)
    const char* e_pair = getenv("GLPROX_GEMM_PAIR");
    const bool tiled = e_pair == nullptr || atoi(e_pair) != 0;
    const int64_t nkb_t = (K + 63) / 64;
    const size_t pe = tiled ? size_t((M + 127) / 128) * 128 * size_t(nkb_t) * 64 : size_t(M) * ldp, qe = size_t(N) * ldq;
    uint8_t* t4 = nullptr;
    CU_TRY(cudaMallocAsync(reinterpret_cast<void**>(&t4), (2 * pe + 2 * qe) * 2 + 64, st));
    if (tiled) CU_TRY(cudaMemsetAsync(t4, 0, (2 * pe + 2 * qe) * 2 + 64, st));
    __half *p0 = reinterpret_cast<__half*>(t4 + 64), *p1 = p0 + pe, *q0 = p1 + pe, *q1 = q0 + qe;
    unsigned* amax = reinterpret_cast<unsigned*>(t4);
    double* inv = reinterpret_cast<double*>(t4 + 16);
    CU_TRY(cudaMemsetAsync(t4, 0, 64, st));
    const float *Pf = static_cast<const float*>(P_hi), *Qf = static_cast<const float*>(Q_hi);
    const int g1 = int(std::min<size_t>((pe + 255) / 256, 148 * 16)), g2 = int(std::min<size_t>((qe + 255) / 256, 148 * 16));
    glp_absmax_kernel<<<g1, 256, 0, st>>>(Pf, M, K, ldp, amax);
    glp_absmax_kernel<<<g2, 256, 0, st>>>(Qf, N, K, ldq, amax + 1);
    if (tiled) glp_split_f16_tiled_dev_scale_kernel<<<g1, 256, 0, st>>>(Pf, p0, p1, M, K, ldp, nkb_t, amax);
    else glp_split_f16_kernel<<<g1, 256, 0, st>>>(Pf, p0, p1, int64_t(pe), amax, nullptr, 0, 1.0f, nullptr, nullptr, nullptr, 0);
    glp_split_f16_kernel<<<g2, 256, 0, st>>>(Qf, q0, q1, int64_t(qe), amax + 1, nullptr, 0, 1.0f, amax, inv, nullptr, 0);
    CU_TRY(cudaGetLastError());
    ws.p_tiled = tiled;
    ws.pair_bytes = pair_ws_bytes(sms, 128);
    CU_TRY(cudaMallocAsync(&ws.pair, ws.pair_bytes, st));
    CU_TRY(cudaMemsetAsync(ws.pair, 0, size_t(PAIR_COUNTER_WORDS) * 4, st));
    rc = gemm_enqueue(st, sms, 4, M, K, N, reinterpret_cast<const float*>(p0), reinterpret_cast<const float*>(p1), ldp,
                      reinterpret_cast<const float*>(q0), reinterpret_cast<const float*>(q1), ldq, D_dev, ldd, kc, nullptr,
                      nullptr, nullptr, 0, ws, bf, GemmGate{nullptr, 0}, inv);
    cudaFreeAsync(ws.pair, st);
    cudaFreeAsync(t4, st);
  } else {
    rc = gemm_enqueue(st, sms, passes, M, K, N, static_cast<const float*>(P_hi), static_cast<const float*>(P_lo), ldp,
                      static_cast<const float*>(Q_hi), static_cast<const float*>(Q_lo), ldq, D_dev, ldd, kc, nullptr,
                      nullptr, nullptr, 0, ws, bf);
  }
  if (ws.buf != nullptr) cudaFreeAsync(ws.buf, st);
  if (tmp != nullptr) cudaFreeAsync(tmp, st);
  return rc;
}

extern "C" int glp_tf32_split(void* stream, const void* src, int64_t rows, int64_t cols, int64_t ld_src, void* hi,
                              void* lo, int64_t ld_dst, int transpose) {
  if (src == nullptr || hi == nullptr || lo == nullptr || rows < 1 || cols < 1) return fail(GLP_ERR_ARG, "bad argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (transpose) {
    dim3 grid(unsigned((cols + 31) / 32), unsigned((rows + 31) / 32));
    if (grid.y > 65535u) return fail(GLP_ERR_ARG, "too many rows for the transposing split");
    glp_transpose_split_kernel<<<grid, dim3(32, 8), 0, st>>>(static_cast<const float*>(src), static_cast<float*>(hi),
                                                             static_cast<float*>(lo), rows, cols, ld_src, ld_dst);
  } else {
    const int grid = int(std::min<int64_t>((rows * cols + 255) / 256, 148 * 16));
    glp_split_tf32_kernel<<<grid, 256, 0, st>>>(static_cast<const float*>(src), static_cast<float*>(hi),
                                                static_cast<float*>(lo), rows, cols, ld_src, ld_dst, nullptr, 0);
  }
  CU_TRY(cudaGetLastError());
  return GLP_OK;
}
