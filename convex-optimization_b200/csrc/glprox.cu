// glprox.cu -- libglprox.so: handle management, the streaming schedule and the C ABI (include/glprox.h).
// Host side of the B200 proximal-gradient path; all numerics run in the sm_100a kernels of
// glp_fwd.cuh / glp_tpass.cuh / glp_prox.cuh / glp_decide.cuh.  There is no CPU fallback: without an
// sm_100 device every entry point that computes returns GLP_ERR_NO_DEVICE.
#include <cuda.h>
#include <cuda_runtime.h>
#include <mutex>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <limits>
#include <thread>
#include <vector>

#include "../../include/glprox.h"
#include "glp_admm.cuh"
#include "glp_admm_primal.cuh"
#include "glp_alm.cuh"
#include "glp_comm.cuh"
#include "glp_decide.cuh"
#include "glp_dense.cuh"
#include "glp_fwd.cuh"
#include "glp_gather_decide.cuh"
#include "glp_gemm.cuh"
#include "glp_gemm_pair.cuh"
#include "glp_metrics.cuh"
#include "glp_misc.cuh"
#include "glp_persist.cuh"
#include "glp_thrfix.cuh"
#include "glp_tpass.cuh"
#include "glp_types.cuh"
#include "glp_wide_comm.cuh"

using namespace glp;

static_assert(sizeof(glp_opts) == 64 && sizeof(glp_result) == 80 && sizeof(glp_step_params) == 40,
              "ABI structs changed: update glprox/_lib.py and tests/test_abi.py");

// ------------------------------------------------------------------------------------------------
// error plumbing
// ------------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CU_TRY(expr)                                                                              \
  do {                                                                                            \
    cudaError_t e_ = (expr);                                                                      \
    if (e_ != cudaSuccess)                                                                        \
      return fail(GLP_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, \
                  __LINE__);                                                                      \
  } while (0)

#define GLP_TRY(expr)        \
  do {                       \
    int rc_ = (expr);        \
    if (rc_ != GLP_OK) return rc_; \
  } while (0)

extern "C" int glp_version(void) { return GLP_VERSION; }
extern "C" const char* glp_last_error(void) { return g_err; }

extern "C" int glp_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  int ok = 0;
  for (int d = 0; d < n; ++d) {
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10) ++ok;
  }
  return ok;
}

extern "C" int glp_default_opts(int algo, glp_opts* o) {
  if (o == nullptr || algo < GLP_PROXGD || algo > GLP_SGD) return fail(GLP_ERR_ARG, "bad algo/opts");
  memset(o, 0, sizeof(*o));
  // gl_ProxGD_primal.py:10-19 / gl_FProxGD_primal.py:10-19 / gl_FGD_primal.py:10-21 / gl_GD_primal.py:10-19 /
  // gl_SGD_primal.py:10-18
  const bool plain = algo == GLP_GD || algo == GLP_SGD;
  o->maxit = algo == GLP_PROXGD ? 2500 : (algo == GLP_GD ? 2500 : (algo == GLP_SGD ? 2100 : 1500));
  o->thres = 1e-3;
  o->step_type = plain ? GLP_STEP_DIMINISHING : GLP_STEP_LINE_SEARCH;
  o->alpha0 = algo == GLP_PROXGD ? 2e-3 : 1e-3;
  o->ftol = plain ? 1e-5 : 1e-6;
  o->stable_len_threshold = plain ? 100 : 70;
  o->line_search_attenuation_coeffi = algo == GLP_PROXGD ? 0.9 : (plain ? 1.0 : 0.98);
  o->maxit_line_search_iter = plain ? 0 : 5;
  o->delta = algo == GLP_FGD ? 1e-6 : (algo == GLP_GD ? 1e-3 : 0.0);
  o->backend = GLP_BACKEND_AUTO;
  o->poll_interval = 0;
  return GLP_OK;
}

// ------------------------------------------------------------------------------------------------
// handle
// ------------------------------------------------------------------------------------------------
struct glp_handle {
  int device = 0;
  cudaStream_t stream = nullptr;
  int dtype = GLP_F32;
  Geom g{};
  const void* A = nullptr;
  int64_t lda = 0;
  CUtensorMap tmap{};
  int evict_first = 1;
  int num_sms = 0;
  void* nccl_comm = nullptr;
  bool xch_nccl = false;        // baseline exchange: ncclAllReduce instead of the peer-mailbox push (A/B measurements)
  double* nccl_stage = nullptr; // [2l*ldm + 8] partial product + scalar tail, [8] max

  // device arena
  void* arena = nullptr;
  double *b = nullptr, *Av = nullptr, *grad = nullptr;
  double* AX[2] = {nullptr, nullptr};
  double* state[2] = {nullptr, nullptr};
  double *scal_partial = nullptr, *cnt_partial = nullptr, *rr_partial = nullptr;
  double* gy_partial[2] = {nullptr, nullptr};
  void* r = nullptr;    // [l][ldm] dtype T
  void* xop = nullptr;  // fp32: [2l][ldn] float
  double *f_partial = nullptr, *t_partial = nullptr;
  unsigned *f_counter = nullptr, *t_counter = nullptr;
  Ctrl* ctrl = nullptr;  // [2]
  int32_t* fault = nullptr;
  double *f_hist = nullptr, *f_hist_best = nullptr, *s_hist = nullptr;   // one allocation: [3][hist_cap]
  int hist_cap = 0;
  int last_k = 0;                  // recorded iterations of the last finished solve (glp_solve_sparsity_hist)
  int decide_grid = 0;
  int n_fchunks = 0;

  // multi-GPU: peer mailboxes (glp_comm.cuh)
  Comm comm{};                     // world == 0: single GPU
  void* mailbox = nullptr;         // this rank's mailbox (cudaMalloc, IPC-exportable)
  size_t mailbox_bytes = 0;
  void* ipc_mapped[MAX_WORLD] = {nullptr};
  uint64_t epoch = 0;              // solves started on this handle (identical on all ranks: SPMD)
  uint64_t idx_f = 0, idx_t = 0;   // forward / prox-producing launches of the current solve
  uint64_t seq_t_cur = 0;          // flag value of the latest prox-producing launch
  int group = 0;                   // forward group of the current solve (mailbox parity)
  unsigned *fin_f = nullptr, *fin_t = nullptr;
  double *rr_g = nullptr, *scal_red = nullptr, *cnt_red = nullptr;
  int gather_grid = 0;
  // launch-latency switches (each read from the environment at glp_create so that A/B runs need no rebuild):
  bool pdl = true;                 // GLPROX_PDL=0: plain stream serialization instead of programmatic dependent launch
  // GLPROX_FUSED_GATHER=1: gather + decide as ONE launch with a grid barrier (glp_gather_decide.cuh).  Measured on
  // 8 B200 at C4 (profiles/r02_ab_latency_switches_c4_n8.log): 1397.9 it/s fused vs 1405.1 with the two launches --
  // under programmatic dependent launch the second launch costs no more than the grid barrier.  Off by default.
  bool fused_gather = false;
  bool push_fence_block = false;   // GLPROX_PUSH_FENCE=block: system fence per pushed row block (round-1 behaviour)
  int fused_grid = 0;              // CTAs of glp_gather_decide_kernel (all co-resident: grid-wide barrier inside)
  unsigned* gd_count = nullptr;
  unsigned long long* gd_flag = nullptr;
  unsigned long long gd_gen = 0;
  int prof_stride = 1;             // profile every prof_stride-th slot (events between launches defeat the overlap)
  bool prof_slot = true;
  const Ctrl* cur_ctrl = nullptr;  // control block the launches of the current slot read
  int backend_used = GLP_BACKEND_STREAM;
  // wide multi-RHS mode (K5, glp_gemm.cuh): resident tensor-core operand sets of A and A^T
  int tensor_mode = 0;             // glp_set_tensor_mode: 0 auto (fp16 pairs), 1 off, 2 TF32, 3 TF32 + bf16, 4 fp16 pairs, 5 3xTF32
  float *A_hi = nullptr, *A_lo = nullptr, *At_hi = nullptr, *At_lo = nullptr;
  float *xop_lo = nullptr, *r_lo = nullptr;
  double* cnt_g = nullptr;
  // tensor mode 3 (tf32 main term + bf16 corrections): bf16 copies of hi and lo instead of the fp32 lo arrays
  __nv_bfloat16 *A_hb = nullptr, *A_lb = nullptr, *At_hb = nullptr, *At_lb = nullptr;
  __nv_bfloat16 *xop_hb = nullptr, *xop_lb = nullptr, *r_hb = nullptr, *r_lb = nullptr;
  // tensor mode 4 (fp16 pairs): A s and A^T s as v0 + v1 (2 x |A| of HBM in all), the small operands likewise
  __half *A_f0 = nullptr, *A_f1 = nullptr, *At_f0 = nullptr, *At_f1 = nullptr;
  __half *xop_f0 = nullptr, *xop_f1 = nullptr, *r_f0 = nullptr, *r_f1 = nullptr;
  unsigned* f16_amax = nullptr;    // [2] float bits of max|xn| / max|r| of the current operands
  double* f16_inv = nullptr;       // [2] 1 / (s_A s_x), 1 / (s_A s_r): what the product epilogues multiply by
  float f16_sA = 1.0f;             // power-of-two scale of the resident A
  bool f16_unusable = false;       // auto mode found non-finite entries or badly scaled columns / rows in A: 3xTF32 instead
  bool f16_tiled = false;          // A_f0.. / At_f0.. are stored as 128 x 64 tiles for the CTA-pair kernel (glp_gemm.cuh)
  // A.xnt from A.xn by a sparse correction (glp_thrfix.cuh); thr_cap < 0: off (GLPROX_WIDE_THRFIX=0)
  int32_t *thr_rows = nullptr, *thr_col_count = nullptr, *thr_count = nullptr;
  float* thr_vals = nullptr;
  int32_t thr_cap = -1;            // list entries per operand column
  double* admm_L = nullptr;    // gl_Admm_dual: m x m Cholesky factor of I + rho A A^T
  double* admm_Y = nullptr;    // blocked path (m > 2048): Y = L^-T, streamed by the inner handle (glp_dense.cuh)
  int64_t admm_ldy = 0;
  glp_handle* admm_inner = nullptr;   // m x m fp64 problem over Y: K1 / K2 apply z = Y (Y^T w)
  double* admm_zero_b = nullptr;
  double* admm_gpart = nullptr;       // [ncb][ADMM_GSTRIDE] per-column-block records of the update kernel
  // gl_Admm_primal / gl_Alm_dual (glp_admm2_host.inl): lazily sized fp64 workspaces in n-space, m-space, per column block
  double* aux_n = nullptr;  size_t aux_n_len = 0;
  double* aux_m = nullptr;  size_t aux_m_len = 0;
  double* aux_rec = nullptr; size_t aux_rec_len = 0;
  void* pair_ws = nullptr;     // K5 on CTA pairs (glp_gemm_pair.cuh): arrival counters + partial-tile slots
  size_t pair_ws_bytes = 0;
  double* gemm_ws = nullptr;   // split-K partial products of K5
  size_t gemm_ws_bytes = 0;
  int64_t n_gemm = 0;              // tensor-core launches of the current solve

  // step-level scratch (lazily sized)
  void* scratch = nullptr;
  size_t scratch_bytes = 0;

  // pinned host mirror + solve bookkeeping
  Ctrl* h_ctrl = nullptr;  // [2] snapshots
  cudaEvent_t ev_snap[2] = {nullptr, nullptr};
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  SolveConst sc{};
  bool solving = false;
  int64_t decide_launches = 0;
  int64_t n_launches = 0;

  // CUDA-event profile of the streaming launches
  bool profiling = false;
  std::vector<cudaEvent_t> ev_pool;
  std::vector<int> ev_tag;  // per pair: 0 forward, 1 transposed
  size_t ev_used = 0;       // pairs used in the current solve
  glp_profile prof{};
};

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency);
// looked up once per process (K5 encodes four maps per product launch).
static PFN_encodeTiled encode_tiled_fn() {
  static PFN_encodeTiled cached = []() -> PFN_encodeTiled {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<PFN_encodeTiled>(fn);
  }();
  return cached;
}

static int make_tensor_map(glp_handle* h) {
  PFN_encodeTiled fn = encode_tiled_fn();
  if (fn == nullptr) return fail(GLP_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  const int esz = h->dtype == GLP_F32 ? 4 : 8;
  const cuuint64_t dims[2] = {cuuint64_t(h->g.n), cuuint64_t(h->g.m)};
  const cuuint64_t strides[1] = {cuuint64_t(h->lda) * esz};
  const cuuint32_t box[2] = {cuuint32_t(TILE_ROW_BYTES / esz), cuuint32_t(TILE_ROWS)};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(
      &h->tmap, h->dtype == GLP_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2,
      const_cast<void*>(h->A), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(GLP_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", int(r));
  return GLP_OK;
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ------------------------------------------------------------------------------------------------
// CUDA-event profile of the streaming launches (roofline evidence measured on the launching stream)
// ------------------------------------------------------------------------------------------------
static void prof_mark(glp_handle* h, int tag, int end) {
  if (!h->profiling || !h->prof_slot) return;
  const size_t idx = 2 * h->ev_used + size_t(end);
  while (h->ev_pool.size() <= idx) {
    cudaEvent_t e = nullptr;
    if (cudaEventCreate(&e) != cudaSuccess) { h->profiling = false; return; }
    h->ev_pool.push_back(e);
  }
  cudaEventRecord(h->ev_pool[idx], h->stream);
  if (!end) {
    if (h->ev_tag.size() <= h->ev_used) h->ev_tag.resize(h->ev_used + 1);
    h->ev_tag[h->ev_used] = tag;
  } else {
    h->ev_used++;
  }
}

static void prof_collect(glp_handle* h) {
  if (!h->profiling || h->ev_used == 0) return;
  std::vector<float> ms(h->ev_used, 0.f);
  float mx[2] = {0.f, 0.f};
  for (size_t i = 0; i < h->ev_used; ++i) {
    if (cudaEventElapsedTime(&ms[i], h->ev_pool[2 * i], h->ev_pool[2 * i + 1]) != cudaSuccess) ms[i] = 0.f;
    mx[h->ev_tag[i]] = std::max(mx[h->ev_tag[i]], ms[i]);
  }
  for (size_t i = 0; i < h->ev_used; ++i) {
    const int tag = h->ev_tag[i];
    if (ms[i] < 0.5f * mx[tag]) continue;  // no-op after `done`, or epilogue-only retry
    if (tag == 0) {
      h->prof.fwd_ms_min = h->prof.fwd_launches == 0 ? ms[i] : std::min<double>(h->prof.fwd_ms_min, ms[i]);
      h->prof.fwd_launches++; h->prof.fwd_ms_sum += ms[i];
    } else {
      h->prof.tpass_ms_min = h->prof.tpass_launches == 0 ? ms[i] : std::min<double>(h->prof.tpass_ms_min, ms[i]);
      h->prof.tpass_launches++; h->prof.tpass_ms_sum += ms[i];
    }
  }
  h->ev_used = 0;
}

// ------------------------------------------------------------------------------------------------
// kernel launch helpers (dtype / chunk-width dispatch)
// ------------------------------------------------------------------------------------------------
// Every kernel of the streaming schedule is launched with the programmatic-stream-serialization attribute: its CTAs
// may become resident while the predecessor drains, run their prologue (mbarrier init, tensor-map prefetch) and block
// in griddepcontrol.wait until the predecessor's memory is visible (glp_device.cuh: pdl_wait).
template <typename... KArgs, typename... Args>
static cudaError_t launch_k(const glp_handle* h, void (*kfn)(KArgs...), dim3 grid, dim3 block, size_t smem, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = h->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = h->pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kfn, std::forward<Args>(args)...);
}

template <typename T, int RC>
static cudaError_t launch_fwd_t(glp_handle* h, const FwdArgs& a) {
  static bool attr_set[64] = {false};
  auto kfn = glp_fwd_kernel<T, RC>;
  if (!attr_set[h->device & 63]) {
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, fwd_smem_bytes<RC>());
    if (e != cudaSuccess) return e;
    attr_set[h->device & 63] = true;
  }
  prof_mark(h, 0, 0);
  const cudaError_t e = launch_k(h, kfn, dim3(h->g.grid), dim3(FWD_THREADS), fwd_smem_bytes<RC>(), h->tmap, a, h->g, h->evict_first);
  prof_mark(h, 0, 1);
  h->n_launches++;
  return e;
}

static cudaError_t launch_fwd(glp_handle* h, int rc, const FwdArgs& a) {
  if (h->dtype == GLP_F32) {
    if (rc == 4) return launch_fwd_t<float, 4>(h, a);
    if (rc == 2) return launch_fwd_t<float, 2>(h, a);
    return launch_fwd_t<float, 1>(h, a);
  }
  if (rc == 4) return launch_fwd_t<double, 4>(h, a);
  if (rc == 2) return launch_fwd_t<double, 2>(h, a);
  return launch_fwd_t<double, 1>(h, a);
}

template <typename T, int LC>
static cudaError_t launch_tpass_t(glp_handle* h, const TpassArgs& a, const ProxArgs& p) {
  static bool attr_set[64] = {false};
  auto kfn = glp_tpass_kernel<T, LC>;
  if (!attr_set[h->device & 63]) {
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, tp_smem_bytes<T, LC>());
    if (e != cudaSuccess) return e;
    attr_set[h->device & 63] = true;
  }
  prof_mark(h, 1, 0);
  const cudaError_t e = launch_k(h, kfn, dim3(h->g.grid), dim3(TP_THREADS), tp_smem_bytes<T, LC>(), h->tmap, a, p, h->g, h->evict_first);
  prof_mark(h, 1, 1);
  h->n_launches++;
  return e;
}

static cudaError_t launch_tpass(glp_handle* h, int lc, const TpassArgs& a, const ProxArgs& p) {
  if (h->dtype == GLP_F32) return lc == 2 ? launch_tpass_t<float, 2>(h, a, p) : launch_tpass_t<float, 1>(h, a, p);
  return lc == 2 ? launch_tpass_t<double, 2>(h, a, p) : launch_tpass_t<double, 1>(h, a, p);
}

static cudaError_t launch_prox(glp_handle* h, const ProxArgs& p, int only_if_retry) {
  const int grid = std::min<int>(h->g.ncb, 4 * h->num_sms);
  h->n_launches++;
  if (h->dtype == GLP_F32) return launch_k(h, glp_prox_kernel<float>, dim3(grid), dim3(CONSUMER_THREADS), 0, p, h->g, only_if_retry);
  return launch_k(h, glp_prox_kernel<double>, dim3(grid), dim3(CONSUMER_THREADS), 0, p, h->g, only_if_retry);
}

static cudaError_t launch_decide(glp_handle* h, const DecideArgs& a) {
  h->n_launches++;
  if (h->dtype == GLP_F32) return launch_k(h, glp_decide_kernel<float>, dim3(h->decide_grid), dim3(DECIDE_THREADS), 0, a, h->g);
  return launch_k(h, glp_decide_kernel<double>, dim3(h->decide_grid), dim3(DECIDE_THREADS), 0, a, h->g);
}

// multi-GPU, l = 2 streaming path: gather of the forward group just launched + decide, one launch
static cudaError_t launch_gather_decide(glp_handle* h, const GatherArgs& ga, const DecideArgs& da) {
  h->n_launches++;
  const GridBarrier bar{h->gd_count, h->gd_flag, ++h->gd_gen};
  if (h->dtype == GLP_F32)
    return launch_k(h, glp_gather_decide_kernel<float>, dim3(h->fused_grid), dim3(DECIDE_THREADS), 0, ga, da, h->g, bar);
  return launch_k(h, glp_gather_decide_kernel<double>, dim3(h->fused_grid), dim3(DECIDE_THREADS), 0, ga, da, h->g, bar);
}

// Operand columns of the forward stream are processed in chunks of 4, 2 or 1.
static int next_chunk(int remaining) { return remaining >= 4 ? 4 : (remaining >= 2 ? 2 : 1); }
static int count_fchunks(int ncols) {
  int n = 0;
  for (int q = 0; q < ncols; q += next_chunk(ncols - q)) ++n;
  return n;
}

template <typename S, typename D>
static cudaError_t to_planar(cudaStream_t st, const S* src, D* dst, int64_t rows, int cols, int64_t ld) {
  const int64_t total = rows * cols;
  if (total == 0) return cudaSuccess;
  const int grid = int(std::min<int64_t>((total + 255) / 256, 2048));
  to_planar_kernel<S, D><<<grid, 256, 0, st>>>(src, dst, rows, cols, ld);
  return cudaGetLastError();
}
template <typename S, typename D>
static cudaError_t from_planar(cudaStream_t st, const S* src, D* dst, int64_t rows, int cols, int64_t ld) {
  const int64_t total = rows * cols;
  if (total == 0) return cudaSuccess;
  const int grid = int(std::min<int64_t>((total + 255) / 256, 2048));
  from_planar_kernel<S, D><<<grid, 256, 0, st>>>(src, dst, rows, cols, ld);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// create / destroy
// ------------------------------------------------------------------------------------------------
static int check_device(int device, int* num_sms) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return fail(GLP_ERR_NO_DEVICE, "no CUDA device visible (%s); libglprox has no CPU fallback",
                e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
  }
  if (device < 0 || device >= n) return fail(GLP_ERR_ARG, "device %d out of range (have %d)", device, n);
  int major = 0;
  CU_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  if (major != 10)
    return fail(GLP_ERR_NO_DEVICE, "device %d has compute capability %d.x; this library is sm_100a only", device, major);
  CU_TRY(cudaDeviceGetAttribute(num_sms, cudaDevAttrMultiProcessorCount, device));
  return GLP_OK;
}

extern "C" int glp_destroy(glp_handle* h) {
  if (h == nullptr) return GLP_OK;
  cudaSetDevice(h->device);
  if (h->stream != nullptr || true) cudaStreamSynchronize(h->stream);
  for (int r = 0; r < MAX_WORLD; ++r)
    if (h->ipc_mapped[r]) cudaIpcCloseMemHandle(h->ipc_mapped[r]);
  if (h->mailbox) cudaFree(h->mailbox);
  if (h->A_hi) cudaFree(h->A_hi);
  if (h->A_lo) cudaFree(h->A_lo);
  if (h->At_hi) cudaFree(h->At_hi);
  if (h->At_lo) cudaFree(h->At_lo);
  if (h->xop_lo) cudaFree(h->xop_lo);
  if (h->r_lo) cudaFree(h->r_lo);
  if (h->cnt_g) cudaFree(h->cnt_g);
  if (h->thr_rows) cudaFree(h->thr_rows);
  if (h->thr_vals) cudaFree(h->thr_vals);
  if (h->thr_col_count) cudaFree(h->thr_col_count);
  if (h->thr_count) cudaFree(h->thr_count);
  if (h->gemm_ws) cudaFree(h->gemm_ws);
  if (h->pair_ws) cudaFree(h->pair_ws);
  if (h->admm_L) cudaFree(h->admm_L);
  if (h->admm_inner) { glp_handle* in = h->admm_inner; h->admm_inner = nullptr; glp_destroy(in); }
  if (h->admm_Y) cudaFree(h->admm_Y);
  if (h->admm_zero_b) cudaFree(h->admm_zero_b);
  if (h->admm_gpart) cudaFree(h->admm_gpart);
  if (h->aux_n) cudaFree(h->aux_n);
  if (h->aux_m) cudaFree(h->aux_m);
  if (h->aux_rec) cudaFree(h->aux_rec);
  if (h->nccl_stage) cudaFree(h->nccl_stage);
  for (__half* q : {h->A_f0, h->A_f1, h->At_f0, h->At_f1, h->xop_f0, h->xop_f1, h->r_f0, h->r_f1})
    if (q) cudaFree(q);
  if (h->f16_amax) cudaFree(h->f16_amax);
  if (h->f16_inv) cudaFree(h->f16_inv);
  for (__nv_bfloat16* q : {h->A_hb, h->A_lb, h->At_hb, h->At_lb, h->xop_hb, h->xop_lb, h->r_hb, h->r_lb})
    if (q) cudaFree(q);
  if (h->arena) cudaFree(h->arena);
  if (h->scratch) cudaFree(h->scratch);
  if (h->f_hist) cudaFree(h->f_hist);
  if (h->h_ctrl) cudaFreeHost(h->h_ctrl);
  for (int i = 0; i < 2; ++i)
    if (h->ev_snap[i]) cudaEventDestroy(h->ev_snap[i]);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
  cudaGetLastError();
  delete h;
  return GLP_OK;
}

extern "C" int glp_create(glp_handle** out, int device, int m, int n_local, int l, int dtype, const void* A_dev,
                          int64_t lda, const void* b_dev, void* nccl_comm, void* stream) {
  if (out == nullptr) return fail(GLP_ERR_ARG, "out is null");
  *out = nullptr;
  if (dtype != GLP_F32 && dtype != GLP_F64) return fail(GLP_ERR_ARG, "dtype must be GLP_F32 or GLP_F64");
  if (m <= 0 || n_local <= 0 || l <= 0) return fail(GLP_ERR_ARG, "m, n, l must be positive (got %d %d %d)", m, n_local, l);
  if (A_dev == nullptr || b_dev == nullptr) return fail(GLP_ERR_ARG, "A_dev / b_dev is null");
  const int esz = dtype == GLP_F32 ? 4 : 8;
  if (lda < n_local) return fail(GLP_ERR_ARG, "lda (%lld) < n (%d)", (long long)lda, n_local);
  if ((lda * esz) % 16 != 0 || (reinterpret_cast<uintptr_t>(A_dev) % 16) != 0)
    return fail(GLP_ERR_ARG, "A_dev and lda*sizeof(T) must be 16-byte aligned for TMA (lda=%lld)", (long long)lda);
  int num_sms = 0;
  GLP_TRY(check_device(device, &num_sms));
  CU_TRY(cudaSetDevice(device));

  glp_handle* h = new glp_handle();
  h->device = device;
  h->stream = static_cast<cudaStream_t>(stream);
  h->dtype = dtype;
  h->A = A_dev;
  h->lda = lda;
  h->num_sms = num_sms;
  h->nccl_comm = nccl_comm;
  const int tc = TILE_ROW_BYTES / esz;
  Geom& g = h->g;
  g.m = m; g.n = n_local; g.l = l;
  g.nrb = (m + TILE_ROWS - 1) / TILE_ROWS;
  g.ncb = (n_local + tc - 1) / tc;
  g.ldm = int64_t(g.nrb) * TILE_ROWS;
  g.ldn = int64_t(g.ncb) * tc;
  const int64_t ntiles = int64_t(g.nrb) * g.ncb;
  g.grid = int(std::min<int64_t>(num_sms, ntiles));
  g.world = 1;
  g.n_total = n_local;
  h->evict_first = (double(m) * double(lda) * esz > 64.0 * 1024 * 1024) ? 1 : 0;
  h->n_fchunks = count_fchunks(2 * l);
  const int64_t vec_len = int64_t(l) * std::max(g.ldm, g.ldn);
  // latency-bound element loops: up to 4 CTAs per SM once the vectors are long (wide multi-RHS mode)
  h->decide_grid = int(std::max<int64_t>(1, std::min<int64_t>(int64_t(num_sms) * 4, (vec_len + 4 * DECIDE_THREADS - 1) / (4 * DECIDE_THREADS))));

  int rc = make_tensor_map(h);
  if (rc != GLP_OK) { glp_destroy(h); return rc; }

  // ---- carve the arena
  size_t off = 0;
  auto carve = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  const size_t lm = size_t(l) * g.ldm, ln = size_t(l) * g.ldn;
  const size_t o_b = carve(lm * 8), o_ax0 = carve(2 * lm * 8), o_ax1 = carve(2 * lm * 8), o_av = carve(lm * 8);
  const size_t o_r = carve(lm * esz);
  const size_t o_s0 = carve(3 * ln * 8), o_s1 = carve(3 * ln * 8), o_grad = carve(ln * 8);
  const size_t o_xop = carve(dtype == GLP_F32 ? 2 * ln * 4 : 0);
  const size_t o_scal = carve(size_t(g.ncb) * SCAL_STRIDE * 8), o_cnt = carve(size_t(g.grid) * 8);
  const size_t o_rr = carve(size_t(h->n_fchunks) * g.nrb * 8);
  h->gather_grid = int(std::max<int64_t>(1, std::min<int64_t>(int64_t(num_sms) * 8, (int64_t(2) * lm + GATHER_THREADS - 1) / GATHER_THREADS)));
  // the fused gather + decide kernel synchronises its whole grid: clamp it to what is resident at once
  {
    int per_sm = 0;
    cudaError_t eo = dtype == GLP_F32
        ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, glp_gather_decide_kernel<float>, DECIDE_THREADS, 0)
        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, glp_gather_decide_kernel<double>, DECIDE_THREADS, 0);
    if (eo != cudaSuccess || per_sm < 1) { cudaGetLastError(); per_sm = 1; }
    h->fused_grid = int(std::min<int64_t>(std::max(h->gather_grid, h->decide_grid), int64_t(per_sm) * num_sms));
  }
  const int part_grid = std::max(std::max(h->decide_grid, h->gather_grid), h->fused_grid);
  const size_t o_gy0 = carve(size_t(part_grid) * 8), o_gy1 = carve(size_t(part_grid) * 8);
  const size_t o_fp = carve(size_t(g.grid + g.nrb) * TILE_ROWS * 4 * 8);
  const size_t o_tp = carve(size_t(g.grid + g.ncb) * 2 * tc * 8);
  const size_t o_fc = carve(size_t(g.nrb) * 4), o_tc = carve(size_t(g.ncb) * 4);
  const size_t o_ctrl = carve(2 * sizeof(Ctrl)), o_fault = carve(16);
  const size_t o_fin = carve(16), o_rrg = carve(size_t(part_grid) * 8), o_sred = carve(SCAL_STRIDE * 8), o_cred = carve(8);
  const size_t o_gd = carve(32);
  cudaError_t e = cudaMalloc(&h->arena, off);
  if (e != cudaSuccess) { glp_destroy(h); return fail(GLP_ERR_CUDA, "cudaMalloc(%zu) failed: %s", off, cudaGetErrorString(e)); }
  e = cudaMemsetAsync(h->arena, 0, off, h->stream);
  if (e != cudaSuccess) { glp_destroy(h); return fail(GLP_ERR_CUDA, "cudaMemsetAsync failed: %s", cudaGetErrorString(e)); }
  char* base = static_cast<char*>(h->arena);
  h->b = reinterpret_cast<double*>(base + o_b);
  h->AX[0] = reinterpret_cast<double*>(base + o_ax0);
  h->AX[1] = reinterpret_cast<double*>(base + o_ax1);
  h->Av = reinterpret_cast<double*>(base + o_av);
  h->r = base + o_r;
  h->state[0] = reinterpret_cast<double*>(base + o_s0);
  h->state[1] = reinterpret_cast<double*>(base + o_s1);
  h->grad = reinterpret_cast<double*>(base + o_grad);
  h->xop = dtype == GLP_F32 ? static_cast<void*>(base + o_xop) : nullptr;
  h->scal_partial = reinterpret_cast<double*>(base + o_scal);
  h->cnt_partial = reinterpret_cast<double*>(base + o_cnt);
  h->rr_partial = reinterpret_cast<double*>(base + o_rr);
  h->gy_partial[0] = reinterpret_cast<double*>(base + o_gy0);
  h->gy_partial[1] = reinterpret_cast<double*>(base + o_gy1);
  h->f_partial = reinterpret_cast<double*>(base + o_fp);
  h->t_partial = reinterpret_cast<double*>(base + o_tp);
  h->f_counter = reinterpret_cast<unsigned*>(base + o_fc);
  h->t_counter = reinterpret_cast<unsigned*>(base + o_tc);
  h->ctrl = reinterpret_cast<Ctrl*>(base + o_ctrl);
  h->fault = reinterpret_cast<int32_t*>(base + o_fault);
  h->fin_f = reinterpret_cast<unsigned*>(base + o_fin);
  h->fin_t = h->fin_f + 1;
  h->rr_g = reinterpret_cast<double*>(base + o_rrg);
  h->scal_red = reinterpret_cast<double*>(base + o_sred);
  h->cnt_red = reinterpret_cast<double*>(base + o_cred);
  h->gd_flag = reinterpret_cast<unsigned long long*>(base + o_gd);
  h->gd_count = reinterpret_cast<unsigned*>(base + o_gd + 16);
  if (const char* e_pdl = getenv("GLPROX_PDL")) h->pdl = atoi(e_pdl) != 0;
  if (const char* e_fg = getenv("GLPROX_FUSED_GATHER")) h->fused_gather = atoi(e_fg) != 0;
  if (const char* e_pf = getenv("GLPROX_PUSH_FENCE")) h->push_fence_block = strcmp(e_pf, "block") == 0;
  h->prof_stride = h->pdl ? 4 : 1;
  if (const char* e_ps = getenv("GLPROX_PROF_STRIDE")) h->prof_stride = std::max(1, atoi(e_ps));

  // b: row-major [m][l] dtype T -> planar fp64 [l][ldm]
  if (dtype == GLP_F32) e = to_planar(h->stream, static_cast<const float*>(b_dev), h->b, m, l, g.ldm);
  else e = to_planar(h->stream, static_cast<const double*>(b_dev), h->b, m, l, g.ldm);
  if (e == cudaSuccess) e = cudaMallocHost(reinterpret_cast<void**>(&h->h_ctrl), 2 * sizeof(Ctrl));
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&h->ev_snap[i], cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreate(&h->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&h->ev1);
  if (e != cudaSuccess) { glp_destroy(h); return fail(GLP_ERR_CUDA, "handle setup failed: %s", cudaGetErrorString(e)); }
  *out = h;
  return GLP_OK;
}

// ------------------------------------------------------------------------------------------------
// stream enqueue helpers
// ------------------------------------------------------------------------------------------------
static int esize(const glp_handle* h) { return h->dtype == GLP_F32 ? 4 : 8; }

static bool multi(const glp_handle* h) { return h->comm.world > 1; }
static bool peer_push(const glp_handle* h) { return h->comm.world > 1 && !h->xch_nccl; }   // mailbox exchange
static int glp_allreduce_f64(glp_handle* h, double* buf, size_t count, bool is_max);

#include "glp_wide_host.inl"

static uint64_t seq_base(const glp_handle* h) { return h->epoch << 32; }

// Forward stream over `ncols` operand columns.  Xb[s] / AXb[s]: planar operand / output bases.
// exchange: multi-GPU solver launch (row blocks are pushed into the peers' slabs).
static int enqueue_fwd(glp_handle* h, const Ctrl* ctrl, int sel, const void* const Xb[2], double* const AXb[2],
                       const double* const xcand[2], int ncols, int count_mode, bool exchange) {
  const Geom& g = h->g;
  int chunk = 0;
  for (int q = 0; q < ncols; ++chunk) {
    const int rc = next_chunk(ncols - q);
    FwdArgs a{};
    for (int s = 0; s < 2; ++s) {
      a.X[s] = static_cast<const char*>(Xb[s]) + size_t(q) * g.ldn * esize(h);
      a.AX[s] = AXb[s] + size_t(q) * g.ldm;
      a.xcand[s] = xcand ? xcand[s] : nullptr;
    }
    a.ctrl = ctrl;
    a.sel = sel;
    a.nb_cols = std::max(0, std::min(rc, g.l - q));
    a.bcol0 = q;
    a.count_mode = (chunk == 0 && count_mode) ? 1 : 0;
    a.b = h->b;
    a.partial = h->f_partial;
    a.counter = h->f_counter;
    a.rr_partial = h->rr_partial + size_t(chunk) * g.nrb;
    a.scal_partial = h->scal_partial;
    a.cnt_partial = h->cnt_partial;
    a.fault = h->fault;
    if (exchange && peer_push(h)) {
      a.comm = h->comm;
      a.seq_f = seq_base(h) | ++h->idx_f;
      a.seq_t = h->seq_t_cur;
      a.par = h->group & 1;
      a.col0 = q;
      a.fin = h->fin_f;
      a.push_fence_per_block = h->push_fence_block ? 1 : 0;
    }
    CU_TRY(launch_fwd(h, rc, a));
    q += rc;
  }
  return GLP_OK;
}

// Transposed stream over the l residual columns + prox epilogue (fused when l <= 2).
static int enqueue_tpass(glp_handle* h, const Ctrl* ctrl, const ProxArgs* p, double* grad, bool exchange) {
  const Geom& g = h->g;
  ProxArgs pa{};
  if (p) pa = *p;
  const bool fuse = p != nullptr && g.l <= 2;
  if (p != nullptr && exchange && peer_push(h)) {
    pa.post.comm = h->comm;
    pa.post.seq_t = h->seq_t_cur = seq_base(h) | ++h->idx_t;
    pa.post.par = h->group & 1;
    pa.post.fin = h->fin_t;
  }
  for (int q = 0; q < g.l;) {
    const int lc = (g.l - q) >= 2 ? 2 : 1;
    TpassArgs a{};
    a.R = static_cast<const char*>(h->r) + size_t(q) * g.ldm * esize(h);
    a.grad = grad;
    a.gcol0 = q;
    a.fuse = fuse ? 1 : 0;
    a.ctrl = ctrl;
    a.partial = h->t_partial;
    a.counter = h->t_counter;
    a.fault = h->fault;
    CU_TRY(launch_tpass(h, lc, a, pa));
    q += lc;
  }
  if (p != nullptr && !fuse) CU_TRY(launch_prox(h, pa, 0));
  return GLP_OK;
}

static ProxArgs solver_prox_args(glp_handle* h, const Ctrl* ctrl) {
  ProxArgs p{};
  p.state[0] = h->state[0];
  p.state[1] = h->state[1];
  p.xop = h->xop;
  p.grad = h->grad;
  p.scal_partial = h->scal_partial;
  p.ctrl = ctrl;
  p.algo = h->sc.algo;
  p.thres = h->sc.thres;
  p.delta = h->sc.delta;
  return p;
}

// NCCL baseline exchange (glp_comm.cuh, bottom): packing kernels + two collectives per forward stream
static NcclXArgs nccl_args(glp_handle* h, const Ctrl* ctrl) {
  NcclXArgs a{};
  a.ctrl = ctrl;
  a.slab_words = int64_t(2) * h->g.l * h->g.ldm;
  a.stage = h->nccl_stage;
  a.stage_m = h->nccl_stage + a.slab_words + 8;
  a.scal_partial = h->scal_partial;
  a.ncb = h->g.ncb;
  a.xcand[0] = h->state[0]; a.xcand[1] = h->state[1];
  a.AX[0] = h->AX[0]; a.AX[1] = h->AX[1];
  a.b = h->b;
  a.rr_partial = h->rr_g;
  a.scal_red = h->scal_red;
  a.cnt_red = h->cnt_red;
  return a;
}
static int nccl_after_prox(glp_handle* h, const Ctrl* ctrl) {   // local scalar partials -> staging; MAX of max|xn|
  const NcclXArgs a = nccl_args(h, ctrl);
  glp_nccl_fold_kernel<<<1, 256, 0, h->stream>>>(a);
  CU_TRY(cudaGetLastError());
  h->n_launches++;
  return glp_allreduce_f64(h, a.stage_m, 1, true);
}

static int enqueue_solver_fwd(glp_handle* h, const Ctrl* ctrl) {
  if (wide_active(h)) return enqueue_wide_fwd(h, ctrl);
  const void* Xb[2];
  const double* xc[2] = {h->state[0], h->state[1]};
  for (int s = 0; s < 2; ++s) Xb[s] = h->dtype == GLP_F32 ? h->xop : static_cast<const void*>(h->state[s]);
  if (h->xch_nccl) {   // partial product into the staging buffer, count against the global max, one SUM all-reduce
    const NcclXArgs a = nccl_args(h, ctrl);
    double* stage2[2] = {a.stage, a.stage};
    GLP_TRY(enqueue_fwd(h, ctrl, 0, Xb, stage2, xc, 2 * h->g.l, 0, false));
    glp_nccl_count_kernel<<<h->gather_grid, 256, 0, h->stream>>>(a, h->g);
    CU_TRY(cudaGetLastError());
    h->n_launches++;
    return glp_allreduce_f64(h, a.stage, size_t(a.slab_words) + 8, false);
  }
  return enqueue_fwd(h, ctrl, 0, Xb, h->AX, xc, 2 * h->g.l, 1, true);
}

// Gather arguments of the forward group that was launched last.
static GatherArgs gather_args(const glp_handle* h, const Ctrl* ctrl) {
  GatherArgs a{};
  a.comm = h->comm;
  a.ctrl = ctrl;
  a.sel = 0;
  a.par = h->group & 1;
  a.seq_f = seq_base(h) | h->idx_f;
  a.seq_t = h->seq_t_cur;
  a.AX[0] = h->AX[0]; a.AX[1] = h->AX[1];
  a.b = h->b;
  a.rr_partial = h->rr_g;
  a.scal_red = h->scal_red;
  a.cnt_red = h->cnt_red;
  a.fault = h->fault;
  return a;
}
// l = 2 streaming path over peer mailboxes: gather and decide run as one launch
static bool fused_gather_active(const glp_handle* h) {
  return h->fused_gather && peer_push(h) && !wide_active(h);
}

// Multi-GPU: sum the W slabs of the forward group that was just launched (rank order) into the
// candidate product buffers, with the residual norm and the exchanged scalars.
static int enqueue_gather(glp_handle* h, const Ctrl* ctrl) {
  if (!multi(h)) return GLP_OK;
  if (h->xch_nccl) {
    glp_nccl_post_kernel<<<h->gather_grid, GATHER_THREADS, 0, h->stream>>>(nccl_args(h, ctrl), h->g);
    CU_TRY(cudaGetLastError());
    h->n_launches++;
    return GLP_OK;
  }
  if (wide_active(h)) return enqueue_wide_exchange(h, ctrl, 1);
  if (fused_gather_active(h)) return GLP_OK;   // the next decide launch gathers first (glp_gather_decide_kernel)
  const GatherArgs a = gather_args(h, ctrl);
  h->n_launches++;
  CU_TRY(launch_k(h, glp_gather_kernel, dim3(h->gather_grid), dim3(GATHER_THREADS), 0, a, h->g));
  return GLP_OK;
}

// ------------------------------------------------------------------------------------------------
// whole solve, as phases: a multi-rank job on ONE device (tests) enqueues each phase for all of its
// logical ranks before the next one, so that no launch ever waits for a later launch of the stream.
// ------------------------------------------------------------------------------------------------
static int phase_prepare(glp_handle* h, int algo, const void* x0_dev, double mu0, const glp_opts* o) {
  if (h == nullptr || x0_dev == nullptr || o == nullptr) return fail(GLP_ERR_ARG, "null argument");
  if (algo < GLP_PROXGD || algo > GLP_SGD) return fail(GLP_ERR_ARG, "unknown algo %d", algo);
  if (o->step_type < GLP_STEP_LINE_SEARCH || o->step_type > GLP_STEP_BB)
    return fail(GLP_ERR_ARG, "unsupported step_type %d", o->step_type);
  if (o->step_type == GLP_STEP_BB && algo != GLP_PROXGD)
    return fail(GLP_ERR_ARG, "step_type bb is an opt-in of gl_ProxGD_primal only");
  if ((algo == GLP_GD || algo == GLP_SGD) && o->step_type == GLP_STEP_LINE_SEARCH)
    return fail(GLP_ERR_ARG, "gl_GD_primal / gl_SGD_primal have no line search (step_type fixed | diminishing | diminishing2)");
  if (o->maxit < 0 || o->maxit > (1 << 26)) return fail(GLP_ERR_ARG, "maxit out of range");
  CU_TRY(cudaSetDevice(h->device));
  const Geom& g = h->g;
  SolveConst& sc = h->sc;
  sc.algo = algo; sc.step_type = o->step_type; sc.maxit = o->maxit; sc.stable_cap = o->stable_len_threshold;
  sc.ls_cap = o->maxit_line_search_iter; sc.thres = o->thres; sc.alpha0 = o->alpha0; sc.ftol = o->ftol;
  sc.atten = o->line_search_attenuation_coeffi; sc.mu0 = mu0; sc.delta = o->delta;
  sc.hist_cap = 3 * o->maxit + 8;
  if (h->hist_cap < sc.hist_cap) {
    if (h->f_hist) CU_TRY(cudaFree(h->f_hist));
    h->f_hist = nullptr;
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->f_hist), size_t(3) * sc.hist_cap * sizeof(double)));
    h->hist_cap = sc.hist_cap;
  }
  h->f_hist_best = h->f_hist + h->hist_cap;
  h->s_hist = h->f_hist + 2 * size_t(h->hist_cap);
  h->last_k = 0;
  h->n_launches = 0;
  h->n_gemm = 0;
  if (wide_active(h)) GLP_TRY(ensure_wide(h));
  h->epoch++;
  h->idx_f = h->idx_t = 0;
  h->group = 0;
  CU_TRY(cudaEventRecord(h->ev0, h->stream));

  Ctrl c0;
  memset(&c0, 0, sizeof(c0));
  ctrl_init(c0, sc);
  c0.n_fwd = 1;
  CU_TRY(cudaMemcpyAsync(&h->ctrl[0], &c0, sizeof(Ctrl), cudaMemcpyHostToDevice, h->stream));
  CU_TRY(cudaMemcpyAsync(&h->ctrl[1], &c0, sizeof(Ctrl), cudaMemcpyHostToDevice, h->stream));
  CU_TRY(cudaMemsetAsync(h->fault, 0, sizeof(int32_t), h->stream));
  const int part_grid = std::max(std::max(h->decide_grid, h->gather_grid), h->fused_grid);
  CU_TRY(cudaMemsetAsync(h->gy_partial[0], 0, size_t(part_grid) * 8, h->stream));
  CU_TRY(cudaMemsetAsync(h->gy_partial[1], 0, size_t(part_grid) * 8, h->stream));
  h->prof_slot = true;
  // x0 -> candidate slot of cur = 0
  if (h->dtype == GLP_F32) CU_TRY(to_planar(h->stream, static_cast<const float*>(x0_dev), h->state[1], g.n, g.l, g.ldn));
  else CU_TRY(to_planar(h->stream, static_cast<const double*>(x0_dev), h->state[1], g.n, g.l, g.ldn));
  if (peer_push(h)) {
    // every rank has drained its previous solve before it gets here: after the epoch barrier the
    // mailboxes may be reused from parity 0
    glp_epoch_post_kernel<<<1, 32, 0, h->stream>>>(h->comm, h->epoch);
    CU_TRY(cudaGetLastError());
  }
  h->decide_launches = 0;
  h->cur_ctrl = &h->ctrl[0];
  h->solving = true;
  return GLP_OK;
}

static int nccl_after_prox(glp_handle* h, const Ctrl* ctrl);
// threshold + scalars of x0 (init epilogue)
static int phase_init_prox(glp_handle* h) {
  if (peer_push(h)) {
    glp_epoch_wait_kernel<<<1, 32, 0, h->stream>>>(h->comm, h->epoch, h->fault);
    CU_TRY(cudaGetLastError());
  }
  ProxArgs p = solver_prox_args(h, nullptr);
  p.cur = 0;
  p.init = 1;
  if (peer_push(h)) {
    p.post.comm = h->comm;
    p.post.seq_t = h->seq_t_cur = seq_base(h) | ++h->idx_t;
    p.post.par = h->group & 1;
    p.post.fin = h->fin_t;
  }
  CU_TRY(launch_prox(h, p, 0));
  if (h->xch_nccl) GLP_TRY(nccl_after_prox(h, nullptr));
  return GLP_OK;
}

static int phase_decide(glp_handle* h) {
  const int64_t j = h->decide_launches++;
  const Ctrl* cin = &h->ctrl[j & 1];
  Ctrl* cout = &h->ctrl[(j + 1) & 1];
  DecideArgs d{};
  d.cin = cin; d.cout = cout; d.sc = h->sc;
  h->prof_slot = (j % h->prof_stride) == 0;
  const bool fused = fused_gather_active(h);
  if (fused) {
    d.rr_partial = h->rr_g; d.n_rr = h->fused_grid;
    d.scal_partial = h->scal_red; d.n_scal = 1;
    d.cnt_partial = h->cnt_red; d.n_cnt = 1;
  } else if (multi(h)) {
    d.rr_partial = h->rr_g; d.n_rr = h->gather_grid;
    d.scal_partial = h->scal_red; d.n_scal = 1;
    d.cnt_partial = h->cnt_red; d.n_cnt = 1;
  } else if (wide_active(h)) {
    d.rr_partial = h->rr_g; d.n_rr = h->gather_grid;
    d.scal_partial = h->scal_partial; d.n_scal = h->g.ncb;
    d.cnt_partial = h->cnt_g; d.n_cnt = h->gather_grid;
  } else {
    d.rr_partial = h->rr_partial; d.n_rr = h->n_fchunks * h->g.nrb;
    d.scal_partial = h->scal_partial; d.n_scal = h->g.ncb;
    d.cnt_partial = h->cnt_partial; d.n_cnt = h->g.grid;
  }
  d.gy_partial[0] = h->gy_partial[0]; d.gy_partial[1] = h->gy_partial[1]; d.n_gy = fused ? h->fused_grid : h->decide_grid;
  d.AX[0] = h->AX[0]; d.AX[1] = h->AX[1]; d.Av = h->Av; d.b = h->b; d.r = h->r;
  d.state[0] = h->state[0]; d.state[1] = h->state[1];
  d.f_hist = h->f_hist; d.f_hist_best = h->f_hist_best; d.s_hist = h->s_hist;
  if (wide_active(h) && wide_passes(h) == 4) {
    CU_TRY(cudaMemsetAsync(h->f16_amax + 1, 0, sizeof(unsigned), h->stream));
    d.r_amax = h->f16_amax + 1;
  }
  if (fused) CU_TRY(launch_gather_decide(h, gather_args(h, cin), d));
  else CU_TRY(launch_decide(h, d));
  h->cur_ctrl = cout;
  return GLP_OK;
}

static int phase_tpass(glp_handle* h) {
  h->group++;
  ProxArgs p = solver_prox_args(h, h->cur_ctrl);
  if (wide_active(h)) {
    GLP_TRY(enqueue_wide_tpass(h, h->cur_ctrl));
    if (multi(h)) {   // the prox tail posts this rank's scalars to every peer (as enqueue_tpass does)
      p.post.comm = h->comm;
      p.post.seq_t = h->seq_t_cur = seq_base(h) | ++h->idx_t;
      p.post.par = h->group & 1;
      p.post.fin = h->fin_t;
    }
    CU_TRY(launch_prox(h, p, 0));
    return GLP_OK;
  }
  GLP_TRY(enqueue_tpass(h, h->cur_ctrl, &p, h->grad, true));
  if (h->xch_nccl) GLP_TRY(nccl_after_prox(h, h->cur_ctrl));
  return GLP_OK;
}
static int phase_fwd(glp_handle* h) { return enqueue_solver_fwd(h, h->cur_ctrl); }
static int phase_gather(glp_handle* h) { return enqueue_gather(h, h->cur_ctrl); }
static int phase_finish(glp_handle* h) {   // third exchange stage of the wide multi-GPU mode, otherwise nothing
  return (multi(h) && wide_active(h)) ? enqueue_wide_exchange(h, h->cur_ctrl, 2) : GLP_OK;
}

extern "C" int glp_solve_begin(glp_handle* h, int algo, const void* x0_dev, double mu0, const glp_opts* o) {
  GLP_TRY(phase_prepare(h, algo, x0_dev, mu0, o));
  h->backend_used = GLP_BACKEND_STREAM;
  GLP_TRY(phase_init_prox(h));
  GLP_TRY(phase_fwd(h));
  GLP_TRY(phase_gather(h));
  GLP_TRY(phase_finish(h));
  return GLP_OK;
}

extern "C" int glp_solve_step(glp_handle* h, int slots) {
  if (h == nullptr || !h->solving) return fail(GLP_ERR_STATE, "glp_solve_step before glp_solve_begin");
  CU_TRY(cudaSetDevice(h->device));
  for (int s = 0; s < slots; ++s) {
    GLP_TRY(phase_decide(h));
    GLP_TRY(phase_tpass(h));
    GLP_TRY(phase_fwd(h));
    GLP_TRY(phase_gather(h));
    GLP_TRY(phase_finish(h));
  }
  return GLP_OK;
}

static void fill_result(const glp_handle* h, const Ctrl& c, glp_result* res) {
  if (res == nullptr) return;
  res->iters = c.k;
  res->done = c.done;
  res->backend_used = h->backend_used;
  res->fval = c.f_cur;
  res->sparsity = c.s_cur;
  res->n_fwd = c.n_fwd;
  res->n_tpass = c.n_tpass;
  res->n_launches = h->n_launches;
  for (int i = 0; i < 3; ++i) res->stage_end[i] = c.stage_end[i];
}

static int check_fault(glp_handle* h, cudaError_t sync_err) {
  if (sync_err != cudaSuccess) {
    h->solving = false;
    return fail(GLP_ERR_DEVICE_FAULT, "kernel fault during solve: %s", cudaGetErrorString(sync_err));
  }
  return GLP_OK;
}

extern "C" int glp_solve_poll(glp_handle* h, glp_result* res) {
  if (h == nullptr || !h->solving) return fail(GLP_ERR_STATE, "glp_solve_poll before glp_solve_begin");
  CU_TRY(cudaSetDevice(h->device));
  CU_TRY(cudaMemcpyAsync(&h->h_ctrl[0], &h->ctrl[h->decide_launches & 1], sizeof(Ctrl), cudaMemcpyDeviceToHost, h->stream));
  GLP_TRY(check_fault(h, cudaStreamSynchronize(h->stream)));
  if (res) { memset(res, 0, sizeof(*res)); fill_result(h, h->h_ctrl[0], res); }
  return GLP_OK;
}

extern "C" int glp_solve_end(glp_handle* h, void* x_out_dev, glp_result* res, double* f_hist_host, int hist_cap) {
  if (h == nullptr || !h->solving) return fail(GLP_ERR_STATE, "glp_solve_end before glp_solve_begin");
  CU_TRY(cudaSetDevice(h->device));
  const Geom& g = h->g;
  CU_TRY(cudaMemcpyAsync(&h->h_ctrl[0], &h->ctrl[h->decide_launches & 1], sizeof(Ctrl), cudaMemcpyDeviceToHost, h->stream));
  int32_t fault = 0;
  CU_TRY(cudaMemcpyAsync(&fault, h->fault, sizeof(fault), cudaMemcpyDeviceToHost, h->stream));
  GLP_TRY(check_fault(h, cudaStreamSynchronize(h->stream)));
  const Ctrl c = h->h_ctrl[0];
  h->solving = false;
  h->last_k = std::min(c.k, h->sc.hist_cap);
  if (res) memset(res, 0, sizeof(*res));
  if (fault != 0) {
    if (res) res->status = GLP_ERR_DEVICE_FAULT;
    return fail(GLP_ERR_DEVICE_FAULT, fault == 2 ? "a peer rank never raised its mailbox flag (20 s watchdog)"
                                                 : "pipeline watchdog fired inside a streaming kernel");
  }
  if (x_out_dev != nullptr) {
    const double* x = h->state[c.cur];
    if (h->dtype == GLP_F32) CU_TRY(from_planar(h->stream, x, static_cast<float*>(x_out_dev), g.n, g.l, g.ldn));
    else CU_TRY(from_planar(h->stream, x, static_cast<double*>(x_out_dev), g.n, g.l, g.ldn));
  }
  if (f_hist_host != nullptr && hist_cap > 0 && c.k > 0)
    CU_TRY(cudaMemcpyAsync(f_hist_host, h->f_hist, sizeof(double) * std::min(hist_cap, std::min(c.k, h->sc.hist_cap)),
                           cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaEventRecord(h->ev1, h->stream));
  GLP_TRY(check_fault(h, cudaStreamSynchronize(h->stream)));
  float ms = 0.f;
  CU_TRY(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
  prof_collect(h);
  fill_result(h, c, res);
  if (res) res->device_ms = ms;
  if ((h->sc.algo == GLP_FPROXGD || h->sc.algo == GLP_FGD) && !c.ran_any) {
    if (res) res->status = GLP_ERR_UNBOUND;
    return fail(GLP_ERR_UNBOUND, "FProxGD ran no iteration (maxit == 0): the reference raises UnboundLocalError");
  }
  if (res) res->status = c.done ? GLP_OK : GLP_ERR_STATE;
  return GLP_OK;
}

#include "glp_comm_host.inl"

static int drive_until_done(glp_handle** hs, int G, const glp_opts* o) {
  // Host runs ahead of the device: batch i+1 is enqueued before the control snapshot taken after
  // batch i is inspected, so the stream never drains while the solve is in progress.
  glp_handle* h0 = hs[0];
  // a negative maxit_line_search_iter is an empty range() in the reference (alpha0 taken untested): one trial per iteration
  const int64_t max_slots = int64_t(3) * o->maxit * (std::max<int64_t>(o->maxit_line_search_iter, 0) + 1) + 8;
  const int batch = o->poll_interval > 0 ? o->poll_interval : 32;
  int64_t issued = 0;
  int pending = -1;  // snapshot slot whose event has not been inspected yet
  bool done = false;
  int cur_slot = 0;
  while (!done && issued < max_slots) {
    for (int s = 0; s < batch; ++s) {
      for (int i = 0; i < G; ++i) { CU_TRY(cudaSetDevice(hs[i]->device)); GLP_TRY(phase_decide(hs[i])); }
      for (int i = 0; i < G; ++i) { CU_TRY(cudaSetDevice(hs[i]->device)); GLP_TRY(phase_tpass(hs[i])); }
      for (int i = 0; i < G; ++i) { CU_TRY(cudaSetDevice(hs[i]->device)); GLP_TRY(phase_fwd(hs[i])); }
      for (int i = 0; i < G; ++i) { CU_TRY(cudaSetDevice(hs[i]->device)); GLP_TRY(phase_gather(hs[i])); }
      for (int i = 0; i < G; ++i) { CU_TRY(cudaSetDevice(hs[i]->device)); GLP_TRY(phase_finish(hs[i])); }
    }
    issued += batch;
    CU_TRY(cudaSetDevice(h0->device));
    CU_TRY(cudaMemcpyAsync(&h0->h_ctrl[cur_slot], &h0->ctrl[h0->decide_launches & 1], sizeof(Ctrl), cudaMemcpyDeviceToHost, h0->stream));
    CU_TRY(cudaEventRecord(h0->ev_snap[cur_slot], h0->stream));
    if (pending >= 0) {
      GLP_TRY(check_fault(h0, cudaEventSynchronize(h0->ev_snap[pending])));
      if (h0->h_ctrl[pending].done) done = true;
    }
    pending = cur_slot;
    cur_slot ^= 1;
  }
  return GLP_OK;
}

// ---- persistent backend: the whole solve in one cluster launch (glp_persist.cuh) -----------------
struct PersistPlan { int C; int nc; size_t smem; };

static bool persist_plan(const glp_handle* h, PersistPlan* out) {
  if (multi(h) || h->g.l > PERSIST_MAX_L) return false;
  int want = 0;
  if (const char* e = getenv("GLPROX_CLUSTER")) want = atoi(e);
  const int cands[2] = {8, 16};
  for (int C : cands) {
    if (want != 0 && want != C) continue;
    int nc = (h->g.n + C - 1) / C;
    nc += nc & 1;
    if (nc < 2) nc = 2;
    if (nc > PERSIST_MAX_NC) continue;
    const PersistLayout lay = persist_layout(h->g.m, nc, h->g.l, h->dtype == GLP_F32 ? 4 : 8);
    if (lay.total > size_t(227) * 1024) continue;
    out->C = C; out->nc = nc; out->smem = lay.total;
    return true;
  }
  return false;
}

template <typename T, int L>
static cudaError_t launch_persist_t(glp_handle* h, const PersistPlan& pl, const PersistArgs& a) {
  auto kfn = glp_persist_kernel<T, L>;
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, int(pl.smem));
  if (e != cudaSuccess) return e;
  if (pl.C > 8) {
    e = cudaFuncSetAttribute(kfn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(pl.C), 1, 1);
  cfg.blockDim = dim3(PERSIST_THREADS, 1, 1);
  cfg.dynamicSmemBytes = pl.smem;
  cfg.stream = h->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = unsigned(pl.C);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  e = cudaLaunchKernelEx(&cfg, kfn, a);
  h->n_launches++;
  return e;
}

static int enqueue_persist(glp_handle* h, const PersistPlan& pl) {
  PersistArgs a{};
  a.A = h->A; a.lda = h->lda; a.b = h->b;
  a.x0 = h->state[1]; a.x_out = h->state[0];
  a.ctrl_out = &h->ctrl[0];
  a.f_hist = h->f_hist; a.f_hist_best = h->f_hist_best; a.s_hist = h->s_hist;
  a.sc = h->sc;
  a.m = h->g.m; a.n = h->g.n; a.l = h->g.l; a.nc = pl.nc;
  a.ldm = h->g.ldm; a.ldn = h->g.ldn;
  cudaError_t e = cudaErrorInvalidValue;
  const bool f32 = h->dtype == GLP_F32;
  switch (h->g.l) {
    case 1: e = f32 ? launch_persist_t<float, 1>(h, pl, a) : launch_persist_t<double, 1>(h, pl, a); break;
    case 2: e = f32 ? launch_persist_t<float, 2>(h, pl, a) : launch_persist_t<double, 2>(h, pl, a); break;
    case 3: e = f32 ? launch_persist_t<float, 3>(h, pl, a) : launch_persist_t<double, 3>(h, pl, a); break;
    case 4: e = f32 ? launch_persist_t<float, 4>(h, pl, a) : launch_persist_t<double, 4>(h, pl, a); break;
    default: break;
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(GLP_ERR_CUDA, "persistent cluster launch (C=%d, %zu B smem) failed: %s", pl.C, pl.smem, cudaGetErrorString(e));
  }
  return GLP_OK;
}

extern "C" int glp_solve(glp_handle* h, int algo, const void* x0_dev, double mu0, const glp_opts* o, void* x_out_dev,
                         glp_result* res, double* f_hist_host, int hist_cap) {
  if (h == nullptr || o == nullptr) return fail(GLP_ERR_ARG, "null argument");
  if (o->backend < GLP_BACKEND_AUTO || o->backend > GLP_BACKEND_PERSISTENT) return fail(GLP_ERR_ARG, "unknown backend %d", o->backend);
  PersistPlan pl{};
  const bool fits = persist_plan(h, &pl);
  if (o->backend == GLP_BACKEND_PERSISTENT && !fits)
    return fail(GLP_ERR_ARG, "backend 'persistent' needs a single-GPU problem with l <= %d whose A fits in one cluster's "
                "shared memory (m=%d n=%d l=%d does not)", PERSIST_MAX_L, h->g.m, h->g.n, h->g.l);
  GLP_TRY(phase_prepare(h, algo, x0_dev, mu0, o));
  int rc;
  if (fits && o->backend != GLP_BACKEND_STREAM) {
    h->backend_used = GLP_BACKEND_PERSISTENT;
    rc = enqueue_persist(h, pl);
  } else {
    h->backend_used = GLP_BACKEND_STREAM;
    rc = phase_init_prox(h);
    if (rc == GLP_OK) rc = phase_fwd(h);
    if (rc == GLP_OK) rc = phase_gather(h);
    if (rc == GLP_OK) rc = phase_finish(h);
    if (rc == GLP_OK) rc = drive_until_done(&h, 1, o);
  }
  if (rc != GLP_OK) { h->solving = false; return rc; }
  return glp_solve_end(h, x_out_dev, res, f_hist_host, hist_cap);
}

// G logical ranks driven by one host thread.  Handles on one device must share a stream (then every
// wait is already satisfied when its kernel starts); handles on different devices run concurrently.
extern "C" int glp_group_solve(glp_handle** hs, int G, int algo, const void* const* x0_dev, double mu0,
                               const glp_opts* o, void* const* x_out_dev, glp_result* res, double* f_hist_host,
                               int hist_cap) {
  if (hs == nullptr || x0_dev == nullptr || o == nullptr || G < 1 || G > MAX_WORLD) return fail(GLP_ERR_ARG, "bad argument");
  for (int i = 0; i < G; ++i) {
    if (hs[i] == nullptr) return fail(GLP_ERR_ARG, "null handle %d", i);
    if (G > 1 && (hs[i]->comm.world != G || hs[i]->comm.rank != i))
      return fail(GLP_ERR_STATE, "handle %d is not attached as rank %d of %d", i, i, G);
    for (int j = 0; j < i; ++j)
      if (hs[j]->device == hs[i]->device && hs[j]->stream != hs[i]->stream)
        return fail(GLP_ERR_ARG, "logical ranks that share a device must share a stream");
  }
  int rc = GLP_OK;
  for (int i = 0; i < G && rc == GLP_OK; ++i) { rc = phase_prepare(hs[i], algo, x0_dev[i], mu0, o); hs[i]->backend_used = GLP_BACKEND_STREAM; }
  for (int i = 0; i < G && rc == GLP_OK; ++i) { cudaSetDevice(hs[i]->device); rc = phase_init_prox(hs[i]); }
  for (int i = 0; i < G && rc == GLP_OK; ++i) { cudaSetDevice(hs[i]->device); rc = phase_fwd(hs[i]); }
  for (int i = 0; i < G && rc == GLP_OK; ++i) { cudaSetDevice(hs[i]->device); rc = phase_gather(hs[i]); }
  for (int i = 0; i < G && rc == GLP_OK; ++i) { cudaSetDevice(hs[i]->device); rc = phase_finish(hs[i]); }
  if (rc == GLP_OK) rc = drive_until_done(hs, G, o);
  if (rc != GLP_OK) {
    for (int i = 0; i < G; ++i) hs[i]->solving = false;
    return rc;
  }
  for (int i = 0; i < G; ++i) {
    int r2 = glp_solve_end(hs[i], x_out_dev ? x_out_dev[i] : nullptr, res ? &res[i] : nullptr,
                           i == 0 ? f_hist_host : nullptr, i == 0 ? hist_cap : 0);
    if (r2 != GLP_OK && rc == GLP_OK) rc = r2;
  }
  return rc;
}

// ------------------------------------------------------------------------------------------------
// step-level entry points (kernel parity tests drive these)
// ------------------------------------------------------------------------------------------------
static int ensure_scratch(glp_handle* h, size_t bytes) {
  if (h->scratch_bytes >= bytes) return GLP_OK;
  if (h->scratch) CU_TRY(cudaFree(h->scratch));
  h->scratch = nullptr;
  h->scratch_bytes = 0;
  CU_TRY(cudaMalloc(&h->scratch, bytes));
  h->scratch_bytes = bytes;
  return GLP_OK;
}

__global__ void sum_partials_kernel(const double* __restrict__ part, int n, int stride, int offset, int is_max,
                                    double* __restrict__ out) {
  // single CTA, fixed order: thread-strided, warp tree, warps in order
  __shared__ double s[8];
  double v = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const double x = part[size_t(i) * stride + offset];
    v = is_max ? ((v != v) ? v : ((x != x) ? x : fmax(v, x))) : v + x;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const double y = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? ((v != v) ? v : ((y != y) ? y : fmax(v, y))) : v + y;
  }
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = s[0];
    for (int w = 1; w < int(blockDim.x >> 5); ++w) t = is_max ? ((t != t) ? t : ((s[w] != s[w]) ? s[w] : fmax(t, s[w]))) : t + s[w];
    *out = t;
  }
}

extern "C" int glp_fwd(glp_handle* h, const void* X_dev, int ncols, double* AX_dev, double* rr_dev) {
  if (h == nullptr || X_dev == nullptr || AX_dev == nullptr || ncols <= 0) return fail(GLP_ERR_ARG, "bad argument");
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  CU_TRY(cudaSetDevice(h->device));
  const Geom& g = h->g;
  const int es = esize(h);
  const size_t xb = align_up(size_t(ncols) * g.ldn * es, 256), ab = align_up(size_t(ncols) * g.ldm * 8, 256);
  const int nch = count_fchunks(ncols);
  const size_t rb = align_up(size_t(nch) * g.nrb * 8, 256);
  GLP_TRY(ensure_scratch(h, xb + ab + rb));
  char* base = static_cast<char*>(h->scratch);
  void* Xp = base;
  double* AXp = reinterpret_cast<double*>(base + xb);
  double* rrp = reinterpret_cast<double*>(base + xb + ab);
  CU_TRY(cudaMemsetAsync(base, 0, xb + ab + rb, h->stream));
  if (h->dtype == GLP_F32) CU_TRY(to_planar(h->stream, static_cast<const float*>(X_dev), static_cast<float*>(Xp), g.n, ncols, g.ldn));
  else CU_TRY(to_planar(h->stream, static_cast<const double*>(X_dev), static_cast<double*>(Xp), g.n, ncols, g.ldn));
  const void* Xb[2] = {Xp, Xp};
  double* AXb[2] = {AXp, AXp};
  double* saved_rr = h->rr_partial;
  h->rr_partial = rrp;  // chunk-indexed inside enqueue_fwd
  int rc = enqueue_fwd(h, nullptr, 0, Xb, AXb, nullptr, ncols, 0, false);
  h->rr_partial = saved_rr;
  GLP_TRY(rc);
  CU_TRY(from_planar(h->stream, AXp, AX_dev, g.m, ncols, g.ldm));
  if (rr_dev != nullptr) {
    sum_partials_kernel<<<1, 256, 0, h->stream>>>(rrp, nch * g.nrb, 1, 0, 0, rr_dev);
    CU_TRY(cudaGetLastError());
  }
  return GLP_OK;
}

extern "C" int glp_tpass_prox(glp_handle* h, const glp_step_params* sp, const double* R_dev, double* G_dev,
                              const double* xt_dev, const double* v_dev, double* xn_dev, double* xnt_dev,
                              double* vn_dev, double* scalars_dev) {
  if (h == nullptr || R_dev == nullptr) return fail(GLP_ERR_ARG, "bad argument");
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  if (sp != nullptr && (xt_dev == nullptr || xn_dev == nullptr || xnt_dev == nullptr || scalars_dev == nullptr))
    return fail(GLP_ERR_ARG, "epilogue requested but xt/xn/xnt/scalars missing");
  if (sp != nullptr && sp->algo == GLP_FPROXGD && (v_dev == nullptr || vn_dev == nullptr))
    return fail(GLP_ERR_ARG, "FProxGD epilogue needs v and vn");
  CU_TRY(cudaSetDevice(h->device));
  const Geom& g = h->g;
  const size_t ln = size_t(g.l) * g.ldn;
  // residual -> planar dtype T
  if (h->dtype == GLP_F32) CU_TRY(to_planar(h->stream, R_dev, static_cast<float*>(h->r), g.m, g.l, g.ldm));
  else CU_TRY(to_planar(h->stream, R_dev, static_cast<double*>(h->r), g.m, g.l, g.ldm));
  ProxArgs p{};
  if (sp != nullptr) {
    CU_TRY(to_planar(h->stream, xt_dev, h->state[0] + ln, g.n, g.l, g.ldn));
    if (v_dev) CU_TRY(to_planar(h->stream, v_dev, h->state[0] + 2 * ln, g.n, g.l, g.ldn));
    p = solver_prox_args(h, nullptr);
    p.algo = sp->algo; p.cur = 0; p.init = 0;
    p.theta = sp->theta; p.t = sp->t; p.mu = sp->mu; p.thres = sp->thres; p.delta = 0.0;
  }
  GLP_TRY(enqueue_tpass(h, nullptr, sp ? &p : nullptr, h->grad, false));
  if (G_dev) CU_TRY(from_planar(h->stream, h->grad, G_dev, g.n, g.l, g.ldn));
  if (sp != nullptr) {
    CU_TRY(from_planar(h->stream, h->state[1], xn_dev, g.n, g.l, g.ldn));
    CU_TRY(from_planar(h->stream, h->state[1] + ln, xnt_dev, g.n, g.l, g.ldn));
    if (vn_dev && sp->algo == GLP_FPROXGD) CU_TRY(from_planar(h->stream, h->state[1] + 2 * ln, vn_dev, g.n, g.l, g.ldn));
    for (int k = 0; k < 4; ++k) {
      sum_partials_kernel<<<1, 256, 0, h->stream>>>(h->scal_partial, g.ncb, SCAL_STRIDE, k, k == 3 ? 1 : 0, scalars_dev + k);
      CU_TRY(cudaGetLastError());
    }
  }
  return GLP_OK;
}

extern "C" int glp_set_tensor_mode(glp_handle* h, int mode) {
  if (h == nullptr) return fail(GLP_ERR_ARG, "null handle");
  if (mode < 0 || mode > 5)
    return fail(GLP_ERR_ARG, "tensor mode must be 0 (auto: fp16 pairs), 1 (off), 2 (plain TF32), 3 (TF32 + bf16 corrections), "
                             "4 (fp16 pairs) or 5 (3xTF32)");
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  h->tensor_mode = mode;
  return GLP_OK;
}

extern "C" int glp_set_profiling(glp_handle* h, int on) {
  if (h == nullptr) return fail(GLP_ERR_ARG, "null handle");
  h->profiling = on != 0;
  h->ev_used = 0;
  memset(&h->prof, 0, sizeof(h->prof));
  return GLP_OK;
}

extern "C" int glp_get_profile(glp_handle* h, glp_profile* out) {
  if (h == nullptr || out == nullptr) return fail(GLP_ERR_ARG, "null argument");
  *out = h->prof;
  out->bytes_per_launch = double(h->g.m) * double(h->g.n) * esize(h);
  return GLP_OK;
}

extern "C" int glp_debug_read(glp_handle* h, int which, double* out_host) {
  if (h == nullptr || out_host == nullptr) return fail(GLP_ERR_ARG, "bad argument");
  CU_TRY(cudaSetDevice(h->device));
  const Geom& g = h->g;
  Ctrl c;
  CU_TRY(cudaMemcpyAsync(&c, &h->ctrl[h->decide_launches & 1], sizeof(Ctrl), cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  const size_t ln = size_t(g.l) * g.ldn, lm = size_t(g.l) * g.ldm;
  const double* src = nullptr;
  bool is_m = false;
  switch (which) {
    case 0: src = h->state[c.cur]; break;
    case 1: src = h->state[c.cur] + ln; break;
    case 2: src = h->state[c.cur] + 2 * ln; break;
    case 3: src = h->AX[c.cur]; is_m = true; break;
    case 4: src = h->AX[c.cur] + lm; is_m = true; break;
    case 5: src = h->Av; is_m = true; break;
    case 7: src = h->grad; break;
    case 8: src = h->state[1 - c.cur]; break;
    case 9: src = h->state[1 - c.cur] + ln; break;
    case 10: src = h->state[1 - c.cur] + 2 * ln; break;
    case 11: src = h->AX[1 - c.cur]; is_m = true; break;
    case 12: src = h->AX[1 - c.cur] + lm; is_m = true; break;
    default: return fail(GLP_ERR_ARG, "unknown vector id %d", which);
  }
  const int64_t rows = is_m ? g.m : g.n, ld = is_m ? g.ldm : g.ldn;
  GLP_TRY(ensure_scratch(h, size_t(rows) * g.l * 8));
  CU_TRY(from_planar(h->stream, src, static_cast<double*>(h->scratch), rows, g.l, ld));
  CU_TRY(cudaMemcpyAsync(out_host, h->scratch, size_t(rows) * g.l * 8, cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  return GLP_OK;
}

#include "glp_host_io.inl"
#include "glp_gemm_host.inl"
#include "glp_admm_host.inl"
#include "glp_admm2_host.inl"

#include "glp_metrics_host.inl"

// ------------------------------------------------------------------------------------------------
// multi-GPU plumbing: NCCL is bound at run time (dlopen) so that the library loads on machines
// without NCCL and has no link-time dependency on torch's bundled copy.
// ------------------------------------------------------------------------------------------------
#include "glp_nccl.inl"

#include "glp_abi_step.inl"

// allocation-free metric reductions (fixed device scratch; no counterpart in the reference)
#include "glp_noalloc.inl"
