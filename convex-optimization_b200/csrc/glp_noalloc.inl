// glp_noalloc.inl -- an allocation-free variant of the driver-metric reductions, included by glprox.cu (same translation
// unit as the array below: cudaGetSymbolAddress must be called from the file that defines the symbol).
//
// No counterpart in the reference.  glp_metrics_eval (glp_metrics_host.inl)
// calls cudaMalloc / cudaFree for its per-CTA records on every call, which code under CUDA graph capture, tight serving
// loops and harnesses that forbid allocation cannot do.  This entry point runs the SAME three reduction kernels
// (glp_metrics.cuh: sums, count, fold -- same grid, same fixed summation order, so the six values are bit-identical to
// glp_metrics_eval's) with the records in a file-scope device array.  It calls no cudaMalloc, cudaFree or allocating
// library routine, takes no handle, does not synchronise and copies nothing to the host: inputs and the six results
// are the caller's device buffers.
//
// The array is shared by every call on a device and keeps the previous call's contents: the records a call uses are
// cleared on the caller's stream before the first kernel.  Calls that overlap on one GPU (two streams, two threads)
// RACE on it -- the caller must order them.  Every GPU that loads the library holds the array until exit.

// one record per CTA of the sums / count kernels: the grid never exceeds NOALLOC_MAX_PARTS CTAs (148 SMs x 4)
constexpr int NOALLOC_MAX_PARTS = 148 * 4;
constexpr size_t NOALLOC_SCRATCH_DOUBLES = size_t(NOALLOC_MAX_PARTS) * glp::MET_STRIDE;   // 4736 doubles = 37,888 bytes
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_SCRATCH_DOUBLES];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, once per device, with that device current.
constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

static int noalloc_scratch(int device, double** out) {
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(GLP_ERR_ARG, "device index %d out of range [0, %d)", device, NOALLOC_MAX_DEVICES);
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_noalloc_ptr[device] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);   // the current device's copy of the array
    if (e != cudaSuccess || p == nullptr)
      return fail(GLP_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) failed on device %d: %s", device, cudaGetErrorString(e));
    g_noalloc_ptr[device] = static_cast<double*>(p);
  }
  *out = g_noalloc_ptr[device];
  return GLP_OK;
}

extern "C" int glp_metrics_sums_noalloc(int device, void* stream, int dtype, const void* x_dev, const void* ref_dev, int64_t n,
                                        int l, double max_abs_override, double* out6_dev) {
  if (x_dev == nullptr || out6_dev == nullptr) return fail(GLP_ERR_ARG, "null argument");
  if (dtype != GLP_F32 && dtype != GLP_F64) return fail(GLP_ERR_ARG, "dtype must be GLP_F32 or GLP_F64");
  if (n <= 0 || l <= 0) return fail(GLP_ERR_ARG, "bad shape (n = %lld, l = %d)", static_cast<long long>(n), l);
  if (n > GLP_NOALLOC_MAX_ROWS || l > GLP_NOALLOC_MAX_L)
    return fail(GLP_ERR_ARG, "glp_metrics_sums_noalloc: n = %lld, l = %d exceeds the fixed-scratch limit (n <= %lld, l <= %d); "
                "use glp_metrics_eval", static_cast<long long>(n), l, static_cast<long long>(GLP_NOALLOC_MAX_ROWS), GLP_NOALLOC_MAX_L);
  CU_TRY(cudaSetDevice(device));
  double* part = nullptr;
  GLP_TRY(noalloc_scratch(device, &part));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t total = n * int64_t(l);
  // the grid formula of glp_metrics_eval: same partition of the rows, same summation order, same bits
  const int grid = int(std::max<int64_t>(1, std::min<int64_t>((n + MET_THREADS - 1) / MET_THREADS, NOALLOC_MAX_PARTS)));
  // the array keeps the last call's records: clear the ones this call folds
  CU_TRY(cudaMemsetAsync(part, 0, size_t(grid) * MET_STRIDE * sizeof(double), st));
  if (dtype == GLP_F32) {
    glp_metrics_sums_kernel<float><<<grid, MET_THREADS, 0, st>>>(static_cast<const float*>(x_dev), static_cast<const float*>(ref_dev), n, l, part);
    glp_metrics_count_kernel<float><<<grid, MET_THREADS, 0, st>>>(static_cast<const float*>(x_dev), total, part, grid, max_abs_override);
  } else {
    glp_metrics_sums_kernel<double><<<grid, MET_THREADS, 0, st>>>(static_cast<const double*>(x_dev), static_cast<const double*>(ref_dev), n, l, part);
    glp_metrics_count_kernel<double><<<grid, MET_THREADS, 0, st>>>(static_cast<const double*>(x_dev), total, part, grid, max_abs_override);
  }
  glp_metrics_fold_kernel<<<1, MET_THREADS, 0, st>>>(part, grid, out6_dev);
  CU_TRY(cudaGetLastError());
  return GLP_OK;
}
