#include <chrono>
// glp_host_io.inl -- the one-shot entry point with HOST buffers (glp_solve_host), included by glprox.cu: pinned or
// bounced upload of A, solve, download of x.
// ------------------------------------------------------------------------------------------------
// one-shot solve with host buffers
// ------------------------------------------------------------------------------------------------
// Host matrix -> HBM.  Page-locked sources (cudaHostAlloc / cudaHostRegister, e.g. torch pinned
// tensors) are DMA'd in place.  Pageable sources go through pinned bounce buffers filled by a few
// host threads, each with its own stream, so that the host-side memcpy (one core moves ~10 GB/s)
// does not cap the PCIe link.
static int upload_rows_bounced(int device, void* dst, size_t dpitch, const char* src, size_t spitch, size_t width,
                               size_t row_begin, size_t row_end, cudaError_t* err_out) {
  const size_t kChunk = size_t(16) << 20;
  cudaError_t e = cudaSetDevice(device);
  cudaStream_t st = nullptr;
  char* pin[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  size_t rows_per = std::max<size_t>(1, kChunk / std::max<size_t>(width, 1));
  rows_per = std::min(rows_per, std::max<size_t>(row_end - row_begin, 1));
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
    e = cudaMallocHost(reinterpret_cast<void**>(&pin[i]), rows_per * width);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming);
  }
  int slot = 0;
  for (size_t r0 = row_begin; r0 < row_end && e == cudaSuccess; r0 += rows_per, slot ^= 1) {
    const size_t nr = std::min(rows_per, row_end - r0);
    e = cudaEventSynchronize(ev[slot]);  // previous DMA out of this bounce buffer finished
    if (e != cudaSuccess) break;
    const char* s = src + r0 * spitch;
    if (spitch == width) memcpy(pin[slot], s, nr * width);
    else for (size_t r = 0; r < nr; ++r) memcpy(pin[slot] + r * width, s + r * spitch, width);
    e = cudaMemcpy2DAsync(static_cast<char*>(dst) + r0 * dpitch, dpitch, pin[slot], width, width, nr,
                          cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaEventRecord(ev[slot], st);
  }
  if (e == cudaSuccess && st) e = cudaStreamSynchronize(st);
  for (int i = 0; i < 2; ++i) {
    if (ev[i]) cudaEventDestroy(ev[i]);
    if (pin[i]) cudaFreeHost(pin[i]);
  }
  if (st) cudaStreamDestroy(st);
  *err_out = e;
  return e == cudaSuccess ? 0 : 1;
}

static int upload_matrix(int device, cudaStream_t st, void* dst, size_t dpitch, const void* src, size_t spitch,
                         size_t width, size_t rows) {
  cudaPointerAttributes at{};
  const bool pinned = cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost;
  cudaGetLastError();
  if (pinned) {
    CU_TRY(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, cudaMemcpyHostToDevice, st));
    CU_TRY(cudaStreamSynchronize(st));
    return GLP_OK;
  }
  CU_TRY(cudaStreamSynchronize(st));  // dst may still be being cleared on st
  const size_t bytes = width * rows;
  int nthr = int(std::min<size_t>(6, std::max<size_t>(1, bytes >> 26)));  // one thread per 64 MiB, up to 6
  nthr = int(std::min<size_t>(size_t(nthr), rows));
  std::vector<std::thread> pool;
  std::vector<cudaError_t> errs(size_t(nthr), cudaSuccess);
  for (int t = 0; t < nthr; ++t) {
    const size_t r0 = rows * size_t(t) / nthr, r1 = rows * size_t(t + 1) / nthr;
    pool.emplace_back([=, &errs]() {
      upload_rows_bounced(device, dst, dpitch, static_cast<const char*>(src), spitch, width, r0, r1, &errs[size_t(t)]);
    });
  }
  for (auto& th : pool) th.join();
  for (cudaError_t e : errs)
    if (e != cudaSuccess) return fail(GLP_ERR_CUDA, "matrix upload failed: %s", cudaGetErrorString(e));
  return GLP_OK;
}

extern "C" int glp_solve_host(int device, int algo, int dtype, int m, int n, int l, const void* A_host, int64_t lda,
                              const void* b_host, const void* x0_host, double mu0, const glp_opts* opts,
                              void* x_out_host, glp_result* res, double* f_hist_host, int hist_cap) {
  if (A_host == nullptr || b_host == nullptr || x0_host == nullptr || x_out_host == nullptr || opts == nullptr)
    return fail(GLP_ERR_ARG, "null argument");
  if (dtype != GLP_F32 && dtype != GLP_F64) return fail(GLP_ERR_ARG, "dtype must be GLP_F32 or GLP_F64");
  if (m <= 0 || n <= 0 || l <= 0 || lda < n) return fail(GLP_ERR_ARG, "bad shape");
  int sms = 0;
  GLP_TRY(check_device(device, &sms));
  CU_TRY(cudaSetDevice(device));
  const size_t es = dtype == GLP_F32 ? 4 : 8;
  const size_t ldd = align_up(size_t(n), 16 / es);  // device pitch in elements, 16-byte aligned rows
  void *dA = nullptr, *db = nullptr, *dx0 = nullptr, *dx = nullptr;
  glp_handle* h = nullptr;
  cudaStream_t st = nullptr;
  int rc = GLP_OK;
  auto cleanup = [&]() {
    if (h) glp_destroy(h);
    if (st) cudaStreamDestroy(st);
    if (dA) cudaFree(dA);
    if (db) cudaFree(db);
    if (dx0) cudaFree(dx0);
    if (dx) cudaFree(dx);
  };
#define HOST_TRY(expr)                                                                        \
  do {                                                                                        \
    cudaError_t e_ = (expr);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      cleanup();                                                                              \
      return fail(GLP_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_));            \
    }                                                                                         \
  } while (0)
  const bool timing = getenv("GLPROX_HOST_TIMING") != nullptr;
  auto now = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t0 = now();
  HOST_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  HOST_TRY(cudaMalloc(&dA, size_t(m) * ldd * es));
  HOST_TRY(cudaMalloc(&db, size_t(m) * l * es));
  HOST_TRY(cudaMalloc(&dx0, size_t(n) * l * es));
  HOST_TRY(cudaMalloc(&dx, size_t(n) * l * es));
  if (ldd != size_t(n)) HOST_TRY(cudaMemsetAsync(dA, 0, size_t(m) * ldd * es, st));
  const double t1 = now();
  rc = upload_matrix(device, st, dA, ldd * es, A_host, size_t(lda) * es, size_t(n) * es, size_t(m));
  if (rc != GLP_OK) { cleanup(); return rc; }
  const double t2 = now();
  HOST_TRY(cudaMemcpyAsync(db, b_host, size_t(m) * l * es, cudaMemcpyHostToDevice, st));
  HOST_TRY(cudaMemcpyAsync(dx0, x0_host, size_t(n) * l * es, cudaMemcpyHostToDevice, st));
  rc = glp_create(&h, device, m, n, l, dtype, dA, int64_t(ldd), db, nullptr, st);
  if (rc == GLP_OK) rc = glp_solve(h, algo, dx0, mu0, opts, dx, res, f_hist_host, hist_cap);
  if (rc == GLP_OK) {
    HOST_TRY(cudaMemcpyAsync(x_out_host, dx, size_t(n) * l * es, cudaMemcpyDeviceToHost, st));
    HOST_TRY(cudaStreamSynchronize(st));
  }
#undef HOST_TRY
  const double t3 = now();
  cleanup();
  if (timing)
    fprintf(stderr, "[glp_solve_host] alloc %.1f ms, upload of A %.1f ms (%.1f GB/s), create + solve + download %.1f ms, free %.1f ms\n",
            t1 - t0, t2 - t1, double(m) * n * es / (t2 - t1) / 1e6, t3 - t2, now() - t3);
  return rc;
}
