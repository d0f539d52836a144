// glp_metrics_host.inl -- driver metrics on the device and the synthetic-data generator entry point, included by
// glprox.cu.
// ------------------------------------------------------------------------------------------------
// caller-side metrics (code/main.py:30-34,48-50) on the device
// ------------------------------------------------------------------------------------------------
extern "C" int glp_metrics_eval(glp_handle* h, const void* x_dev, const void* ref_dev, double mu,
                                double max_abs_override, glp_metrics* out) {
  if (h == nullptr || x_dev == nullptr || out == nullptr) return fail(GLP_ERR_ARG, "bad argument");
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  CU_TRY(cudaSetDevice(h->device));
  const Geom& g = h->g;
  const int64_t n = g.n, total = int64_t(g.n) * g.l;
  const int grid = int(std::max<int64_t>(1, std::min<int64_t>((n + MET_THREADS - 1) / MET_THREADS, 148 * 4)));
  const bool sharded = h->comm.world > 1;
  // [grid][MET_STRIDE] records, 8 folded values, rr, then (unsharded) the m x l product
  const size_t words = size_t(grid) * MET_STRIDE + 16 + (sharded ? 0 : size_t(g.m) * g.l);
  double* buf = nullptr;
  CU_TRY(cudaMalloc(&buf, words * sizeof(double)));
  double* part = buf;
  double* folded = buf + size_t(grid) * MET_STRIDE;
  double* rr_dev = folded + 8;
  double* ax = folded + 16;
  int rc = GLP_OK;
  cudaError_t ce = cudaMemsetAsync(buf, 0, (size_t(grid) * MET_STRIDE + 16) * sizeof(double), h->stream);
  if (ce == cudaSuccess) {
    if (h->dtype == GLP_F32) {
      glp_metrics_sums_kernel<float><<<grid, MET_THREADS, 0, h->stream>>>(static_cast<const float*>(x_dev), static_cast<const float*>(ref_dev), n, g.l, part);
      glp_metrics_count_kernel<float><<<grid, MET_THREADS, 0, h->stream>>>(static_cast<const float*>(x_dev), total, part, grid, max_abs_override);
    } else {
      glp_metrics_sums_kernel<double><<<grid, MET_THREADS, 0, h->stream>>>(static_cast<const double*>(x_dev), static_cast<const double*>(ref_dev), n, g.l, part);
      glp_metrics_count_kernel<double><<<grid, MET_THREADS, 0, h->stream>>>(static_cast<const double*>(x_dev), total, part, grid, max_abs_override);
    }
    glp_metrics_fold_kernel<<<1, MET_THREADS, 0, h->stream>>>(part, grid, folded);
    ce = cudaGetLastError();
  }
  if (ce == cudaSuccess && !sharded) rc = glp_fwd(h, x_dev, g.l, ax, rr_dev);
  double host[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (ce == cudaSuccess && rc == GLP_OK) ce = cudaMemcpyAsync(host, folded, 9 * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
  if (ce == cudaSuccess && rc == GLP_OK) ce = cudaStreamSynchronize(h->stream);
  cudaFree(buf);
  if (rc != GLP_OK) return rc;
  CU_TRY(ce);
  out->sumsq_x = host[0]; out->sumsq_ref = host[1]; out->sumsq_diff = host[2];
  out->reg = host[3]; out->max_abs = host[4]; out->count = host[5];
  const double nan = std::numeric_limits<double>::quiet_NaN();
  out->rr = sharded ? nan : host[8];
  out->objective = sharded ? nan : 0.5 * host[8] + mu * host[3];
  out->sparsity = total > 0 ? host[5] / double(total) : nan;
  out->err_exact = ref_dev ? std::sqrt(host[2]) / (1.0 + std::sqrt(host[0])) : nan;
  out->err_ref = ref_dev ? std::sqrt(host[2]) / (1.0 + std::sqrt(host[1])) : nan;
  return GLP_OK;
}

// ------------------------------------------------------------------------------------------------
// synthetic data
// ------------------------------------------------------------------------------------------------
extern "C" int glp_randn(void* dst_dev, int dtype, int64_t rows, int64_t cols, int64_t ld, uint64_t seed, int64_t row0,
                         int64_t col0, void* stream) {
  if (dst_dev == nullptr || rows < 0 || cols < 0 || ld < cols) return fail(GLP_ERR_ARG, "bad argument");
  if (dtype != GLP_F32 && dtype != GLP_F64) return fail(GLP_ERR_ARG, "dtype must be GLP_F32 or GLP_F64");
  if (rows * cols == 0) return GLP_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = int(std::min<int64_t>((rows * cols + 255) / 256, 148 * 16));
  if (dtype == GLP_F32) randn_kernel<float><<<grid, 256, 0, st>>>(static_cast<float*>(dst_dev), rows, cols, ld, seed, row0, col0);
  else randn_kernel<double><<<grid, 256, 0, st>>>(static_cast<double*>(dst_dev), rows, cols, ld, seed, row0, col0);
  CU_TRY(cudaGetLastError());
  return GLP_OK;
}
