// glp_abi_step.inl -- step-level ABI around K3 and the recorded histories (included at the end of glprox.cu).
//
// glp_decide: the DEVICE build of the solver state machine (glp_decide.h: decide_step, the function the decide
// kernel and the persistent kernel call) applied to `count` independent (control block, reduced scalars) pairs, one
// per thread.  Replaces, per call, one pass through the scalar control flow of gl_ProxGD_primal.py:53,76,103-132 /
// gl_FProxGD_primal.py:55,68-69,79,108-147.  tests/test_gpu_decide.py feeds it the states the g++ build sees
// in tests/gpu_emulator.py and compares the outputs bit for bit.

static_assert(sizeof(glp_ctrl) == sizeof(Ctrl) && sizeof(glp_solve_const) == sizeof(SolveConst) &&
              sizeof(glp_reduced) == sizeof(Reduced) && sizeof(glp_decide_out) == sizeof(DecideOut),
              "include/glprox.h mirrors of the control PODs are out of date");

__global__ void glp_decide_probe_kernel(const SolveConst sc, const Ctrl* __restrict__ in, const Reduced* __restrict__ red,
                                        double size_total, Ctrl* __restrict__ out, DecideOut* __restrict__ act,
                                        double* __restrict__ hist, int count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const Ctrl cin = in[i];
  Ctrl cout;
  DecideOut a;
  // decide_step records at index k, k+1, ... of the history arrays (one entry, or one per stage it crosses: a stage
  // that ends on the stop rule is followed by the first record of the next stage in the same step): give every state
  // a private window per array and report the first entry
  const double qnan = __longlong_as_double(0x7ff8000000000000ll);
  double wf[4] = {qnan, qnan, qnan, qnan}, wb[4] = {qnan, qnan, qnan, qnan}, ws[4] = {qnan, qnan, qnan, qnan};
  SolveConst scl = sc;
  const bool fits = cin.k >= 0 && cin.k < sc.hist_cap;
  scl.hist_cap = fits ? min(sc.hist_cap, cin.k + 4) : sc.hist_cap;
  decide_step(scl, cin, red[i], size_total, cout, a, fits ? wf - cin.k : nullptr, fits ? wb - cin.k : nullptr,
              fits ? ws - cin.k : nullptr);
  hist[size_t(3) * i + 0] = wf[0];
  hist[size_t(3) * i + 1] = wb[0];
  hist[size_t(3) * i + 2] = ws[0];
  out[i] = cout;
  act[i] = a;
}

extern "C" int glp_decide(int device, const glp_solve_const* sc, const glp_ctrl* ctrl_in_host, const glp_reduced* red_host,
                          double size_total, glp_ctrl* ctrl_out_host, glp_decide_out* act_host, double* hist3_host,
                          int count) {
  if (sc == nullptr || ctrl_in_host == nullptr || red_host == nullptr || ctrl_out_host == nullptr || act_host == nullptr ||
      count <= 0)
    return fail(GLP_ERR_ARG, "bad argument");
  int sms = 0;
  GLP_TRY(check_device(device, &sms));
  CU_TRY(cudaSetDevice(device));
  const size_t n = size_t(count);
  const size_t b_ctrl = align_up(n * sizeof(Ctrl), 256), b_red = align_up(n * sizeof(Reduced), 256);
  const size_t b_act = align_up(n * sizeof(DecideOut), 256), b_hist = align_up(n * 3 * sizeof(double), 256);
  char* buf = nullptr;
  CU_TRY(cudaMalloc(reinterpret_cast<void**>(&buf), 2 * b_ctrl + b_red + b_act + b_hist));
  Ctrl* d_in = reinterpret_cast<Ctrl*>(buf);
  Ctrl* d_out = reinterpret_cast<Ctrl*>(buf + b_ctrl);
  Reduced* d_red = reinterpret_cast<Reduced*>(buf + 2 * b_ctrl);
  DecideOut* d_act = reinterpret_cast<DecideOut*>(buf + 2 * b_ctrl + b_red);
  double* d_hist = reinterpret_cast<double*>(buf + 2 * b_ctrl + b_red + b_act);
  SolveConst c;
  memcpy(&c, sc, sizeof(c));
  cudaError_t e = cudaMemcpy(d_in, ctrl_in_host, n * sizeof(Ctrl), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_red, red_host, n * sizeof(Reduced), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    glp_decide_probe_kernel<<<(count + 127) / 128, 128>>>(c, d_in, d_red, size_total, d_out, d_act, d_hist, count);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(ctrl_out_host, d_out, n * sizeof(Ctrl), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(act_host, d_act, n * sizeof(DecideOut), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && hist3_host != nullptr) e = cudaMemcpy(hist3_host, d_hist, n * 3 * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(buf);
  if (e != cudaSuccess) return fail(GLP_ERR_CUDA, "glp_decide failed: %s", cudaGetErrorString(e));
  return GLP_OK;
}

// The reference's sparsity_hist (gl_ProxGD_primal.py:116 / gl_FProxGD_primal.py:119) of the last finished solve on
// this handle: sparsity rule evaluated at every recorded iterate.  Feeds the `sparsity=` field of the DEBUG log line.
extern "C" int glp_solve_sparsity_hist(glp_handle* h, double* out_host, int cap) {
  if (h == nullptr || out_host == nullptr || cap < 0) return fail(GLP_ERR_ARG, "bad argument");
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  if (h->s_hist == nullptr || h->last_k <= 0) return 0;
  CU_TRY(cudaSetDevice(h->device));
  const int n = std::min(cap, h->last_k);
  CU_TRY(cudaMemcpyAsync(out_host, h->s_hist, sizeof(double) * size_t(n), cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  return n;
}

// Right-hand side of the synthetic recipe, bit-identical to the CPU restatement (see rhs_exact_kernel).
// u_dev: [n_local][l] fp64 row-major; nz_dev: ascending indices of its non-zero rows; y_dev: [m][l] fp64 row-major.
extern "C" int glp_rhs_exact(const void* A_dev, int dtype, int64_t m, int64_t lda, const double* u_dev, int l,
                             const int64_t* nz_dev, int64_t nnz, double* y_dev, void* stream) {
  if (A_dev == nullptr || u_dev == nullptr || y_dev == nullptr || (nnz > 0 && nz_dev == nullptr) || m < 0 || l < 1 || nnz < 0)
    return fail(GLP_ERR_ARG, "bad argument");
  if (dtype != GLP_F32 && dtype != GLP_F64) return fail(GLP_ERR_ARG, "dtype must be GLP_F32 or GLP_F64");
  if (m == 0) return GLP_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = int((m + 127) / 128);
  if (dtype == GLP_F32) rhs_exact_kernel<float><<<grid, 128, 0, st>>>(static_cast<const float*>(A_dev), m, lda, u_dev, l, nz_dev, nnz, y_dev);
  else rhs_exact_kernel<double><<<grid, 128, 0, st>>>(static_cast<const double*>(A_dev), m, lda, u_dev, l, nz_dev, nnz, y_dev);
  CU_TRY(cudaGetLastError());
  return GLP_OK;
}
