// glp_admm_host.inl -- host side of gl_Admm_dual (code/gl_ADMM_dual.py:10-103), included by
// glprox.cu.  Two schedules behind one entry point:
//   m <= 2048   the factor is built and applied by single-CTA kernels (glp_admm.cuh), as in round 1;
//   m >  2048   (or GLPROX_ADMM_BLOCKED=1) blocked set-up in fp64 (glp_dense.cuh): Gram matrix -- on the tensor cores
//               through K5 for fp32 data --, blocked Cholesky, Y = L^-T; every iteration then applies
//               z = Y (Y^T w) with one transposed and one forward stream of Y through K2 / K1 on an inner handle.
extern "C" int glp_admm_default_opts(glp_admm_opts* o) {
  if (o == nullptr) return fail(GLP_ERR_ARG, "null opts");
  o->maxit = 100; o->converge_len = 20; o->thres = 1e-3; o->tau = (1.0 + std::sqrt(5.0)) * 0.5; o->rho = 1e2;
  return GLP_OK;
}


template <typename TA>
static cudaError_t dense_gemm_nt(cudaStream_t st, int64_t M, int64_t N, int64_t K, double alpha, const TA* A, int64_t lda,
                                 const TA* B, int64_t ldb, double beta, double* C, int64_t ldc, bool lower_only, bool add_eye) {
  if (M <= 0 || N <= 0) return cudaSuccess;
  const dim3 grid(unsigned((N + DENSE_TILE - 1) / DENSE_TILE), unsigned((M + DENSE_TILE - 1) / DENSE_TILE));
  dense_gemm_nt_kernel<TA><<<grid, DENSE_THREADS, 0, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lower_only ? 1 : 0,
                                                          add_eye ? 1 : 0);
  return cudaGetLastError();
}

__global__ void dense_transpose_block_kernel(const double* __restrict__ W, int nb, double* __restrict__ Y, int64_t ldy) {
  for (int e = threadIdx.x; e < nb * nb; e += blockDim.x) {
    const int j = e / nb, i = e - j * nb;
    Y[int64_t(j) * ldy + i] = W[i * 128 + j];   // Y[j][i] = W[i][j]
  }
}

// G = I + rho A A^T for fp32 data on the tensor cores: A s = A0 + A1 in fp16 (row-major temporaries), 256 columns of
// the Gram matrix per K5 launch (P = all rows of A, Q = 256 of them), fp64 result scaled by 1 / s^2.
static int admm_gram_tensor(glp_handle* h, double rho, double* G, int64_t ldg) {
  const Geom& g = h->g;
  const float* A = static_cast<const float*>(h->A);
  const int64_t m = g.m, n = g.n, ldk = (n + 7) / 8 * 8;
  unsigned* amax = nullptr;
  __half *a0 = nullptr, *a1 = nullptr;
  double *inv = nullptr, *D = nullptr;
  const int64_t ldd = (m + 7) / 8 * 8;
  auto cleanup = [&]() {
    for (void* q : {static_cast<void*>(amax), static_cast<void*>(a0), static_cast<void*>(a1), static_cast<void*>(inv), static_cast<void*>(D)})
      if (q) cudaFree(q);
  };
#define ADMM_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { cleanup(); return fail(GLP_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_)); } } while (0)
  ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&amax), 16));
  ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&inv), 8));
  ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&a0), size_t(m) * ldk * 2));
  ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&a1), size_t(m) * ldk * 2));
  ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&D), size_t(256) * ldd * 8));
  ADMM_TRY(cudaMemsetAsync(amax, 0, 16, h->stream));
  ADMM_TRY(cudaMemsetAsync(a0, 0, size_t(m) * ldk * 2, h->stream));
  ADMM_TRY(cudaMemsetAsync(a1, 0, size_t(m) * ldk * 2, h->stream));
  const int ag = int(std::min<int64_t>((m * n + 255) / 256, int64_t(h->num_sms) * 16));
  glp_absmax_kernel<<<ag, 256, 0, h->stream>>>(A, m, n, h->lda, amax);
  unsigned bits = 0;
  ADMM_TRY(cudaMemcpyAsync(&bits, amax, sizeof(bits), cudaMemcpyDeviceToHost, h->stream));
  ADMM_TRY(cudaStreamSynchronize(h->stream));
  float mx;
  memcpy(&mx, &bits, sizeof(mx));
  if (!(mx < 3.0e38f)) { cleanup(); return fail(GLP_ERR_ARG, "A holds non-finite values"); }
  const float s = f16_scale_for(mx);
  const double inv_h = 1.0 / (double(s) * double(s));
  ADMM_TRY(cudaMemcpyAsync(inv, &inv_h, 8, cudaMemcpyHostToDevice, h->stream));
  glp_split_f16_2d_kernel<<<ag, 256, 0, h->stream>>>(A, a0, a1, m, n, h->lda, ldk, s);
  ADMM_TRY(cudaGetLastError());
  if (h->pair_ws == nullptr) {
    h->pair_ws_bytes = pair_ws_bytes(h->num_sms, 128);
    int rc0 = alloc_zero(&h->pair_ws, h->pair_ws_bytes, h->stream);
    if (rc0 != GLP_OK) { cleanup(); return rc0; }
  }
  const GemmWs ws{nullptr, 0, h->pair_ws, h->pair_ws_bytes, false};
  for (int64_t jb = 0; jb < m; jb += 256) {
    const int N = int(std::min<int64_t>(256, m - jb));
    int rc1 = gemm_enqueue(h->stream, h->num_sms, 4, m, n, N, reinterpret_cast<const float*>(a0), reinterpret_cast<const float*>(a1),
                           ldk, reinterpret_cast<const float*>(a0 + jb * ldk), reinterpret_cast<const float*>(a1 + jb * ldk), ldk,
                           D, ldd, 4, h->fault, nullptr, nullptr, 0, ws, GemmBf16{nullptr, nullptr, nullptr, nullptr},
                           GemmGate{nullptr, 0}, inv);
    if (rc1 != GLP_OK) { cleanup(); return rc1; }
    const int cg_ = int(std::min<int64_t>((m * N + 255) / 256, int64_t(h->num_sms) * 8));
    dense_from_planar_kernel<<<cg_, 256, 0, h->stream>>>(D, ldd, m, N, jb, rho, G, ldg);
    ADMM_TRY(cudaGetLastError());
    h->n_launches += 2;
  }
  ADMM_TRY(cudaStreamSynchronize(h->stream));   // the temporaries are freed below
#undef ADMM_TRY
  cleanup();
  return GLP_OK;
}

// Blocked set-up: G = I + rho A A^T, L = chol(G), Y = L^-T (upper triangular, row-major [m][ldy]), inner handle over Y.
static int admm_setup_blocked(glp_handle* h, double rho) {
  const Geom& g = h->g;
  const int64_t m = g.m;
  const int64_t ldy = (m + 1) & ~int64_t(1);
  const int nblk = int((m + DENSE_NB - 1) / DENSE_NB);
  double *G = nullptr, *Lm = nullptr, *W = nullptr, *R = nullptr;
  auto cleanup = [&]() { for (double* q : {G, Lm, W, R}) if (q) cudaFree(q); };
#define ADMM_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { cleanup(); return fail(GLP_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_)); } } while (0)
  ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&G), size_t(m) * m * 8));
  ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&Lm), size_t(m) * m * 8));
  ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&W), size_t(nblk) * 128 * 128 * 8));
  ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&R), size_t(m) * 128 * 8));
  if (h->admm_Y == nullptr || h->admm_ldy != ldy) {
    if (h->admm_inner) { glp_handle* in = h->admm_inner; h->admm_inner = nullptr; glp_destroy(in); }
    if (h->admm_Y) cudaFree(h->admm_Y);
    h->admm_Y = nullptr;
    ADMM_TRY(cudaMalloc(reinterpret_cast<void**>(&h->admm_Y), size_t(m) * ldy * 8));
    h->admm_ldy = ldy;
  }
  double* Y = h->admm_Y;
  ADMM_TRY(cudaMemsetAsync(Y, 0, size_t(m) * ldy * 8, h->stream));
  ADMM_TRY(cudaMemsetAsync(Lm, 0, size_t(m) * m * 8, h->stream));
  // ---- Gram matrix
  const char* e_tc = getenv("GLPROX_ADMM_GRAM");   // "cuda": fp64 FMAs also for fp32 data (A/B and a fallback)
  if (h->dtype == GLP_F32 && !(e_tc != nullptr && strcmp(e_tc, "cuda") == 0)) {
    int rc0 = admm_gram_tensor(h, rho, G, m);
    if (rc0 != GLP_OK) { cleanup(); return rc0; }
  } else if (h->dtype == GLP_F32) {
    ADMM_TRY(dense_gemm_nt<float>(h->stream, m, m, g.n, rho, static_cast<const float*>(h->A), h->lda, static_cast<const float*>(h->A),
                                  h->lda, 0.0, G, m, true, true));
    h->n_launches++;
  } else {
    ADMM_TRY(dense_gemm_nt<double>(h->stream, m, m, g.n, rho, static_cast<const double*>(h->A), h->lda, static_cast<const double*>(h->A),
                                   h->lda, 0.0, G, m, true, true));
    h->n_launches++;
  }
  // ---- blocked right-looking Cholesky (lower), W_k = L_kk^-1 kept for the triangular solves with the panels
  ADMM_TRY(cudaFuncSetAttribute(dense_potrf_trtri_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (128 * 129 + 128) * 8));
  for (int k = 0; k < nblk; ++k) {
    const int64_t k0 = int64_t(k) * DENSE_NB, nbk = std::min<int64_t>(DENSE_NB, m - k0), rest = m - k0 - nbk;
    double* Wk = W + size_t(k) * 128 * 128;
    dense_potrf_trtri_kernel<<<1, 512, (128 * 129 + 128) * 8, h->stream>>>(G + k0 * m + k0, m, int(nbk), Lm + k0 * m + k0, m, Wk, h->fault);
    ADMM_TRY(cudaGetLastError());
    h->n_launches++;
    if (rest > 0) {
      // L_ik = G_ik W_k^T
      ADMM_TRY(dense_gemm_nt<double>(h->stream, rest, nbk, nbk, 1.0, G + (k0 + nbk) * m + k0, m, Wk, 128, 0.0, Lm + (k0 + nbk) * m + k0, m, false, false));
      // G_ij -= L_ik L_jk^T on the lower tiles of the trailing block
      ADMM_TRY(dense_gemm_nt<double>(h->stream, rest, rest, nbk, -1.0, Lm + (k0 + nbk) * m + k0, m, Lm + (k0 + nbk) * m + k0, m, 1.0,
                                     G + (k0 + nbk) * m + (k0 + nbk), m, true, false));
      h->n_launches += 2;
    }
  }
  // ---- Y = L^-T, block column k: Y[<k0, k] = -(Y[<k0, <k0] L[k, <k0]^T) W_k^T ; Y[k, k] = W_k^T
  for (int k = 0; k < nblk; ++k) {
    const int64_t k0 = int64_t(k) * DENSE_NB, nbk = std::min<int64_t>(DENSE_NB, m - k0);
    const double* Wk = W + size_t(k) * 128 * 128;
    if (k0 > 0) {
      ADMM_TRY(dense_gemm_nt<double>(h->stream, k0, nbk, k0, -1.0, Y, ldy, Lm + k0 * m, m, 0.0, R, 128, false, false));
      ADMM_TRY(dense_gemm_nt<double>(h->stream, k0, nbk, nbk, 1.0, R, 128, Wk, 128, 0.0, Y + k0, ldy, false, false));
      h->n_launches += 2;
    }
    dense_transpose_block_kernel<<<1, 512, 0, h->stream>>>(Wk, int(nbk), Y + k0 * ldy + k0, ldy);
    ADMM_TRY(cudaGetLastError());
    h->n_launches++;
  }
  ADMM_TRY(cudaStreamSynchronize(h->stream));
#undef ADMM_TRY
  cleanup();
  // ---- inner handle over Y (fp64, m x m): its transposed / forward streams apply L^-1 and L^-T
  if (h->admm_inner == nullptr) {
    if (h->admm_zero_b == nullptr) {
      CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->admm_zero_b), size_t(m) * g.l * 8));
      CU_TRY(cudaMemsetAsync(h->admm_zero_b, 0, size_t(m) * g.l * 8, h->stream));
    }
    GLP_TRY(glp_create(&h->admm_inner, h->device, int(m), int(m), g.l, GLP_F64, Y, ldy, h->admm_zero_b, nullptr, h->stream));
  }
  return GLP_OK;
}

extern "C" int glp_admm_dual(glp_handle* h, const void* x0_dev, double mu, const glp_admm_opts* o, void* x_out_dev,
                             glp_result* res, double* f_hist_host, int hist_cap) {
  if (h == nullptr || x0_dev == nullptr || o == nullptr) return fail(GLP_ERR_ARG, "null argument");
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  if (multi(h)) return fail(GLP_ERR_ARG, "gl_Admm_dual runs on one GPU");
  const Geom& g = h->g;
  if (g.l > ADMM_MAX_L) return fail(GLP_ERR_ARG, "gl_Admm_dual supports l <= %d (got l = %d)", ADMM_MAX_L, g.l);
  if (h->admm_gpart == nullptr) {
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->admm_gpart), size_t(g.ncb) * ADMM_GSTRIDE * 8));
    CU_TRY(cudaMemsetAsync(h->admm_gpart, 0, size_t(g.ncb) * ADMM_GSTRIDE * 8, h->stream));
  }
  const char* e_blk = getenv("GLPROX_ADMM_BLOCKED");
  const bool blocked = g.m > ADMM_MAX_M || (e_blk != nullptr && atoi(e_blk) != 0);
  if (o->maxit < 0 || !(o->rho > 0.0)) return fail(GLP_ERR_ARG, "bad opts (maxit >= 0, rho > 0)");
  CU_TRY(cudaSetDevice(h->device));
  const bool f32 = h->dtype == GLP_F32;
  const int64_t ll = int64_t(g.l) * g.ldn, lm = int64_t(g.l) * g.ldm;
  AdmmConst ac{};
  ac.maxit = o->maxit; ac.converge_len = o->converge_len; ac.hist_cap = std::max(o->maxit, 1) + 8;
  ac.thres = o->thres; ac.tau = o->tau; ac.rho = o->rho; ac.mu = mu;
  if (h->hist_cap < ac.hist_cap) {
    if (h->f_hist) CU_TRY(cudaFree(h->f_hist));
    h->f_hist = nullptr;
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->f_hist), size_t(3) * ac.hist_cap * sizeof(double)));
    h->hist_cap = ac.hist_cap;
  }
  h->f_hist_best = h->f_hist + h->hist_cap;
  h->n_launches = 0;
  h->backend_used = GLP_BACKEND_STREAM;
  CU_TRY(cudaEventRecord(h->ev0, h->stream));
  // ---- one-time factor (inside the call, like :57 of the reference)
  CU_TRY(cudaMemsetAsync(h->fault, 0, sizeof(int32_t), h->stream));
  if (blocked) {
    GLP_TRY(admm_setup_blocked(h, ac.rho));
  } else {
    if (h->admm_L == nullptr) CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->admm_L), size_t(g.m) * g.m * 8));
    CU_TRY(cudaMemsetAsync(h->admm_L, 0, size_t(g.m) * g.m * 8, h->stream));
    const int tb = (g.m + 31) / 32;
    if (f32) glp_admm_gram_kernel<float><<<dim3(tb, tb), dim3(32, 32), 0, h->stream>>>(static_cast<const float*>(h->A), h->lda, g.m, g.n, ac.rho, h->admm_L);
    else glp_admm_gram_kernel<double><<<dim3(tb, tb), dim3(32, 32), 0, h->stream>>>(static_cast<const double*>(h->A), h->lda, g.m, g.n, ac.rho, h->admm_L);
    glp_admm_chol_kernel<<<1, ADMM_THREADS, 0, h->stream>>>(h->admm_L, g.m, h->fault);
    CU_TRY(cudaGetLastError());
    h->n_launches += 2;
  }
  // ---- state: x, u in state[1]; operands [x - rho u | u] in xop (fp32) or state[0] (fp64); A u_prev in Av
  double* x = h->state[1];
  double* u = h->state[1] + ll;
  void* xop = f32 ? h->xop : static_cast<void*>(h->state[0]);
  if (f32) CU_TRY(to_planar(h->stream, static_cast<const float*>(x0_dev), x, g.n, g.l, g.ldn));
  else CU_TRY(to_planar(h->stream, static_cast<const double*>(x0_dev), x, g.n, g.l, g.ldn));
  CU_TRY(cudaMemsetAsync(h->Av, 0, size_t(lm) * 8, h->stream));
  Ctrl c0;
  memset(&c0, 0, sizeof(c0));
  c0.cur = 1;   // forward stream: sel = 1 - cur = 0
  c0.f_best = INFINITY;
  c0.n_fwd = 1;
  CU_TRY(cudaMemcpyAsync(&h->ctrl[0], &c0, sizeof(Ctrl), cudaMemcpyHostToDevice, h->stream));
  CU_TRY(cudaMemcpyAsync(&h->ctrl[1], &c0, sizeof(Ctrl), cudaMemcpyHostToDevice, h->stream));
  h->solving = true;
  const int upd_threads = f32 ? Elem<float>::TILE_COLS : Elem<double>::TILE_COLS;
  auto launch_update = [&](const Ctrl* ctrl, int init) -> cudaError_t {
    AdmmUpdArgs ua{};
    ua.ctrl = ctrl; ua.init = init; ua.x = x; ua.u = u; ua.grad = h->grad; ua.xop = xop;
    ua.gram_partial = h->admm_gpart; ua.scal_partial = h->scal_partial; ua.c = ac;
    if (f32) glp_admm_update_kernel<float><<<g.ncb, upd_threads, 0, h->stream>>>(ua, g);
    else glp_admm_update_kernel<double><<<g.ncb, upd_threads, 0, h->stream>>>(ua, g);
    h->n_launches++;
    return cudaGetLastError();
  };
  int rc = GLP_OK;
  auto body = [&]() -> int {
    CU_TRY(launch_update(nullptr, 1));
    const void* Xb[2] = {xop, xop};
    double* AXb[2] = {h->AX[0], h->AX[0]};
    const double* xc[2] = {x, x};
    const int64_t max_slots = int64_t(o->maxit) + 1;
    int64_t issued = 0, launches = 0;
    bool done = false;
    while (!done && issued < max_slots + 16) {
      for (int s = 0; s < 16; ++s) {
        const Ctrl* cin = &h->ctrl[launches & 1];
        Ctrl* cout = &h->ctrl[(launches + 1) & 1];
        GLP_TRY(enqueue_fwd(h, cin, 0, Xb, AXb, xc, 2 * g.l, 1, false));
        AdmmMidArgs ma{};
        ma.cin = cin; ma.cout = cout; ma.c = ac;
        ma.P = h->AX[0]; ma.b = h->b; ma.Au_prev = h->Av; ma.L = h->admm_L;
        ma.gram_partial = h->admm_gpart; ma.n_scal = g.ncb;
        ma.cnt_partial = h->cnt_partial; ma.n_cnt = g.grid;
        ma.r = h->r; ma.f_hist = h->f_hist; ma.f_hist_best = h->f_hist_best;
        ma.size_total = double(g.n) * double(g.l);
        glp_handle* in = h->admm_inner;
        if (blocked) { ma.w_out = static_cast<double*>(in->r); ma.w_ld = in->g.ldm; }
        const size_t smem = blocked ? 0 : size_t(g.l) * g.m * 8;
        if (smem > 40 * 1024) {   // l > 2 at m near 2048: beyond the default dynamic shared-memory limit
          if (f32) CU_TRY(cudaFuncSetAttribute(glp_admm_mid_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
          else CU_TRY(cudaFuncSetAttribute(glp_admm_mid_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        }
        if (f32) glp_admm_mid_kernel<float><<<1, ADMM_THREADS, smem, h->stream>>>(ma, g);
        else glp_admm_mid_kernel<double><<<1, ADMM_THREADS, smem, h->stream>>>(ma, g);
        CU_TRY(cudaGetLastError());
        h->n_launches++;
        ++launches;
        if (blocked) {
          // z = Y (Y^T w): transposed stream of Y (y = L^-1 w lands in the inner gradient buffer), forward stream of Y
          // over y, then z in the matrix dtype as the operand of the transposed stream of A
          GLP_TRY(enqueue_tpass(in, cout, nullptr, in->grad, false));
          const void* Yx[2] = {in->grad, in->grad};
          double* Yz[2] = {in->AX[0], in->AX[0]};
          GLP_TRY(enqueue_fwd(in, cout, 0, Yx, Yz, nullptr, g.l, 0, false));
          const int64_t tot = int64_t(g.l) * g.ldm;
          const int cgrid = int(std::min<int64_t>((tot + 255) / 256, int64_t(h->num_sms) * 4));
          if (f32) dense_cast_planar_kernel<float><<<cgrid, 256, 0, h->stream>>>(in->AX[0], static_cast<float*>(h->r), tot, cout);
          else dense_cast_planar_kernel<double><<<cgrid, 256, 0, h->stream>>>(in->AX[0], static_cast<double*>(h->r), tot, cout);
          CU_TRY(cudaGetLastError());
          h->n_launches += in->n_launches + 1;
          in->n_launches = 0;
        }
        GLP_TRY(enqueue_tpass(h, cout, nullptr, h->grad, false));
        CU_TRY(launch_update(cout, 0));
      }
      issued += 16;
      CU_TRY(cudaMemcpyAsync(&h->h_ctrl[0], &h->ctrl[launches & 1], sizeof(Ctrl), cudaMemcpyDeviceToHost, h->stream));
      GLP_TRY(check_fault(h, cudaStreamSynchronize(h->stream)));
      done = h->h_ctrl[0].done != 0;
    }
    return GLP_OK;
  };
  rc = body();
  h->solving = false;
  if (rc != GLP_OK) return rc;
  int32_t fault = 0;
  CU_TRY(cudaMemcpyAsync(&fault, h->fault, sizeof(fault), cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  if (res) memset(res, 0, sizeof(*res));
  if (fault != 0) {
    if (res) res->status = GLP_ERR_DEVICE_FAULT;
    return fail(GLP_ERR_DEVICE_FAULT, fault == 3 ? "I + rho A A^T is not positive definite (NaN / inf in A?)"
                                                 : "pipeline watchdog fired inside a streaming kernel");
  }
  const Ctrl c = h->h_ctrl[0];
  if (x_out_dev != nullptr) {
    if (f32) CU_TRY(from_planar(h->stream, x, static_cast<float*>(x_out_dev), g.n, g.l, g.ldn));
    else CU_TRY(from_planar(h->stream, x, static_cast<double*>(x_out_dev), g.n, g.l, g.ldn));
  }
  if (f_hist_host != nullptr && hist_cap > 0 && c.k > 0)
    CU_TRY(cudaMemcpyAsync(f_hist_host, h->f_hist, sizeof(double) * std::min(hist_cap, std::min(c.k, ac.hist_cap)),
                           cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaEventRecord(h->ev1, h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  float ms = 0.f;
  CU_TRY(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
  fill_result(h, c, res);
  if (res) { res->device_ms = ms; res->status = c.done ? GLP_OK : GLP_ERR_STATE; }
  return GLP_OK;
}

