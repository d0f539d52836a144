// glp_admm2_host.inl -- host side of gl_Admm_primal (glp_admm_primal.cuh) and gl_Alm_dual (glp_alm.cuh), included by
// glprox.cu after glp_admm_host.inl: both reuse its blocked set-up (G = I + rho' A A^T, L = chol(G), Y = L^-T behind an
// inner m x m fp64 handle) and apply G^-1 as one transposed and one forward stream of Y.

// dst [l][ldm] (fp64 planar) = G^-1 w, w = the inner handle's residual operand (in->r).  No-op once ctrl->done.
static int admm_apply_ginv(glp_handle* h, const Ctrl* ctrl, double* dst) {
  glp_handle* in = h->admm_inner;
  GLP_TRY(enqueue_tpass(in, ctrl, nullptr, in->grad, false));      // y = L^-1 w = Y^T w
  const void* Yx[2] = {in->grad, in->grad};
  double* Yz[2] = {dst, dst};
  GLP_TRY(enqueue_fwd(in, ctrl, 0, Yx, Yz, nullptr, h->g.l, 0, false));   // L^-T y = Y y
  h->n_launches += in->n_launches;
  in->n_launches = 0;
  return GLP_OK;
}

static int admm2_common_checks(glp_handle* h, const void* x0_dev, const char* who) {
  if (h == nullptr || x0_dev == nullptr) return fail(GLP_ERR_ARG, "null argument");
  if (h->solving) return fail(GLP_ERR_STATE, "a solve is in progress on this handle");
  if (multi(h)) return fail(GLP_ERR_ARG, "%s runs on one GPU", who);
  if (h->dtype != GLP_F64) return fail(GLP_ERR_ARG, "%s needs an fp64 handle (the m x m form cancels ~1e5 of the leading digits)", who);
  if (h->g.l > ADMM_MAX_L) return fail(GLP_ERR_ARG, "%s supports l <= %d (got l = %d)", who, ADMM_MAX_L, h->g.l);
  return GLP_OK;
}

static int admm2_workspace(glp_handle* h, size_t n_doubles, size_t m_doubles, size_t rec_doubles) {
  if (h->aux_n_len < n_doubles) {
    if (h->aux_n) cudaFree(h->aux_n);
    h->aux_n = nullptr; h->aux_n_len = 0;
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->aux_n), n_doubles * 8));
    h->aux_n_len = n_doubles;
  }
  if (h->aux_m_len < m_doubles) {
    if (h->aux_m) cudaFree(h->aux_m);
    h->aux_m = nullptr; h->aux_m_len = 0;
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->aux_m), m_doubles * 8));
    h->aux_m_len = m_doubles;
  }
  if (h->aux_rec_len < rec_doubles) {
    if (h->aux_rec) cudaFree(h->aux_rec);
    h->aux_rec = nullptr; h->aux_rec_len = 0;
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->aux_rec), rec_doubles * 8));
    h->aux_rec_len = rec_doubles;
  }
  CU_TRY(cudaMemsetAsync(h->aux_n, 0, h->aux_n_len * 8, h->stream));
  CU_TRY(cudaMemsetAsync(h->aux_m, 0, h->aux_m_len * 8, h->stream));
  CU_TRY(cudaMemsetAsync(h->aux_rec, 0, h->aux_rec_len * 8, h->stream));
  return GLP_OK;
}

static int admm2_hist(glp_handle* h, int cap) {
  if (h->hist_cap < cap) {
    if (h->f_hist) CU_TRY(cudaFree(h->f_hist));
    h->f_hist = nullptr;
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&h->f_hist), size_t(3) * cap * sizeof(double)));
    h->hist_cap = cap;
  }
  h->f_hist_best = h->f_hist + h->hist_cap;
  return GLP_OK;
}

// the common tail: fault word, x, history, timing, result
static int admm2_finish(glp_handle* h, const double* x, void* x_out_dev, glp_result* res, double* f_hist_host, int hist_cap,
                        int dev_hist_cap, const char* not_pd) {
  const Geom& g = h->g;
  int32_t fault = 0;
  CU_TRY(cudaMemcpyAsync(&fault, h->fault, sizeof(fault), cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  if (res) memset(res, 0, sizeof(*res));
  if (fault != 0) {
    if (res) res->status = GLP_ERR_DEVICE_FAULT;
    return fail(GLP_ERR_DEVICE_FAULT, fault == 3 ? not_pd : "pipeline watchdog fired inside a streaming kernel");
  }
  const Ctrl c = h->h_ctrl[0];
  if (x_out_dev != nullptr) CU_TRY(from_planar(h->stream, x, static_cast<double*>(x_out_dev), g.n, g.l, g.ldn));
  if (f_hist_host != nullptr && hist_cap > 0 && c.k > 0)
    CU_TRY(cudaMemcpyAsync(f_hist_host, h->f_hist, sizeof(double) * std::min(hist_cap, std::min(c.k, dev_hist_cap)),
                           cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaEventRecord(h->ev1, h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  float ms = 0.f;
  CU_TRY(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
  fill_result(h, c, res);
  if (res) { res->device_ms = ms; res->status = c.done ? GLP_OK : GLP_ERR_STATE; }
  return GLP_OK;
}

static int admm2_ctrl_init(glp_handle* h) {
  Ctrl c0;
  memset(&c0, 0, sizeof(c0));
  c0.cur = 1;   // forward streams: sel = 1 - cur = 0
  c0.f_best = INFINITY;
  c0.n_fwd = 1;
  CU_TRY(cudaMemcpyAsync(&h->ctrl[0], &c0, sizeof(Ctrl), cudaMemcpyHostToDevice, h->stream));
  CU_TRY(cudaMemcpyAsync(&h->ctrl[1], &c0, sizeof(Ctrl), cudaMemcpyHostToDevice, h->stream));
  return GLP_OK;
}

extern "C" int glp_admm_primal_default_opts(glp_admm_primal_opts* o) {
  if (o == nullptr) return fail(GLP_ERR_ARG, "null argument");
  o->maxit = 100; o->converge_len = 10; o->step_type = GLP_STEP_FIXED; o->pad = 0;
  o->thres = 1e-3; o->tau = (1.0 + sqrt(5.0)) * 0.5; o->rho = 1e-2; o->eta_0 = 100.0;
  return GLP_OK;
}

extern "C" int glp_admm_primal(glp_handle* h, const void* x0_dev, double mu, const glp_admm_primal_opts* o, void* x_out_dev,
                               glp_result* res, double* f_hist_host, int hist_cap) {
  if (o == nullptr) return fail(GLP_ERR_ARG, "null argument");
  GLP_TRY(admm2_common_checks(h, x0_dev, "gl_Admm_primal"));
  if (o->maxit < 0 || !(o->rho > 0.0)) return fail(GLP_ERR_ARG, "bad opts (maxit >= 0, rho > 0)");
  if (o->step_type != GLP_STEP_FIXED && o->step_type != GLP_STEP_DIMINISHING && o->step_type != GLP_STEP_DIMINISHING2)
    return fail(GLP_ERR_ARG, "step_type must be fixed, diminishing or diminishing2");
  const Geom& g = h->g;
  CU_TRY(cudaSetDevice(h->device));
  const int64_t ln = int64_t(g.l) * g.ldn, lm = int64_t(g.l) * g.ldm;
  AdmmPConst ac{};
  ac.maxit = o->maxit; ac.converge_len = o->converge_len; ac.hist_cap = std::max(o->maxit, 1) + 8;
  ac.step_type = o->step_type == GLP_STEP_DIMINISHING ? 1 : (o->step_type == GLP_STEP_DIMINISHING2 ? 2 : 0);
  ac.thres = o->thres; ac.tau = o->tau; ac.rho = o->rho; ac.mu = mu; ac.eta0 = o->eta_0;
  GLP_TRY(admm2_hist(h, ac.hist_cap));
  GLP_TRY(admm2_workspace(h, size_t(ln), 0, size_t(g.ncb) * ADMMP_GSTRIDE));
  h->n_launches = 0;
  h->backend_used = GLP_BACKEND_STREAM;
  CU_TRY(cudaEventRecord(h->ev0, h->stream));
  CU_TRY(cudaMemsetAsync(h->fault, 0, sizeof(int32_t), h->stream));
  // ---- one-time factor (inside the call, like :62 of the reference): G = I + A A^T / rho
  GLP_TRY(admm_setup_blocked(h, 1.0 / ac.rho));
  glp_handle* in = h->admm_inner;
  double* x = h->state[1];
  double* z = h->state[1] + ln;
  double* yprev = h->state[1] + 2 * ln;
  double* atb = h->aux_n;
  double* xop = h->state[0];
  CU_TRY(to_planar(h->stream, static_cast<const double*>(x0_dev), x, g.n, g.l, g.ldn));
  GLP_TRY(admm2_ctrl_init(h));
  // A^T b (:63): b is the residual operand of one transposed stream
  CU_TRY(cudaMemcpyAsync(h->r, h->b, size_t(lm) * 8, cudaMemcpyDeviceToDevice, h->stream));
  GLP_TRY(enqueue_tpass(h, nullptr, nullptr, h->grad, false));
  CU_TRY(cudaMemcpyAsync(atb, h->grad, size_t(ln) * 8, cudaMemcpyDeviceToDevice, h->stream));
  h->solving = true;
  auto launch_update = [&](const Ctrl* ctrl, int init) -> cudaError_t {
    AdmmPUpdArgs ua{};
    ua.ctrl = ctrl; ua.init = init; ua.x = x; ua.z = z; ua.yprev = yprev; ua.atb = atb; ua.grad = h->grad; ua.xop = xop;
    ua.rec = h->aux_rec; ua.scal_partial = h->scal_partial; ua.c = ac;
    glp_admm_primal_update_kernel<<<g.ncb, Elem<double>::TILE_COLS, 0, h->stream>>>(ua, g);
    h->n_launches++;
    return cudaGetLastError();
  };
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
        AdmmPMidArgs ma{};
        ma.cin = cin; ma.cout = cout; ma.c = ac;
        ma.P = h->AX[0]; ma.b = h->b; ma.rec = h->aux_rec; ma.n_rec = g.ncb;
        ma.cnt_partial = h->cnt_partial; ma.n_cnt = g.grid;
        ma.f_hist = h->f_hist; ma.f_hist_best = h->f_hist_best;
        ma.size_total = double(g.n) * double(g.l);
        ma.w_out = static_cast<double*>(in->r); ma.w_ld = in->g.ldm;
        glp_admm_primal_mid_kernel<<<1, ADMM_THREADS, 0, h->stream>>>(ma, g);
        CU_TRY(cudaGetLastError());
        h->n_launches++;
        ++launches;
        GLP_TRY(admm_apply_ginv(h, cout, static_cast<double*>(h->r)));   // q = G^-1 A v, straight into the residual operand
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
  const int rc = body();
  h->solving = false;
  if (rc != GLP_OK) return rc;
  return admm2_finish(h, x, x_out_dev, res, f_hist_host, hist_cap, ac.hist_cap,
                      "I + A A^T / rho is not positive definite (NaN / inf in A?)");
}

// ------------------------------------------------------------------------------------------------
// gl_Alm_dual
// ------------------------------------------------------------------------------------------------
extern "C" int glp_alm_dual(glp_handle* h, const void* x0_dev, double mu, const glp_admm_opts* o, void* x_out_dev,
                            glp_result* res, double* f_hist_host, int hist_cap) {
  if (o == nullptr) return fail(GLP_ERR_ARG, "null argument");
  GLP_TRY(admm2_common_checks(h, x0_dev, "gl_Alm_dual"));
  if (o->maxit < 0 || !(o->rho > 0.0)) return fail(GLP_ERR_ARG, "bad opts (maxit >= 0, rho > 0)");
  const Geom& g = h->g;
  CU_TRY(cudaSetDevice(h->device));
  const int64_t ln = int64_t(g.l) * g.ldn, lm = int64_t(g.l) * g.ldm;
  AlmConst ac{};
  ac.maxit = o->maxit; ac.converge_len = o->converge_len; ac.hist_cap = std::max(o->maxit, 1) + 8;
  ac.thres = o->thres; ac.tau = o->tau; ac.rho = o->rho; ac.mu = mu;
  GLP_TRY(admm2_hist(h, ac.hist_cap));
  GLP_TRY(admm2_workspace(h, size_t(ln), size_t(4) * lm, size_t(g.ncb) * ADMM_GSTRIDE));
  h->n_launches = 0;
  h->backend_used = GLP_BACKEND_STREAM;
  CU_TRY(cudaEventRecord(h->ev0, h->stream));
  CU_TRY(cudaMemsetAsync(h->fault, 0, sizeof(int32_t), h->stream));
  // ---- one-time factor (inside the call, like :114 of the reference)
  GLP_TRY(admm_setup_blocked(h, ac.rho));
  glp_handle* in = h->admm_inner;
  double* x = h->state[1];
  double* uu = h->state[1] + ln;
  double* vv = h->state[1] + 2 * ln;
  double* J = h->aux_n;
  double* xop = h->state[0];
  double* D = h->aux_m;
  double* Au = h->aux_m + lm;
  double* Av = h->aux_m + 2 * lm;
  double* Au_prev = h->aux_m + 3 * lm;
  CU_TRY(to_planar(h->stream, static_cast<const double*>(x0_dev), x, g.n, g.l, g.ldn));
  GLP_TRY(admm2_ctrl_init(h));
  h->solving = true;
  const int m_grid = int(std::max<int64_t>(1, std::min<int64_t>((lm + ALM_THREADS - 1) / ALM_THREADS, int64_t(h->num_sms) * 4)));
  auto launch_n = [&](const Ctrl* ctrl, int mode, int kk) -> cudaError_t {
    AlmNArgs na{};
    na.ctrl = ctrl; na.mode = mode; na.kk = kk; na.x = x; na.u = uu; na.v = vv; na.J = J; na.grad = h->grad; na.xop = xop;
    na.rec = h->aux_rec; na.scal_partial = h->scal_partial; na.c = ac;
    glp_alm_n_kernel<<<g.ncb, Elem<double>::TILE_COLS, 0, h->stream>>>(na, g);
    h->n_launches++;
    return cudaGetLastError();
  };
  auto launch_m = [&](const Ctrl* ctrl, int mode, int kk) -> cudaError_t {
    AlmMArgs ma{};
    ma.ctrl = ctrl; ma.mode = mode; ma.kk = kk; ma.P = h->AX[0]; ma.Au = Au; ma.Av = Av; ma.D = D; ma.rho = ac.rho;
    ma.w_out = static_cast<double*>(in->r); ma.w_ld = in->g.ldm;
    glp_alm_m_kernel<<<m_grid, ALM_THREADS, 0, h->stream>>>(ma, g);
    h->n_launches++;
    return cudaGetLastError();
  };
  const void* Xb[2] = {xop, xop};
  double* AXb[2] = {h->AX[0], h->AX[0]};
  const double* xc[2] = {x, x};
  double* rdst = static_cast<double*>(h->r);
  auto body = [&]() -> int {
    CU_TRY(launch_n(&h->ctrl[0], 3, 0));
    int64_t slots = 0;
    bool done = false;
    while (!done && slots < int64_t(o->maxit) + 2) {
      const Ctrl* cin = &h->ctrl[slots & 1];
      Ctrl* cout = &h->ctrl[(slots + 1) & 1];
      // A x (objective of the outer iteration just finished, D of the next one) + sparsity count
      GLP_TRY(enqueue_fwd(h, cin, 0, Xb, AXb, xc, g.l, 1, false));
      AlmEvalArgs ea{};
      ea.cin = cin; ea.cout = cout; ea.c = ac;
      ea.P = h->AX[0]; ea.b = h->b; ea.Au = Au; ea.Au_prev = Au_prev;
      ea.rec = h->aux_rec; ea.n_rec = g.ncb; ea.cnt_partial = h->cnt_partial; ea.n_cnt = g.grid;
      ea.f_hist = h->f_hist; ea.f_hist_best = h->f_hist_best; ea.size_total = double(g.n) * double(g.l);
      ea.D = D; ea.w_out = static_cast<double*>(in->r); ea.w_ld = in->g.ldm;
      glp_alm_eval_kernel<<<1, ADMM_THREADS, 0, h->stream>>>(ea, g);
      CU_TRY(cudaGetLastError());
      h->n_launches++;
      ++slots;
      // E = T D, J = rho (A^T E - x / rho), u = v = 0
      GLP_TRY(admm_apply_ginv(h, cout, rdst));
      GLP_TRY(enqueue_tpass(h, cout, nullptr, h->grad, false));
      CU_TRY(launch_n(cout, 0, 0));
      // 500 accelerated projected-gradient steps on u (:50-61)
      for (int kk = 1; kk <= ALM_INNER_STEPS; ++kk) {
        if (kk > 1) {
          CU_TRY(launch_m(cout, 0, kk));
          GLP_TRY(admm_apply_ginv(h, cout, rdst));
          GLP_TRY(enqueue_tpass(h, cout, nullptr, h->grad, false));
        }
        CU_TRY(launch_n(cout, 1, kk));
        GLP_TRY(enqueue_fwd(h, cout, 0, Xb, AXb, nullptr, g.l, 0, false));
      }
      // z = T (D - rho A u), x <- x - tau rho (u + A^T z)
      CU_TRY(launch_m(cout, 1, 0));
      GLP_TRY(admm_apply_ginv(h, cout, rdst));
      GLP_TRY(enqueue_tpass(h, cout, nullptr, h->grad, false));
      CU_TRY(launch_n(cout, 2, 0));
      CU_TRY(cudaMemcpyAsync(&h->h_ctrl[0], cout, sizeof(Ctrl), cudaMemcpyDeviceToHost, h->stream));
      GLP_TRY(check_fault(h, cudaStreamSynchronize(h->stream)));
      done = h->h_ctrl[0].done != 0;
    }
    return GLP_OK;
  };
  const int rc = body();
  h->solving = false;
  if (rc != GLP_OK) return rc;
  return admm2_finish(h, x, x_out_dev, res, f_hist_host, hist_cap, ac.hist_cap,
                      "I + rho A A^T is not positive definite (NaN / inf in A?)");
}
