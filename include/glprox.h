/* glprox.h -- C ABI of libglprox.so: B200 (sm_100a) proximal-gradient group-LASSO solvers.
 *
 *   min_x 0.5*||A x - b||_F^2 + mu * sum_i ||x(i,:)||_2 ,   A: m x n,  x: n x l,  b: m x l
 *
 * The reference (magic3007/convex-optimization) has NO FFI layer: its boundary for this path is the
 * Python callable  [x, iter, out] = gl_ProxGD_primal(x0, A, b, mu, opts)
 *   (code/gl_ProxGD_primal.py:9,146 ; code/gl_FProxGD_primal.py:9,161 ; invoked at code/main.py:114).
 * The drop-in Python modules in convex-optimization_b200/ keep that signature and bind the entry
 * points below with ctypes.  Everything here is plain C: pointers, sizes,
 * PODs; no torch / C++ types cross the boundary, nothing throws across it.
 *
 * Conventions
 *   - Every function returns 0 on success and a negative glp_status on failure;
 *     glp_last_error() returns a thread-local human readable message.
 *   - "matrix dtype" T is GLP_F32 or GLP_F64.  A, b, x0, x_out use T and the reference's
 *     row-major layouts (A[m][lda], b[m][l], x[n][l]).  Vector state inside the library is fp64.
 *   - *_dev pointers are device pointers on the handle's device, *_host are host pointers.
 *   - A handle is single-threaded; all work is enqueued on the stream given at creation.
 */
#ifndef GLPROX_H_
#define GLPROX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GLP_VERSION 100 /* 0.1.0 */

typedef enum glp_status {
  GLP_OK = 0,
  GLP_ERR_ARG = -1,      /* bad argument (shape, dtype, alignment, unknown enum) */
  GLP_ERR_CUDA = -2,     /* a CUDA runtime/driver call failed */
  GLP_ERR_NO_DEVICE = -3,/* no sm_100 device: there is NO CPU fallback */
  GLP_ERR_NCCL = -4,
  GLP_ERR_STATE = -5,    /* call sequence error (e.g. step before begin) */
  GLP_ERR_DEVICE_FAULT = -6, /* a kernel reported an internal fault (pipeline watchdog) */
  GLP_ERR_UNBOUND = -7   /* FProxGD with maxit == 0: the reference raises UnboundLocalError
                            (gl_FProxGD_primal.py:156,161) */
} glp_status;

enum { GLP_F32 = 0, GLP_F64 = 1 };
/* 0/1: the hot path; 2..4: SURVEY.md 8(f) row 1 -- the smoothed / subgradient solvers on the same two streams
 * (gl_FGD_primal.py, gl_GD_primal.py, gl_SGD_primal.py) */
enum { GLP_PROXGD = 0, GLP_FPROXGD = 1, GLP_FGD = 2, GLP_GD = 3, GLP_SGD = 4 };
/* step_type strings of gl_ProxGD_primal.py:78-101 / gl_FProxGD_primal.py:81-106 */
enum { GLP_STEP_LINE_SEARCH = 0, GLP_STEP_FIXED = 1, GLP_STEP_DIMINISHING = 2, GLP_STEP_DIMINISHING2 = 3,
       /* additive opt-in (the reference has no such rule), gl_ProxGD_primal only: the line search of each
        * iteration starts from the Barzilai-Borwein step ||d||^2/||A d||^2 of the displacement just
        * accepted instead of alpha0; both norms come from the line-search scalars already on the device */
       GLP_STEP_BB = 4 };
enum { GLP_BACKEND_AUTO = 0, GLP_BACKEND_STREAM = 1, GLP_BACKEND_PERSISTENT = 2 };

/* POD mirror of the reference's `opts` dict after the merge with default_opts
 * (gl_ProxGD_primal.py:10-21,38-43 ; gl_FProxGD_primal.py:10-21,38-43).  glp_default_opts fills
 * the per-solver defaults. */
typedef struct glp_opts {
  int32_t maxit;                          /* per mu-stage (3 stages) */
  int32_t step_type;                      /* GLP_STEP_* */
  int32_t stable_len_threshold;
  int32_t maxit_line_search_iter;
  double thres;
  double alpha0;
  double ftol;
  double line_search_attenuation_coeffi;
  int32_t backend;                        /* GLP_BACKEND_* (additive key, default AUTO) */
  int32_t poll_interval;                  /* iterations enqueued between host polls (0 = auto) */
  double delta;                           /* smoothing parameter (FGD 1e-6, GD 1e-3; gl_FGD_primal.py:20, gl_GD_primal.py:18) */
} glp_opts;

typedef struct glp_result {
  int32_t status;        /* glp_status of the solve */
  int32_t iters;         /* k: number of recorded objective values (== len(f_hist)) */
  int32_t done;          /* 1 once all three continuation stages have finished */
  int32_t backend_used;  /* GLP_BACKEND_STREAM / GLP_BACKEND_PERSISTENT */
  double fval;           /* out["fval"]: objective (with mu_0) at the returned x */
  double device_ms;      /* CUDA-event time of the solve on this device */
  double sparsity;       /* reference sparsity rule evaluated at the returned x */
  int64_t n_fwd;         /* forward streams of A executed */
  int64_t n_tpass;       /* transposed streams of A executed */
  int64_t n_launches;    /* kernels launched by the library for this solve */
  int32_t stage_end[3];  /* k at the end of each continuation stage */
  int32_t reserved;
} glp_result;

/* Parameters of one transposed-pass epilogue (K2), used by the step-level entry point. */
typedef struct glp_step_params {
  int32_t algo;     /* GLP_PROXGD / GLP_FPROXGD */
  int32_t reserved;
  double theta;     /* FProxGD extrapolation weight 2/(inner+1); ignored for ProxGD */
  double t;         /* trial step */
  double mu;        /* stage mu */
  double thres;
} glp_step_params;

typedef struct glp_handle glp_handle;

/* ---- library ---------------------------------------------------------------------------------*/
int glp_version(void);
const char* glp_last_error(void);
/* Number of visible sm_100 devices (0 when there is none; negative on CUDA error). */
int glp_device_count(void);
int glp_default_opts(int algo, glp_opts* out);

/* ---- handle: one per (device, problem).  A and b stay resident in HBM --------------------------
 * Replaces: the closure over (A, b) inside the reference solvers (gl_ProxGD_primal.py:24-27,60-61).
 * A_dev: m x n_local, row-major, leading dimension lda (elements); lda*sizeof(T) and A_dev must be
 * 16-byte aligned (TMA).  b_dev: m x l row-major, dtype T.  In a multi-GPU job each rank holds a
 * column block A[:, J_g] and passes the ncclComm_t it created with glp_nccl_* (or NULL).
 * stream: cudaStream_t (NULL = legacy default stream). */
int glp_create(glp_handle** out, int device, int m, int n_local, int l, int dtype,
               const void* A_dev, int64_t lda, const void* b_dev, void* nccl_comm, void* stream);
int glp_destroy(glp_handle* h);

/* ---- K1: forward skinny product with fused residual epilogue ---------------------------------
 * Replaces `A @ x` / `0.5*np.sum((A @ x - b)**2)` (gl_ProxGD_primal.py:25,61 ; gl_FProxGD_primal.py:25,63).
 * X_dev: n_local x ncols row-major dtype T.  AX_dev: m x ncols row-major fp64 (this rank's partial
 * product in a multi-GPU job).  rr_dev (nullable): one fp64 = sum((AX[:, :l] - b)^2) over the
 * first min(ncols, l) columns. */
int glp_fwd(glp_handle* h, const void* X_dev, int ncols, double* AX_dev, double* rr_dev);

/* ---- K2: transposed product, optionally with the fused prox epilogue ---------------------------
 * Replaces `A.T @ r` (gl_ProxGD_primal.py:129 ; gl_FProxGD_primal.py:66) and, fused, prox_th /
 * extrapolation / line-search inner products (gl_ProxGD_primal.py:65-74 ; gl_FProxGD_primal.py:71-77,
 * 96-99,139,144-145).  R_dev: m x l fp64 row-major.  G_dev: n_local x l fp64 (gradient, nullable).
 * With p != NULL the epilogue runs on (xt, v) [n_local x l fp64 row-major; v nullable for ProxGD] and
 * writes xn, xnt, vn (nullable) and scalars[4] = {gd, dd, sum_i||xn_i||, max|xn|}. */
int glp_tpass_prox(glp_handle* h, const glp_step_params* p, const double* R_dev, double* G_dev,
                   const double* xt_dev, const double* v_dev, double* xn_dev, double* xnt_dev,
                   double* vn_dev, double* scalars_dev);

/* ---- whole solve ------------------------------------------------------------------------------
 * Replaces gl_ProxGD_primal / gl_FProxGD_primal bodies.  x0_dev, x_out_dev: n_local x l row-major
 * dtype T.  f_hist_host (nullable): receives min(iters, hist_cap) objective values.
 * glp_solve == begin + (step/poll)* + end. */
int glp_solve(glp_handle* h, int algo, const void* x0_dev, double mu0, const glp_opts* opts,
              void* x_out_dev, glp_result* res, double* f_hist_host, int hist_cap);

int glp_solve_begin(glp_handle* h, int algo, const void* x0_dev, double mu0, const glp_opts* opts);
/* Enqueue `slots` decide/transposed/forward rounds (one iteration each unless a line-search trial
 * is rejected); no host synchronisation. */
int glp_solve_step(glp_handle* h, int slots);
/* Synchronise and report progress (res->done, res->iters, res->fval so far). */
int glp_solve_poll(glp_handle* h, glp_result* res);
int glp_solve_end(glp_handle* h, void* x_out_dev, glp_result* res, double* f_hist_host, int hist_cap);

/* ---- K3 step level: the control state machine of the solvers on the device -----------------------------
 * Replaces one pass through the scalar control flow of gl_ProxGD_primal.py:53,76,103-132 /
 * gl_FProxGD_primal.py:55,68-69,79,108-147: sufficient-decrease test of the pending trial, step shrink,
 * commit, f_hist / f_hist_best / sparsity bookkeeping, the two-ratio stop rule, the mu-continuation stage
 * switch and the step rule of the next iteration.  The PODs mirror the library's internal control blocks. */
typedef struct glp_solve_const {
  int32_t algo, step_type, maxit, stable_cap, ls_cap, hist_cap;
  double thres, alpha0, ftol, atten, mu0, delta;
} glp_solve_const;
typedef struct glp_ctrl {
  int32_t done, stage, inner, k, stable, trial, cur, retry, tested, first_commit, ran_any, fault;
  int32_t stage_end[3];
  int32_t gy_sel;
  double t, t_k, theta, mu, f_cur, s_cur, f_prev, s_prev, f_best, g_base;
  int64_t n_fwd, n_tpass;
} glp_ctrl;
typedef struct glp_reduced { double rr, gd, dd, reg, cnt, gbase, sx, sy; } glp_reduced;
typedef struct glp_decide_out {
  int32_t action;       /* 0 none, 1 retry (trial rejected), 2 iterate (candidate committed), 3 finish */
  int32_t reset_v, first_commit, new_cur;
  double theta_old;
} glp_decide_out;
/* Applies the device build of the state machine to `count` independent (ctrl_in, reduced) pairs (host arrays in,
 * host arrays out).  hist3_host (nullable): [count][3] = the first (f_hist, f_hist_best, sparsity_hist) entry the
 * step recorded, NaN where it recorded none.  size_total = n_total * l (sparsity denominator). */
int glp_decide(int device, const glp_solve_const* sc, const glp_ctrl* ctrl_in_host, const glp_reduced* red_host,
               double size_total, glp_ctrl* ctrl_out_host, glp_decide_out* act_host, double* hist3_host, int count);
/* sparsity_hist of the last finished solve (gl_ProxGD_primal.py:116 / gl_FProxGD_primal.py:119): copies
 * min(cap, iters) values, returns how many (>= 0) or a negative glp_status. */
int glp_solve_sparsity_hist(glp_handle* h, double* out_host, int cap);

/* ---- measurement: CUDA-event timing of every streaming launch on the handle's stream ------------
 * With profiling on, each forward / transposed launch of a solve is bracketed by a pair of events;
 * glp_get_profile (after glp_solve/glp_solve_end) aggregates the launches that actually streamed A
 * (duration >= half of the longest of their kind; no-op and epilogue-only launches are left out). */
typedef struct glp_profile {
  int64_t fwd_launches;   /* forward launches that streamed A */
  int64_t tpass_launches; /* transposed launches that streamed A */
  double fwd_ms_sum, fwd_ms_min;
  double tpass_ms_sum, tpass_ms_min;
  double bytes_per_launch; /* algorithmic bytes of one stream of A on this device: m*n_local*sizeof(T) */
} glp_profile;
int glp_set_profiling(glp_handle* h, int on);
int glp_get_profile(glp_handle* h, glp_profile* out);

/* Test/debug: copy an internal fp64 vector to the host in the reference's row-major layout.
 * which: 0 x, 1 xt, 2 v (n_local x l);  3 Ax, 4 Axt, 5 Av (m x l);  7 grad (n_local x l);
 *        8 xn, 9 xnt, 10 vn (candidates);  11 Axn, 12 Axnt. */
int glp_debug_read(glp_handle* h, int which, double* out_host);

/* ---- one-shot convenience with HOST buffers (what the NumPy drop-in path calls) ---------------
 * Uploads A (m x n row-major, leading dimension lda), b, x0 from host memory, solves on `device`,
 * downloads x.  host<->device copies are part of the call. */
int glp_solve_host(int device, int algo, int dtype, int m, int n, int l, const void* A_host,
                   int64_t lda, const void* b_host, const void* x0_host, double mu0,
                   const glp_opts* opts, void* x_out_host, glp_result* res, double* f_hist_host,
                   int hist_cap);

/* ---- multi-GPU: column-block shards of A, exchange through peer mailboxes over NVLink -------------
 * The reference is single-process; this is what sharding its products needs:
 * rank g holds A[:, J_g] (glp_create with n_local = |J_g|) and the row groups x[J_g, :]; the prox is
 * rank-local; per forward stream every rank pushes its m x 2l partial product and five scalars into
 * every peer's mailbox from inside the streaming kernel, and each rank sums the W contributions in
 * rank order (bit-identical on all ranks).
 *   1. glp_comm_export on every rank: allocates this rank's mailbox, returns its 64-byte CUDA IPC
 *      handle (and/or the raw device pointer, for ranks that live in one process);
 *   2. the host exchanges the handles (torch.distributed.all_gather in the Python host);
 *   3. glp_comm_attach(rank, world, n_total, peers, kind): kind 0 = world x 64-byte IPC handles,
 *      kind 1 = world device pointers.  n_total = sum of n_local (sparsity denominator).
 * After that glp_solve* on every rank (same algo/opts/mu0, SPMD) runs the sharded solve; x0 / x_out
 * are the local row groups.  glp_fwd / glp_tpass_prox stay rank-local (partial products). */
int glp_comm_export(glp_handle* h, int world, void* ipc_handle64_out, void** mailbox_dev_out);
int glp_comm_attach(glp_handle* h, int rank, int world, int64_t n_total, const void* peers, int kind);
/* G logical ranks driven by one host thread (one process, G devices -- or G shards on ONE device and
 * one stream, which is how the exchange is tested on a single GPU).  Arrays have G entries; f_hist
 * is rank 0's (identical on all ranks). */
int glp_group_solve(glp_handle** hs, int G, int algo, const void* const* x0_dev, double mu0,
                    const glp_opts* opts, void* const* x_out_dev, glp_result* res, double* f_hist_host,
                    int hist_cap);

/* ---- NCCL plumbing (baseline exchange for A/B measurements; rendezvous of the id is the host's job) */
int glp_nccl_unique_id(void* id128 /* 128 bytes out */);
int glp_nccl_comm_create(void** comm_out, const void* id128, int rank, int world, int device);
int glp_nccl_comm_destroy(void* comm);
/* Baseline exchange for A/B measurements: shard sums with two ncclAllReduce calls per forward stream on the
 * solver's stream instead of the peer-mailbox push (l <= 8 streaming path only; bit-identical on all ranks like
 * the product path, slower: see DESIGN.md section 5). */
int glp_comm_attach_nccl(glp_handle* h, void* nccl_comm, int rank, int world, int64_t n_total);

/* ---- K5: tensor-core products of the wide multi-RHS mode (tcgen05 / TMEM, kind::tf32) -------------
 * D[N][ldd] (fp64, planar) = P[M x K] . Q[N x K]^T with P row-major (pitch ldp) and Q planar (pitch ldq),
 * both fp32 device arrays, N <= 256.  Batched form of `A @ x` / `A.T @ r` (gl_*_primal.py:25,61,129 / :63,66)
 * for l >= 16 right-hand sides.  passes = 1: plain TF32 operands; passes = 3: error-compensated
 * Ph.Qh + Ph.Ql + Pl.Qh with the hi/lo splits of glp_tf32_split (fp32-class accuracy); passes = 2: the same
 * two corrections computed on bf16 copies of hi / lo (kind::f16 at twice the TF32 rate; the corrections are
 * 2^-11 of the result, so their 8-bit mantissas cost ~1e-7 of scale); passes = 4: P_hi and Q_hi are the fp32
 * operands themselves (the lo arguments are ignored): each is scaled by a power of two taken from its largest
 * magnitude and held as a pair of fp16 numbers v0 + v1 (the same 11-bit significands as TF32 hi/lo at half the
 * bytes), P0.Q0 + P0.Q1 + P1.Q0 run on kind::f16 at twice the TF32 rate and the epilogue undoes the scales.
 * k_chunk_blocks < 0: |k_chunk_blocks| with the split-K schedule (partial products summed in split order).
 * k_chunk_blocks > 0: flush the fp32 accumulator to fp64 every k_chunk_blocks*32 elements of K. */
int glp_gemm_tf32(int device, void* stream, int passes, int64_t M, int64_t K, int N, const void* P_hi,
                  const void* P_lo, int64_t ldp, const void* Q_hi, const void* Q_lo, int64_t ldq,
                  double* D_dev, int64_t ldd, int k_chunk_blocks);
/* Wide multi-RHS mode of the solvers (fp32 data, 16 <= l <= 128, one GPU or column shards over several): both products of an
 * iteration run on glp_gemm_tf32 with A and A^T resident as fp16 pairs (2 x |A| of HBM) or TF32 hi/lo splits (4 x |A|).
 * mode 0 (default): on, fp16 pairs (passes = 4 above; measured 2x closer to the fp64 oracle than 3xTF32 and 1.55x
 *   faster at m=16384 n=65536 l=128); falls back to 3xTF32 when A holds non-finite entries, which cannot be scaled;
 * 1: off (CUDA-core streams);  2: on, plain TF32;  3: on, TF32 main term + bf16 correction products (passes = 2);
 * 4: on, fp16 pairs, an error instead of the fallback;  5: on, error-compensated 3xTF32 (passes = 3). */
int glp_set_tensor_mode(glp_handle* h, int mode);
/* hi = src rounded to TF32 (10 mantissa bits), lo = src - hi; transpose != 0 writes the transposes
 * (hi[j][i], lo[j][i]) -- the resident A^T of the transposed pass. */
int glp_tf32_split(void* stream, const void* src, int64_t rows, int64_t cols, int64_t ld_src, void* hi,
                   void* lo, int64_t ld_dst, int transpose);

/* ---- gl_Admm_dual (code/gl_ADMM_dual.py:10-103) ------------------------------------
 * ADMM on the dual problem, no mu-continuation.  Option semantics of default_opts (:11-17): maxit 100,
 * thres 1e-3 (residual test), tau (1+sqrt(5))/2, rho 1e2, converge_len 20.  Replaces the whole function body:
 * the Cholesky factor of I + rho A A^T (:57) is built once on the device, every iteration is one forward and
 * one transposed stream of A plus two triangular solves; objective history, residual norms and the
 * converge_len stop rule (:87-99) are evaluated on the device.  Scope: l <= 8, one GPU; m > 2048 takes the blocked fp64 Cholesky and applies the factor as two streams of L^-T.
 * res->iters = k, res->fval = out["fval"]; f_hist_host receives min(k, hist_cap) values. */
typedef struct glp_admm_opts {
  int32_t maxit;
  int32_t converge_len;
  double thres;
  double tau;
  double rho;
} glp_admm_opts;
int glp_admm_default_opts(glp_admm_opts* out);
int glp_admm_dual(glp_handle* h, const void* x0_dev, double mu, const glp_admm_opts* opts, void* x_out_dev,
                  glp_result* res, double* f_hist_host, int hist_cap);

/* ---- gl_Admm_primal (code/gl_ADMM_primal.py:10-117; registered in code/main.py:181) --------------
 * Linearised ADMM on  min 0.5||A y - b||^2 + mu sum ||x_i||  s.t. x = y.  The reference factors the n x n matrix
 * rho I + A^T A (:62) and solves with it every iteration (:78).  Here the solve goes through the m x m matrix
 * G = I + A A^T / rho:  (rho I + A^T A)^-1 v = (v - A^T G^-1 A v / rho) / rho, so that an iteration is one forward
 * stream of A over [x | v], G^-1 as two streams of the inverse Cholesky factor (the blocked set-up of glp_admm_dual),
 * one transposed stream and a row-local prox / multiplier kernel (csrc/glp_admm_primal.cuh).  Objective history,
 * residual 2-norms and the converge_len stop rule (:98-107) on the device.  converge_thres (:18) is read by the
 * reference but never used; it has no field here.  Scope: fp64 handles (the identity cancels ~cond(rho I + A^T A) of
 * the leading digits), l <= 8, one GPU.  res / f_hist_host as for glp_admm_dual. */
typedef struct glp_admm_primal_opts {
  int32_t maxit;
  int32_t converge_len;
  int32_t step_type;        /* GLP_STEP_FIXED (eta_0), GLP_STEP_DIMINISHING (eta_0 / sqrt k), GLP_STEP_DIMINISHING2 (eta_0 / k) :66-72 */
  int32_t pad;
  double thres;
  double tau;
  double rho;
  double eta_0;
} glp_admm_primal_opts;
int glp_admm_primal_default_opts(glp_admm_primal_opts* out);
int glp_admm_primal(glp_handle* h, const void* x0_dev, double mu, const glp_admm_primal_opts* opts, void* x_out_dev,
                    glp_result* res, double* f_hist_host, int hist_cap);

/* ---- gl_Alm_dual (code/gl_ALM_dual.py:66-158; registered in code/main.py:179) ---------------------
 * Augmented Lagrangian method on the dual: every outer iteration solves for u by 500 accelerated projected-gradient
 * steps (solve_sub_problem, :10-63), then z, the multiplier step and the shared stop rule (:123-147).  The reference
 * forms the n x n matrices Q and G of the inner problem explicitly (:33-40); here Q = rho (I - rho A^T T A) and
 * J = rho (A^T T (A x - b) - x / rho), T = (I + rho A A^T)^-1, are applied as streams (csrc/glp_alm.cuh), so one inner
 * step is one forward stream, T as two streams of the inverse Cholesky factor, one transposed stream and a row-local
 * projection kernel.  Options and defaults are glp_admm_opts / glp_admm_default_opts (the reference's two solvers
 * share them, gl_ALM_dual.py:67-73 == gl_ADMM_dual.py:11-17).  Scope: fp64 handles, l <= 8, one GPU. */
int glp_alm_dual(glp_handle* h, const void* x0_dev, double mu, const glp_admm_opts* opts, void* x_out_dev,
                 glp_result* res, double* f_hist_host, int hist_cap);

/* ---- caller-side metrics of the reference driver, evaluated on the device ----------------------
 * Replaces, for an iterate that stays in HBM, the host lambdas of code/main.py:48-50 and obj_func
 * (code/main.py:30-34); these are the columns of the per-solver log line / Markdown row
 * (code/main.py:113-125, doc/report.md:439-447):
 *   sparsity     = #{ |x| > 1e-6 max|x| } / (n l)                      main.py:50
 *   err_exact    = ||x - ref||_F / (1 + ||x||_F)                        main.py:49  (ref = u)
 *   err_ref      = ||ref - x||_F / (1 + ||ref||_F)                      main.py:48  errfun(ref, x)
 *   objective    = 0.5 ||A x - b||_F^2 + mu sum_i ||x(i,:)||_2          main.py:30-34
 * x_dev / ref_dev: n_local x l row-major arrays of the handle's dtype (ref_dev may be NULL).
 * The raw sums are returned too so that column shards can be combined across ranks (sum / max);
 * for a shard, pass the cross-rank max|x| as max_abs_override (>= 0) to get the global count, and
 * leave the objective to the solve result (it is NaN here when the handle is attached to peers). */
typedef struct glp_metrics {
  double sumsq_x;    /* ||x||_F^2 of this shard */
  double sumsq_ref;  /* ||ref||_F^2 */
  double sumsq_diff; /* ||x - ref||_F^2 */
  double reg;        /* sum_i ||x(i,:)||_2 */
  double max_abs;    /* max |x| of this shard */
  double count;      /* #{ |x| > 1e-6 * max } (max = max_abs_override if >= 0) */
  double rr;         /* ||A x - b||_F^2 (NaN for a shard) */
  double objective, sparsity, err_exact, err_ref; /* finished values for an unsharded handle */
} glp_metrics;
int glp_metrics_eval(glp_handle* h, const void* x_dev, const void* ref_dev, double mu,
                     double max_abs_override, glp_metrics* out);

/* ---- allocation-free variant of the metric reductions -------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  glp_metrics_eval
 * allocates and frees its per-CTA records on every call; this entry point runs the same three reduction kernels with
 * the records in a file-scope `static __device__` array of the library (37,888 bytes; every GPU that loads the
 * library holds it until exit).  It calls no cudaMalloc / cudaFree / allocating library routine, takes no handle,
 * does not synchronise and copies nothing to the host.
 *   x_dev / ref_dev : n x l row-major arrays of `dtype` on `device` (ref_dev may be NULL: sumsq_ref = sumsq_diff = 0)
 *   out6_dev        : 6 doubles on `device`, written on `stream`:
 *                     { sumsq_x, sumsq_ref, sumsq_diff, reg = sum_i ||x(i,:)||_2, max|x|, #{|x| > 1e-6 max} }
 *                     -- bit-identical to the same fields of glp_metrics_eval (same grid, same summation order);
 *                     sparsity = out[5] / (n l), err_exact = sqrt(out[2]) / (1 + sqrt(out[0])) (code/main.py:48-50)
 *   max_abs_override: >= 0 to count against a cross-rank max, < 0 to use this array's own max
 * Limits: n <= GLP_NOALLOC_MAX_ROWS, l <= GLP_NOALLOC_MAX_L; larger inputs are refused with GLP_ERR_ARG (use
 * glp_metrics_eval).  All calls on a device share the one array, which keeps the previous call's contents (the
 * records a call uses are cleared first, on `stream`): calls that overlap on the same GPU -- two streams, two
 * threads -- RACE and must be ordered by the caller. */
#define GLP_NOALLOC_MAX_ROWS ((int64_t)1 << 24)   /* 16,777,216 row groups */
#define GLP_NOALLOC_MAX_L 128
int glp_metrics_sums_noalloc(int device, void* stream, int dtype, const void* x_dev, const void* ref_dev, int64_t n,
                             int l, double max_abs_override, double* out6_dev);

/* ---- synthetic data: counter-based N(0,1) generator (bench / large fixtures) --------------------
 * Fills dst_dev[rows][ld] (dtype T) with value(seed, global_row, global_col) so that any
 * row/column block of the same logical matrix can be generated independently on any rank. */
int glp_randn(void* dst_dev, int dtype, int64_t rows, int64_t cols, int64_t ld, uint64_t seed,
              int64_t row0, int64_t col0, void* stream);

/* b = A u of the same recipe (code/main.py:45): fp64 accumulation over the non-zero row groups of u in ascending
 * column order (nz_dev: their indices), rounded multiply then rounded add -- bit-identical to the CPU restatement
 * (oracle/glprox_gen.c), so that a CPU arm that never loads this library works on the very same right-hand side.
 * u_dev [n_local][l] and y_dev [m][l] are fp64 row-major; a column shard yields its partial product. */
int glp_rhs_exact(const void* A_dev, int dtype, int64_t m, int64_t lda, const double* u_dev, int l,
                  const int64_t* nz_dev, int64_t nnz, double* y_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GLPROX_H_ */
