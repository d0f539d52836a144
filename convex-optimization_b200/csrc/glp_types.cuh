// glp_types.cuh -- PODs shared by the host-side scheduler (solver.cu) and the kernels.
#pragma once
#include <stdint.h>

namespace glp {

// Device-resident control block of one solve.  It is double buffered: decide launch j reads
// ctrl[j & 1] and writes ctrl[(j + 1) & 1]; the transposed / forward launches that follow read the
// latter.  The host never has to know whether a line-search trial was accepted.
struct Ctrl {
  int32_t done;          // all three continuation stages finished: later launches are no-ops
  int32_t stage;         // 0..2  (mu = {100, 10, 1} * mu0)
  int32_t inner;         // inner_iter of the reference (post-increment value)
  int32_t k;             // recorded objective values so far
  int32_t stable;        // stable_len
  int32_t trial;         // failed sufficient-decrease tests of the current iteration
  int32_t cur;           // index (0/1) of the committed state buffers; candidates live in 1-cur
  int32_t retry;         // 1: transposed launch re-runs only the prox epilogue with the new t
  int32_t tested;        // pending candidate carries a sufficient-decrease test
  int32_t first_commit;  // pending candidate is x0
  int32_t ran_any;       // at least one iteration ran (FProxGD maxit==0 -> UnboundLocalError)
  int32_t fault;
  int32_t stage_end[3];
  int32_t gy_sel;        // which of the two base-residual partial buffers is current
  double t;              // trial step
  double t_k;            // FProxGD: last accepted step (carried within a stage)
  double theta;          // FProxGD: 2/(inner+1) of the running iteration
  double mu;             // stage mu
  double f_cur;          // objective (mu0) at the committed x
  double s_cur;          // sparsity at the committed x
  double f_prev, s_prev; // last recorded pair (stop rule)
  double f_best;
  double g_base;         // 0.5*||A base - b||^2 at the extrapolated / thresholded base point
  int64_t n_fwd, n_tpass;
};

// Per-column-block scalar record written by the prox epilogue: {gd, dd, reg, max|xn|, sx, sy, -, -}
// (sx / sy: smoothed regulariser at the candidate / base point, FGD only).
constexpr int SCAL_STRIDE = 8;
enum { ALGO_PROXGD = 0, ALGO_FPROXGD = 1, ALGO_FGD = 2, ALGO_GD = 3, ALGO_SGD = 4 };
enum { STEP_LINE_SEARCH = 0, STEP_FIXED = 1, STEP_DIMINISHING = 2, STEP_DIMINISHING2 = 3, STEP_BB = 4 };

struct SolveConst {
  int32_t algo;          // ALGO_*
  int32_t step_type;
  int32_t maxit;
  int32_t stable_cap;
  int32_t ls_cap;
  int32_t hist_cap;
  double thres, alpha0, ftol, atten, mu0;
  double delta;          // smoothing parameter of FGD / GD (gl_FGD_primal.py:20, gl_GD_primal.py:18)
};

// Geometry of one problem shard on one device.
struct Geom {
  int32_t m, n, l;       // n = local columns of A == local row groups of x
  int32_t nrb, ncb;      // row blocks of TILE_ROWS, column blocks of TILE_COLS
  int64_t ldm, ldn;      // padded vector lengths: nrb*TILE_ROWS, ncb*TILE_COLS
  int32_t grid;          // CTAs of the streaming kernels (<= #SMs)
  int32_t world;         // ranks sharing the rows of x (1 = single GPU)
  int64_t n_total;       // sum of n over ranks (sparsity denominator)
};

// ---- multi-GPU: peer mailboxes ---------------------------------------------------------------------
// Every rank owns one mailbox in its HBM; all ranks map all mailboxes (CUDA IPC between processes,
// plain device pointers inside one process).  Producers STORE into the consumers' mailboxes over
// NVLink (push model) and then raise a monotonically increasing 64-bit flag; consumers only ever
// read their own HBM.  Layout of one mailbox in 8-byte words:
//   [ 0,  8)  flag_ax[src]   sequence number of src's latest finished forward launch
//   [ 8, 16)  flag_sc[src]   sequence number of src's latest prox-producing launch
//   [16, 24)  flag_ep[src]   solve epoch (barrier at glp_solve_begin)
//   [32, 32 + 2*8*8)         sc[par][src][8] = {max|xn|, gd, dd, reg, cnt, -, -, -}
//   [256, 256 + 2*8*slab)    slab[par][src][ncols_total * ldm]   partial products A[:, J_src] * X[J_src]
constexpr int MAX_WORLD = 8;
constexpr int MAIL_FLAG_AX = 0, MAIL_FLAG_SC = 8, MAIL_FLAG_EP = 16, MAIL_SC = 32, MAIL_SLAB = 256;
struct Comm {
  int32_t world;               // 0 / 1: single GPU, no mailbox traffic
  int32_t rank;
  int64_t slab_words;          // 2*l*ldm
  unsigned long long* mail[MAX_WORLD];
};

// Forward stream  AX[:, q] = A * X[:, q]  for RC operand columns.
struct FwdArgs {
  const void* X[2];      // operand [RC][ldn] (dtype T), selected by sel
  double* AX[2];         // output  [RC][ldm] fp64, selected by sel
  const Ctrl* ctrl;      // nullable: when set, sel = 1 - ctrl->cur and done => no-op
  int32_t sel;           // used when ctrl == nullptr
  int32_t nb_cols;       // first nb_cols operand columns are compared against b[bcol0 + q]
  int32_t bcol0;
  int32_t count_mode;    // 1: aux warp evaluates the sparsity count of the candidate x
  const double* b;       // [l][ldm]
  double* partial;       // [(grid + nrb)][TILE_ROWS][RC]
  unsigned* counter;     // [nrb], zero between launches
  double* rr_partial;    // [nrb]  sum over rows of the block of (AX - b)^2
  const double* scal_partial;  // [ncb][SCAL_STRIDE] from the transposed epilogue (slot 3 = max|xn|)
  const double* xcand[2];      // candidate x [l][ldn] fp64, selected like X
  double* cnt_partial;   // [grid]
  int32_t* fault;
  // multi-GPU (comm.world > 1): the completed row blocks are pushed into every rank's slab instead
  // of AX / rr_partial; the last CTA (and aux warp) to finish raises flag_ax = seq_f everywhere.
  Comm comm;
  unsigned long long seq_f;    // this launch
  unsigned long long seq_t;    // prox-producing launch whose scalars (max|xn|) the count needs
  int32_t par;                 // mailbox double-buffer index of this forward group
  int32_t col0;                // first column of this chunk inside the 2l-column slab
  unsigned* fin;               // local finish counter, zero between launches
  int32_t push_fence_per_block;  // 1: system fence after every pushed row block (round-1 behaviour, A/B only)
};

// Transposed stream  G[:, c] = A^T * R[:, c]  for LC residual columns, optional fused epilogue.
struct TpassArgs {
  const void* R;         // [LC][ldm] (dtype T), zero padded
  double* grad;          // [l][ldn] fp64, column gcol0 + c
  int32_t gcol0;
  int32_t fuse;          // 1: run the prox epilogue in the last-arriving CTA of each column block
  const Ctrl* ctrl;      // nullable
  double* partial;       // [(grid + ncb)][LC][TILE_COLS]
  unsigned* counter;     // [ncb]
  int32_t* fault;
};

// Where a prox-producing launch publishes its block-reduced scalars in a multi-GPU job.
struct ProxPost {
  Comm comm;                   // world <= 1: nothing is posted
  unsigned long long seq_t;
  int32_t par;
  int32_t pad;
  unsigned* fin;               // local finish counter, zero between launches
};

// Prox epilogue operands (fused in the transposed stream or run stand-alone for retries / init).
struct ProxArgs {
  double* state[2];      // each: [3][l][ldn] = {x, xt, v}; committed = state[cur], candidates = state[1-cur]
  void* xop;             // fp32 mode: operand copy [2l][ldn] float {xn | xnt}; fp64 mode: nullptr
  const double* grad;    // [l][ldn]
  double* scal_partial;  // [ncb][SCAL_STRIDE] = {gd, dd, reg, max, sx, sy}
  const Ctrl* ctrl;      // nullable: when null the explicit fields below are used
  int32_t algo;
  int32_t cur;
  int32_t init;          // 1: candidate = x0 already stored in state[1-cur][0]; only threshold + scalars
  int32_t pad;
  double theta, t, mu, thres;
  double delta;
  ProxPost post;
};

}  // namespace glp
