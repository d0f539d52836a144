// Host build of the solver control state machine (csrc/glp_decide.h) and of the tile partition
// algebra (csrc/glp_partition.h) for CPU unit tests.  g++ -O1 -ffp-contract=off -shared -fPIC.
#include "glp_decide.h"
#include "glp_partition.h"

extern "C" {
int shim_sizeof_ctrl() { return int(sizeof(glp::Ctrl)); }
void shim_ctrl_init(glp::Ctrl* c, const glp::SolveConst* sc) { glp::ctrl_init(*c, *sc); }
void shim_decide(const glp::SolveConst* sc, const glp::Ctrl* in, const glp::Reduced* red, double size_total,
                 glp::Ctrl* out, glp::DecideOut* act, double* f_hist, double* f_hist_best) {
  glp::decide_step(*sc, *in, *red, size_total, *out, *act, f_hist, f_hist_best);
}
long long shim_tile_lo(long long c, long long G, long long N) { return glp::tile_lo(c, G, N); }
long long shim_tile_owner(long long t, long long G, long long N) { return glp::tile_owner(t, G, N); }
long long shim_contrib_first(long long b, long long per, long long G, long long N) { return glp::contrib_first(b, per, G, N); }
long long shim_contrib_last(long long b, long long per, long long G, long long N) { return glp::contrib_last(b, per, G, N); }
}
