// glp_partition.h -- static "stream-K" partition of a linear tile sequence over the persistent grid.
// Plain C++ (host + device) so that the index algebra is unit-tested on the CPU
// (tests/test_partition.py builds it with g++).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define GLP_HD __host__ __device__ __forceinline__
#else
#define GLP_HD inline
#endif

namespace glp {

// CTA c of G owns tiles [tile_lo(c), tile_lo(c+1)) of N.
GLP_HD int64_t tile_lo(int64_t c, int64_t G, int64_t N) { return (c * N) / G; }

// Owner of tile t: the unique c with tile_lo(c) <= t < tile_lo(c+1).
GLP_HD int64_t tile_owner(int64_t t, int64_t G, int64_t N) { return ((t + 1) * G - 1) / N; }

// The tiles of output block `blk` (a row block in the forward pass, a column block in the
// transposed pass) are [blk*per, (blk+1)*per).  Its contributors are the consecutive CTAs
// first..last; contributor c stores its partial in slot c + blk (unique per (c, blk) because both
// indices are non-decreasing along the tile sequence and at least one increases).
GLP_HD int64_t contrib_first(int64_t blk, int64_t per, int64_t G, int64_t N) {
  return tile_owner(blk * per, G, N);
}
GLP_HD int64_t contrib_last(int64_t blk, int64_t per, int64_t G, int64_t N) {
  return tile_owner((blk + 1) * per - 1, G, N);
}

}  // namespace glp
