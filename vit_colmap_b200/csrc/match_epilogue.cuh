// Epilogue of the single-pass fused matchers (match_fused.cu, match_fused2.cu): one 32-row x 32-column
// chunk of the fp32 score tile, read from tensor memory and folded into
//   * the thread's running row state (best, second, index) -- COLMAP/feature/sift.cc:745-757, and
//   * the CTA-wide column state in shared memory: a 64-bit key (score bits << 32 | ~row: the larger key
//     is the better score, the lower row on ties) plus a 32-bit second-best per column.
// A score takes part in the column state only if it beats the column's current second-best.  That test is
// two-level: the maximum of each group of 8 scores (needed for the rows anyway) is first compared with the
// smallest threshold of the group's 8 columns (colmin8, refreshed by the epilogue warps at every row tile;
// a stale value is merely less selective), and only groups that pass for some lane compare score by score
// against the per-column thresholds.  Once the thresholds have risen -- or start at the pruning bound --
// most groups stop there.  Two merge paths:
//   sparse : few columns of the chunk are hit -- every hit lane merges its own score with two shared-
//            memory atomics (the usual case once the thresholds have risen);
//   dense  : many columns are hit at once (the first row tiles of a work item, when the thresholds are
//            still low: with S scores seen, a new one enters the top-2 with probability 2/S, and more
//            than half of all hits of a sweep fall into its first 128 rows) -- per hit column the
//            32-row segment's top-2 comes from two warp-wide integer maxima (REDUX) and ONE lane
//            merges it, so the 64-bit compare-and-swap loops do not pile up on one address.
// Both paths produce the same state: best = maximum, lowest row on ties, second = maximum of the rest.
#pragma once
#include <cstdint>

#include "match_kernel.cuh"
#include "vc_ptx.cuh"

namespace vc {

// Refresh of colmin8 for the groups [g_begin, g_end) of 8 columns (lane l takes group g_begin + l).  Called by
// every epilogue warp for its own chunks at each row tile; concurrent writers store lower bounds that
// were valid when read, and the thresholds only rise.
__device__ __forceinline__ void refresh_colmin8(const unsigned int* colsec, unsigned int* colmin8, int g_begin,
                                                int g_end, int lane) {
  const int g = g_begin + lane;
  if (g < g_end) {
    const uint4 a = *reinterpret_cast<const uint4*>(colsec + 8 * g);
    const uint4 b = *reinterpret_cast<const uint4*>(colsec + 8 * g + 4);
    // thresholds are non-negative floats: their bit patterns order like unsigned integers
    colmin8[g] = min(min(min(a.x, a.y), min(a.z, a.w)), min(min(b.x, b.y), min(b.z, b.w)));
  }
  __syncwarp();
}

struct RowTop2 {
  float best, second;
  int idx;
};

// row_tag = 0xFFFFFFFF - row of THIS lane; rows of a warp are consecutive (row = base + lane).
__device__ __forceinline__ void fused_epilogue_chunk(uint32_t taddr, int c, int ncols, bool row_ok,
                                                     unsigned long long row_tag, float thr, int colbase, int lane,
                                                     int dense_min, unsigned long long* colkey, unsigned int* colsec,
                                                     const unsigned int* colmin8, RowTop2& rs) {
  uint32_t vr[32];
  tmem_ld_32x32(taddr, vr);
  tmem_wait_ld();
  float v[32];
#pragma unroll
  for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(vr[e]);
  if (!row_ok) {  // padding rows of the last row tile must not reach the column state
#pragma unroll
    for (int e = 0; e < 32; ++e) v[e] = 0.f;
  }
  if (c * 32 + 32 > ncols) {
#pragma unroll
    for (int e = 0; e < 32; ++e)
      if (c * 32 + e >= ncols) v[e] = 0.f;
  }
  // ---- group maxima: they decide whether the row's running top-2 can change, and pre-filter the columns
  float gm[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    float m = v[8 * g];
#pragma unroll
    for (int e = 1; e < 8; ++e) m = fmaxf(m, v[8 * g + e]);
    gm[g] = m;
  }
  // ---- columns: per-lane mask of scores that beat the column's current second-best
  const uint4 tm = *reinterpret_cast<const uint4*>(colmin8 + c * 4);
  unsigned gpass = (gm[0] > __uint_as_float(tm.x) ? 1u : 0u) | (gm[1] > __uint_as_float(tm.y) ? 2u : 0u) |
                   (gm[2] > __uint_as_float(tm.z) ? 4u : 0u) | (gm[3] > __uint_as_float(tm.w) ? 8u : 0u);
  gpass = __reduce_or_sync(0xffffffffu, gpass);
  const uint4* thr4 = reinterpret_cast<const uint4*>(colsec + c * 32);
  unsigned cm = 0;
#pragma unroll
  for (int g2 = 0; g2 < 4; ++g2) {
    if (gpass & (1u << g2)) {  // warp-uniform
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int g = 2 * g2 + k;
        const uint4 t4 = thr4[g];
        cm |= (v[4 * g + 0] > __uint_as_float(t4.x)) ? (1u << (4 * g + 0)) : 0u;
        cm |= (v[4 * g + 1] > __uint_as_float(t4.y)) ? (1u << (4 * g + 1)) : 0u;
        cm |= (v[4 * g + 2] > __uint_as_float(t4.z)) ? (1u << (4 * g + 2)) : 0u;
        cm |= (v[4 * g + 3] > __uint_as_float(t4.w)) ? (1u << (4 * g + 3)) : 0u;
      }
    }
  }
  const float cmax = fmaxf(fmaxf(gm[0], gm[1]), fmaxf(gm[2], gm[3]));
  const unsigned any_col = gpass ? __reduce_or_sync(0xffffffffu, cm) : 0u;
  const bool any_row = __any_sync(0xffffffffu, cmax > rs.second && cmax >= thr);
  if (any_col) {
    if (__popc(any_col) >= dense_min) {
      // dense: lane e ends up with the segment top-2 of column e and merges it
      unsigned my_m1 = 0u, my_m2 = 0u;
      int my_w = 0;
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        if (any_col & (1u << e)) {  // warp-uniform
          const unsigned xb = (cm & (1u << e)) ? __float_as_uint(v[e]) : 0u;  // hits are > threshold >= 0
          const unsigned m1 = __reduce_max_sync(0xffffffffu, xb);
          const int w = __ffs(__ballot_sync(0xffffffffu, xb == m1)) - 1;      // lowest lane = lowest row
          const unsigned m2 = __reduce_max_sync(0xffffffffu, lane == w ? 0u : xb);
          if (lane == e) {
            my_m1 = m1;
            my_m2 = m2;
            my_w = w;
          }
        }
      }
      if (my_m1 != 0u) {
        const int jl = c * 32 + lane;
        const unsigned long long key =
            ((unsigned long long)my_m1 << 32) | (unsigned long long)(unsigned)((unsigned)row_tag + lane - my_w);
        const unsigned long long old = atomicMax(colkey + jl, key);
        const unsigned cand = key > old ? (unsigned)(old >> 32) : my_m1;
        atomicMax(colsec + jl, max(cand, my_m2));
      }
    } else {
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        if (any_col & (0xFu << (4 * g))) {  // warp-uniform
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            if (cm & (1u << (4 * g + e))) {
              // every score above the threshold is merged: the key keeps the best score
              // and, on ties, the lowest row; each loser's score feeds the second-best
              const int jl = c * 32 + 4 * g + e;
              const unsigned xb = __float_as_uint(v[4 * g + e]);
              const unsigned long long key = ((unsigned long long)xb << 32) | row_tag;
              const unsigned long long old = atomicMax(colkey + jl, key);
              const unsigned cand = key > old ? (unsigned)(old >> 32) : xb;
              atomicMax(colsec + jl, cand);
            }
          }
        }
      }
    }
  }
  if (any_row) {
    // A group of 8 is folded in when ANY lane's group maximum can enter its top-2 (warp-uniform test), and
    // then with selects only: a per-lane branchy update (if / else if per score) cost 30 % of the kernel at
    // d = 128 on data the pruning bound does not remove (profiles/r02_match_d128_noprune_hot_lines.txt).
    // A lane whose state cannot change is unaffected by the selects (sc <= second leaves all three as is).
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (__any_sync(0xffffffffu, gm[g] > rs.second && gm[g] >= thr)) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float sc = v[8 * g + e];
          const bool gt = sc > rs.best;  // strict: the lowest column keeps a tie (sift.cc:747-751)
          rs.second = gt ? rs.best : fmaxf(rs.second, sc);
          rs.idx = gt ? colbase + 8 * g + e : rs.idx;
          rs.best = fmaxf(rs.best, sc);
        }
      }
    }
  }
}

}  // namespace vc
