// Single-pass fused matcher for sm_100a: ONE tcgen05 GEMM per pair yields both the row and the
// column (best, second, index), so the cross check costs no second pass over the tensor cores.
//
// Work item = (pair, 256-descriptor block of image 2).  The CTA keeps that block on the UMMA N
// side and sweeps image 1 in 128-row tiles (TMEM lanes).  Epilogue, per 128x256 fp32 tile in TMEM:
//   rows    : each thread owns one row of the tile; a chunk of 8 scores is skipped unless its max
//             can still change the row's top-2.  The bound is max(local second, global hint); the
//             hint (rowthr) is a monotone lower bound of the row's final second-best published by
//             CTAs working on other column blocks (a stale value is merely less selective).
//             Per-tile partial (best, second, idx) goes to HBM: 12 B per row per 256 columns.
//   columns : the CTA-wide state lives in shared memory as a 64-bit key (score bits << 32 |
//             ~row) plus a 32-bit second-best; a 32-row column segment touches it only when one
//             of its scores beats the current second-best (warp vote); then a warp-wide
//             integer max (REDUX) extracts the segment's top-2 and two shared-memory atomics merge it.
// Exactness: scores <= 0 never match (COLMAP/feature/sift.cc:742-757 start best/second at 0), so
// positive floats are compared as unsigned bit patterns; ties keep the lowest index, as the
// reference's strict '>' scan does.
#include "match_kernel.cuh"

#include <math.h>

#include "vc_ptx.cuh"
#include "vc_launch.cuh"
#include "match_epilogue.cuh"

namespace vc {

namespace {

constexpr int BM = 128, BN = 256, BK = 64, kStages = 4;
constexpr int kABytes = BM * BK * 2, kBBytes = BN * BK * 2, kStageBytes = kABytes + kBBytes;
// warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps 4-19 epilogue: kMatchEpiGroups (4) warps per TMEM lane quarter,
// each on two of the tile's eight 32-column chunks.  The epilogue, not the MMA, paces the kernel below d = 768: two
// warps per quarter were too few (profiles/r01_v8_match_fused_ncu_summary.json), three left the 2-chunk warp idle a
// third of every tile (3 / 3 / 2).
constexpr int kEpiGroups = kMatchEpiGroups;  // epilogue warps per TMEM lane quarter
constexpr int kEpiThreads = kEpiGroups * 128;
constexpr int kThreads = 128 + kEpiThreads;
constexpr int kTmemCols = 512;
constexpr int kColStateBytes = BN * 8 + BN * 4 + (BN / 8) * 4;  // keys, seconds, per-8-column minima
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256 + kColStateBytes;

struct FItem {
  int pair, cb, img1, img2, n1, n2, nrt, start;
  bool valid;
};

__device__ __forceinline__ FItem decode_fitem(const MatchParams& p, int item) {
  FItem it;
  it.pair = item / p.cb_per_img;
  it.cb = item - it.pair * p.cb_per_img;
  it.img1 = __ldg(p.pairs + 2 * it.pair);
  it.img2 = __ldg(p.pairs + 2 * it.pair + 1);
  it.n1 = min(__ldg(p.counts + it.img1), p.kmax);
  it.n2 = min(__ldg(p.counts + it.img2), p.kmax);
  it.nrt = (it.n1 + BM - 1) / BM;
  it.valid = it.cb * BN < it.n2 && it.nrt > 0;
  // stagger the sweep so CTAs on different column blocks of a pair sit on different row tiles:
  // their published hints are then useful to each other
  it.start = it.nrt > 0 ? (it.cb * p.stagger) % it.nrt : 0;
  return it;
}

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory"); }

}  // namespace

__global__ void __launch_bounds__(kThreads, 1)
match_fused_kernel(const __grid_constant__ CUtensorMap tmap, const MatchParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_addr);
  const uint32_t sA = base;
  const uint32_t sB = base + kStages * kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (kStages + s); };
  auto tfull_bar = [&](int s) { return bar0 + 8u * (2 * kStages + s); };
  auto tempty_bar = [&](int s) { return bar0 + 8u * (2 * kStages + 2 + s); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
  unsigned long long* colkey = reinterpret_cast<unsigned long long*>(smem + kStages * kStageBytes + 256);
  unsigned int* colsec = reinterpret_cast<unsigned int*>(colkey + BN);
  unsigned int* colmin8 = colsec + BN;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nkb = (p.d + BK - 1) / BK;

  if (warp == 0 && lane == 0) tma_prefetch_desc(&tmap);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), kEpiThreads);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer
    uint32_t stage = 0, phase = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const FItem it = decode_fitem(p, item);
      if (!it.valid) continue;
      for (int s = 0; s < it.nrt; ++s) {
        int rt = it.start + s;
        if (rt >= it.nrt) rt -= it.nrt;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          mbar_arrive_expect_tx(full_bar(stage), kStageBytes);
          tma_load_3d(sA + stage * kABytes, &tmap, full_bar(stage), kb * BK, rt * BM, it.img1);
          tma_load_3d(sB + stage * kBBytes, &tmap, full_bar(stage), kb * BK, it.cb * BN, it.img2);
          tma_load_3d(sB + stage * kBBytes + kABytes, &tmap, full_bar(stage), kb * BK,
                      it.cb * BN + 128, it.img2);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
    uint32_t stage = 0, phase = 0, tile = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const FItem it = decode_fitem(p, item);
      if (!it.valid) continue;
      for (int s = 0; s < it.nrt; ++s, ++tile) {
        const uint32_t as = tile & 1, aph = (tile >> 1) & 1;
        mbar_wait(tempty_bar(as), aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint64_t da = umma_desc_k_sw128(sA + stage * kABytes);
          const uint64_t db = umma_desc_k_sw128(sB + stage * kBBytes);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            tc_mma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          tc_commit(empty_bar(stage));
          if (kb == nkb - 1) tc_commit(tfull_bar(as));
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue (12 warps)
    // warp -> TMEM lane quarter q = warp % 4 (hardware rule) and column group h of the tile
    const int q = warp & 3;
    const int h = (warp - 4) >> 2;
    const int et = threadIdx.x - 128;  // 0..kEpiThreads-1
    uint32_t tile = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const FItem it = decode_fitem(p, item);
      if (!it.valid) continue;
      if (et < BN) {
        colkey[et] = 0ull;
        colsec[et] = __float_as_uint(p.prune);  // scores at or below the bound never reach the column state
        if (et < BN / 8) colmin8[et] = __float_as_uint(p.prune);
      }
      epi_bar_sync();
      const int ncols = min(BN, it.n2 - it.cb * BN);  // valid columns of this block
      // this warp's chunks (32 columns each): group 0 -> 0..2, group 1 -> 3..5, group 2 -> 6..7
      const int c_begin = h * kMatchChunksPerGroup;
      const int c_end = min(min(c_begin + kMatchChunksPerGroup, BN / 32), (ncols + 31) >> 5);
      const size_t pair_row0 = (size_t)it.pair * p.kmax;
      // hint: lower bound of a row's final second-best (L2 load, may be stale).  The load for the NEXT row
      // tile is issued before this tile's chunks are processed, so its latency is off the critical path.
      auto row_of = [&](int s) {
        int rt = it.start + s;
        if (rt >= it.nrt) rt -= it.nrt;
        return rt * BM + q * 32 + lane;
      };
      unsigned hint_next = 0u;
      if (it.nrt > 0 && row_of(0) < it.n1) hint_next = __ldcg(p.rowthr + pair_row0 + row_of(0));
      for (int s = 0; s < it.nrt; ++s, ++tile) {
        const int row = row_of(s);
        const bool row_ok = row < it.n1;
        float thr = 0.f;
        if (row_ok) thr = fmaxf(p.prune, __uint_as_float(hint_next));
        if (s + 1 < it.nrt && row_of(s + 1) < it.n1) hint_next = __ldcg(p.rowthr + pair_row0 + row_of(s + 1));
        const unsigned long long row_tag = 0xFFFFFFFFu - (unsigned)row;
        RowTop2 rs{0.f, 0.f, -1};
        const uint32_t as = tile & 1, aph = (tile >> 1) & 1;
        if (s > 0) refresh_colmin8(colsec, colmin8, 4 * c_begin, 4 * c_end, lane);
        mbar_wait(tfull_bar(as), aph);
        tc_fence_after();
        const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
#pragma unroll 1
        for (int c = c_begin; c < c_end; ++c)
          fused_epilogue_chunk(t_addr + c * 32, c, ncols, row_ok, row_tag, thr, it.cb * BN + c * 32, lane, p.dense_min,
                               colkey, colsec, colmin8, rs);
        tc_fence_before();
        mbar_arrive(tempty_bar(as));
        if (row_ok) {
          const size_t o = ((size_t)it.pair * p.cb_per_img * kEpiGroups + it.cb * kEpiGroups + h) * p.kmax + row;
          p.part_best[o] = rs.best;
          p.part_second[o] = rs.second;
          p.part_idx[o] = rs.idx;
          if (rs.second > thr) atomicMax(p.rowthr + pair_row0 + row, __float_as_uint(rs.second));
        }
      }
      // ---- column block finished: every row of image 1 has been seen
      epi_bar_sync();
      if (et < ncols) {
        const unsigned long long key = colkey[et];
        const float cbest = __uint_as_float((unsigned)(key >> 32));
        const float csec = __uint_as_float(colsec[et]);
        const int cidx = key ? (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull)) : -1;
        const size_t o = ((size_t)it.pair * 2 + 1) * p.kmax + it.cb * BN + et;
        p.m_out[o] = accept_scores(cbest, csec, cidx, p);
        if (p.top_best != nullptr) {
          p.top_best[o] = cbest;
          p.top_second[o] = csec;
          p.top_idx[o] = cidx;
        }
      }
      epi_bar_sync();  // state is re-initialised at the top of the next item
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

// Merge the per-column-block row partials in ascending column order (the reference's scan order,
// sift.cc:745-757) and apply the distance / ratio tests -> m_out[pair][0][row].
__global__ void merge_rows_kernel(const MatchParams p) {
  const int pair = blockIdx.x;  // pairs on x: no 65535 limit
  const int row = blockIdx.y * blockDim.x + threadIdx.x;
  const int i0 = p.pairs[2 * pair], i1 = p.pairs[2 * pair + 1];
  const int n1 = min(p.counts[i0], p.kmax), n2 = min(p.counts[i1], p.kmax);
  if (row >= n1) return;
  const int nparts = ((n2 + BN - 1) / BN) * kEpiGroups;  // one partial per (column block, epilogue column group)
  float best = 0.f, second = 0.f;
  int idx = -1;
  for (int cb = 0; cb < nparts; ++cb) {
    const size_t o = ((size_t)pair * p.cb_per_img * kEpiGroups + cb) * p.kmax + row;
    const float pb = p.part_best[o], ps = p.part_second[o];
    const int pi = p.part_idx[o];
    if (pi < 0) continue;
    if (pb > best) {
      second = fmaxf(best, ps);
      best = pb;
      idx = pi;
    } else {
      second = fmaxf(second, pb);
    }
  }
  const size_t o = ((size_t)pair * 2) * p.kmax + row;
  p.m_out[o] = accept_scores(best, second, idx, p);
  if (p.top_best != nullptr) {
    p.top_best[o] = best;
    p.top_second[o] = second;
    p.top_idx[o] = idx;
  }
}

int match_fused_smem_bytes() { return kSmemBytes; }

cudaError_t launch_match_fused(const CUtensorMap& tmap, const MatchParams& p, int num_sms,
                               cudaStream_t stream) {
  static SmemOptIn opt_in;
  if (cudaError_t e = ensure_dynamic_smem(match_fused_kernel, kSmemBytes, opt_in); e != cudaSuccess) return e;
  if (p.n_items <= 0) return cudaSuccess;
  cudaError_t e = cudaMemsetAsync(p.rowthr, 0, (size_t)p.n_pairs * p.kmax * sizeof(unsigned int), stream);
  if (e != cudaSuccess) return e;
  const int grid = p.n_items < num_sms ? p.n_items : num_sms;
  match_fused_kernel<<<grid, kThreads, kSmemBytes, stream>>>(tmap, p);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  dim3 g(p.n_pairs, (p.kmax + 255) / 256);
  merge_rows_kernel<<<g, 256, 0, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace vc
