// Two-SM (cta_group::2) form of the single-pass fused matcher in match_fused.cu.  A cluster of two
// CTAs on one TPC works on one (pair, 256-descriptor block of image 2) item: each CTA stages its own
// 128-row tile of image 1 and HALF of the column block (128 descriptors), the leader CTA issues one
// tcgen05.mma.cta_group::2 of M = 256, N = 256 per k-step, and each CTA finds the fp32 scores of its
// own 128 rows x all 256 columns in its TMEM.  Per 512 tensor cycles a CTA now pulls 32 KB through
// L2 instead of 48 KB (the single-CTA kernel is bound by that feed: 600 GB of L2->SM traffic per 496
// pairs, profiles/r01_v8_match_fused_ncu_summary.json), and the shared-memory ring holds 6 stages
// instead of 4.  Epilogue as in match_fused.cu; the two CTAs' column states are written to HBM and
// merged by merge_cols2_kernel (no cross-CTA synchronisation inside the sweep).
//
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

constexpr int BM = 128, BN = 256, BK = 64, kStages = 6;
constexpr int kABytes = BM * BK * 2, kBBytes = (BN / 2) * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kEpiGroups = kMatchEpiGroups;  // epilogue warps per TMEM lane quarter
constexpr int kEpiWarps = 4 * kEpiGroups;
constexpr int kEpiThreads = 32 * kEpiWarps;
constexpr int kThreads = 128 + kEpiThreads;  // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps 4-19 epilogue
constexpr int kTmemCols = 512;
constexpr int kColStateBytes = BN * 8 + BN * 4 + (BN / 8) * 4;  // keys, seconds, per-8-column minima
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256 + kColStateBytes;

struct FItem {
  int pair, cb, img1, img2, n1, n2, nrt, npt, start;
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
  it.npt = (it.nrt + 1) / 2;  // row-tile pairs: CTA rank r of the cluster takes tile 2 * pair + r
  it.valid = it.cb * BN < it.n2 && it.nrt > 0;
  it.start = it.npt > 0 ? (it.cb * p.stagger) % it.npt : 0;
  return it;
}

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory"); }

// ---- cluster / two-SM PTX (cta_group::2)
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_rank(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  // default semantics (.release.cta): no cluster-scope fence, which would wait for this warp's
  // outstanding global stores.  TMEM hand-over is ordered by the tcgen05 fences on either side.
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion bytes are counted on the LEADER CTA's mbarrier (bar_leader is a
// shared::cluster address of rank 0); the data lands in this CTA's shared memory.
__device__ __forceinline__ void tma_load_3d_2sm(uint32_t dst_smem, const CUtensorMap* m, uint32_t bar_leader,
                                                int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_leader), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_2sm() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// arrive (once all earlier MMAs of this thread are done) on the barrier at this offset in both CTAs
__device__ __forceinline__ void tc_commit_2sm(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(bar), "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void tc_mma_bf16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

}  // namespace

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
match_fused2_kernel(const __grid_constant__ CUtensorMap tmap, const MatchParams p) {
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
  const uint32_t rank = cluster_rank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) tma_prefetch_desc(&tmap);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 2 * kEpiWarps);  // one arrival per epilogue warp of either CTA
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_2sm(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  cluster_sync_all();  // barriers of both CTAs initialised before any remote arrival / multicast
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer
    uint32_t stage = 0, phase = 0;
    for (int item = cluster_id; item < p.n_items; item += n_clusters) {
      const FItem it = decode_fitem(p, item);
      if (!it.valid) continue;
      for (int s = 0; s < it.npt; ++s) {
        int rtp = it.start + s;
        if (rtp >= it.npt) rtp -= it.npt;
        const int rt = 2 * rtp + (int)rank;  // may be one past the last tile: rows are masked, TMA zero-fills
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t lead_full = map_to_rank(full_bar(stage), 0);
          if (rank == 0) mbar_arrive_expect_tx(full_bar(stage), 2 * kStageBytes);  // both CTAs' bytes
          tma_load_3d_2sm(sA + stage * kABytes, &tmap, lead_full, kb * BK, rt * BM, it.img1);
          tma_load_3d_2sm(sB + stage * kBBytes, &tmap, lead_full, kb * BK, it.cb * BN + (int)rank * (BN / 2),
                          it.img2);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1 && lane == 0 && rank == 0) {
    // ------------------------------------------------------------ MMA issuer (leader CTA only)
    constexpr uint32_t idesc = umma_idesc_bf16(2 * BM, BN);
    uint32_t stage = 0, phase = 0, tile = 0;
    for (int item = cluster_id; item < p.n_items; item += n_clusters) {
      const FItem it = decode_fitem(p, item);
      if (!it.valid) continue;
      for (int s = 0; s < it.npt; ++s, ++tile) {
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
            tc_mma_bf16_2sm(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          tc_commit_2sm(empty_bar(stage));
          if (kb == nkb - 1) tc_commit_2sm(tfull_bar(as));
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue (8 warps)
    // warp -> TMEM lane quarter q = warp % 4 (hardware rule) and column half h of the tile
    const int q = warp & 3;
    const int h = (warp - 4) >> 2;
    const int et = threadIdx.x - 128;  // 0..255
    uint32_t tile = 0;
    const uint32_t lead_tempty[2] = {map_to_rank(tempty_bar(0), 0), map_to_rank(tempty_bar(1), 0)};
    for (int item = cluster_id; item < p.n_items; item += n_clusters) {
      const FItem it = decode_fitem(p, item);
      if (!it.valid) continue;
      if (et < BN) {
        colkey[et] = 0ull;
        colsec[et] = __float_as_uint(p.prune);  // scores at or below the bound never reach the column state
        if (et < BN / 8) colmin8[et] = __float_as_uint(p.prune);
      }
      epi_bar_sync();
      const int ncols = min(BN, it.n2 - it.cb * BN);  // valid columns of this block
      // this warp's chunks: columns [h*128, h*128+128) of the block, 32 per chunk
      const int c_begin = h * kMatchChunksPerGroup;
      const int c_end = min(min(c_begin + kMatchChunksPerGroup, BN / 32), (ncols + 31) >> 5);
      const size_t pair_row0 = (size_t)it.pair * p.kmax;
      // hint: lower bound of a row's final second-best (L2 load, may be stale).  The load for the NEXT row
      // tile is issued before this tile's chunks are processed, so its latency is off the critical path.
      auto row_of = [&](int s) {
        int rtp = it.start + s;
        if (rtp >= it.npt) rtp -= it.npt;
        return (2 * rtp + (int)rank) * BM + q * 32 + lane;
      };
      unsigned hint_next = 0u;
      if (it.npt > 0 && row_of(0) < it.n1) hint_next = __ldcg(p.rowthr + pair_row0 + row_of(0));
      for (int s = 0; s < it.npt; ++s, ++tile) {
        const int row = row_of(s);
        const bool row_ok = row < it.n1;
        float thr = 0.f;
        if (row_ok) thr = fmaxf(p.prune, __uint_as_float(hint_next));
        if (s + 1 < it.npt && row_of(s + 1) < it.n1) hint_next = __ldcg(p.rowthr + pair_row0 + row_of(s + 1));
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
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(lead_tempty[as]);
        if (row_ok) {
          const size_t o = ((size_t)it.pair * p.cb_per_img * kEpiGroups + it.cb * kEpiGroups + h) * p.kmax + row;
          p.part_best[o] = rs.best;
          p.part_second[o] = rs.second;
          p.part_idx[o] = rs.idx;
          if (rs.second > thr) atomicMax(p.rowthr + pair_row0 + row, __float_as_uint(rs.second));
        }
      }
      // ---- column block finished in this CTA (its half of the row tiles): state -> HBM, merged with
      // the other CTA's by merge_cols2_kernel
      epi_bar_sync();
      if (et < ncols) {
        const size_t o = ((size_t)it.pair * 2 + rank) * p.kmax + it.cb * BN + et;
        p.col_key[o] = colkey[et];
        p.col_sec[o] = colsec[et];
      }
      epi_bar_sync();  // state is re-initialised at the top of the next item
    }
  }

  tc_fence_before();
  cluster_sync_all();  // the peer may still multicast into this CTA's barriers / read its shared memory
  if (warp == 2) tmem_dealloc_2sm(tmem_base, kTmemCols);
}

// Column results: merge the two CTAs' (key, second) states.  key = score bits << 32 | ~row, so the
// larger key is the better score (lowest row on ties); the loser's best score competes for second.
__global__ void merge_cols2_kernel(const MatchParams p) {
  const int pair = blockIdx.x;  // pairs on x: no 65535 limit
  const int col = blockIdx.y * blockDim.x + threadIdx.x;
  const int i1 = p.pairs[2 * pair + 1];
  const int n2 = min(p.counts[i1], p.kmax);
  const int n1 = min(p.counts[p.pairs[2 * pair]], p.kmax);
  if (col >= n2 || n1 <= 0) return;
  const size_t o0 = ((size_t)pair * 2) * p.kmax + col, o1 = o0 + p.kmax;
  const unsigned long long k0 = p.col_key[o0], k1 = p.col_key[o1];
  const unsigned long long key = k0 > k1 ? k0 : k1, lose = k0 > k1 ? k1 : k0;
  unsigned sec = max(p.col_sec[o0], p.col_sec[o1]);
  sec = max(sec, (unsigned)(lose >> 32));
  const float cbest = __uint_as_float((unsigned)(key >> 32));
  const float csec = __uint_as_float(sec);
  const int cidx = key ? (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull)) : -1;
  const size_t o = ((size_t)pair * 2 + 1) * p.kmax + col;
  p.m_out[o] = accept_scores(cbest, csec, cidx, p);
  if (p.top_best != nullptr) {
    p.top_best[o] = cbest;
    p.top_second[o] = csec;
    p.top_idx[o] = cidx;
  }
}

// Merge the per-column-block row partials in ascending column order (the reference's scan order,
// sift.cc:745-757) and apply the distance / ratio tests -> m_out[pair][0][row].
__global__ void merge_rows2_kernel(const MatchParams p) {
  const int pair = blockIdx.x;  // pairs on x: no 65535 limit
  const int row = blockIdx.y * blockDim.x + threadIdx.x;
  const int i0 = p.pairs[2 * pair], i1 = p.pairs[2 * pair + 1];
  const int n1 = min(p.counts[i0], p.kmax), n2 = min(p.counts[i1], p.kmax);
  if (row >= n1) return;
  const int nparts = ((n2 + BN - 1) / BN) * kEpiGroups;
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

int match_fused2_smem_bytes() { return kSmemBytes; }

cudaError_t launch_match_fused2(const CUtensorMap& tmap, const MatchParams& p, int num_sms,
                               cudaStream_t stream) {
  static SmemOptIn opt_in;
  if (cudaError_t e = ensure_dynamic_smem(match_fused2_kernel, kSmemBytes, opt_in); e != cudaSuccess) return e;
  if (p.n_items <= 0) return cudaSuccess;
  cudaError_t e = cudaMemsetAsync(p.rowthr, 0, (size_t)p.n_pairs * p.kmax * sizeof(unsigned int), stream);
  if (e != cudaSuccess) return e;
  const int clusters = p.n_items < num_sms / 2 ? p.n_items : num_sms / 2;
  match_fused2_kernel<<<2 * clusters, kThreads, kSmemBytes, stream>>>(tmap, p);  // __cluster_dims__(2,1,1)
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  dim3 g(p.n_pairs, (p.kmax + 255) / 256);
  merge_rows2_kernel<<<g, 256, 0, stream>>>(p);
  merge_cols2_kernel<<<g, 256, 0, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace vc
