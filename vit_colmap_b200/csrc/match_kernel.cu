// Fused exhaustive descriptor matcher for sm_100a (B200).
//
// One persistent CTA per SM.  A work item is (pair, direction, 128-row block): the rows of the
// "M" image of the pair sit on the 128 TMEM lanes, the CTA sweeps the "N" image in 256-column
// tiles.  Per tile: TMA (128B swizzle) stages bf16 operands in shared memory, one thread issues
// tcgen05.mma (M=128, N=256, K=16) into a double-buffered fp32 TMEM accumulator, four epilogue
// warps pull the tile with tcgen05.ld and fold it into a per-row running (best, second, index).
// The score matrix never leaves TMEM.  Direction 1 swaps the roles of the two images and yields
// the column results, so the cross check (COLMAP/feature/sift.cc:803-813) needs no transpose.
//
// Semantics restated from the reference (not copied):
//   COLMAP/feature/sift.cc:733-787  one-way best/second, init 0, strict '>', acos tests
//   COLMAP/feature/sift.cc:789-825  cross check, ascending idx1 emit order
//   SIFTGPU/ProgramCU.cu:1411-1495,1696-1797  (the reference GPU kernels this replaces; they
//   write the int32 dot matrix to HBM and re-read it twice)
#include "match_kernel.cuh"

#include <math.h>

#include "vc_ptx.cuh"
#include "vc_launch.cuh"

namespace vc {

namespace {

constexpr int BM = 128;            // rows per work item (TMEM lanes)
constexpr int BN = 256;            // columns per tile (UMMA N)
constexpr int BK = 64;             // bf16 per k-block = one 128-byte swizzle row
constexpr int kStages = 4;         // smem ring depth
constexpr int kABytes = BM * BK * 2;
constexpr int kBBytes = BN * BK * 2;
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kThreads = 256;      // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps4-7 epilogue
constexpr int kTmemCols = 512;     // 2 accumulator stages x 256 fp32 columns
constexpr int kGuidedFeatBytes = 4 * 32 * 2 * 16;  // guided matching: per epilogue warp, 32 columns x 2 float4
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/ + kGuidedFeatBytes;

struct Item {
  int pair, dir, rb, img_m, img_n, n_m, n_n;
  bool valid;
};

__device__ __forceinline__ Item decode_item(const MatchParams& p, int item) {
  Item it;
  const int per_pair = 2 * p.rb_per_img;
  it.pair = item / per_pair;
  const int r = item - it.pair * per_pair;
  it.dir = r / p.rb_per_img;
  it.rb = r - it.dir * p.rb_per_img;
  const int i0 = __ldg(p.pairs + 2 * it.pair), i1 = __ldg(p.pairs + 2 * it.pair + 1);
  it.img_m = it.dir ? i1 : i0;
  it.img_n = it.dir ? i0 : i1;
  it.n_m = min(__ldg(p.counts + it.img_m), p.kmax);
  it.n_n = min(__ldg(p.counts + it.img_n), p.kmax);
  it.valid = it.rb * BM < it.n_m;
  return it;
}

}  // namespace

template <bool kDebug, bool kGuided>
__global__ void __launch_bounds__(kThreads, 1)
match_top2_kernel(const __grid_constant__ CUtensorMap tmap, const MatchParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;  // 128B swizzle atoms need 1024B alignment
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
  float4* feat_all = reinterpret_cast<float4*>(smem + kStages * kStageBytes + 256);
  (void)feat_all;

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
      mbar_init(tempty_bar(s), 128);
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
      const Item it = decode_item(p, item);
      if (!it.valid) continue;
      const int ntiles = (it.n_n + BN - 1) / BN;
      for (int ct = 0; ct < ntiles; ++ct) {
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          mbar_arrive_expect_tx(full_bar(stage), kStageBytes);
          tma_load_3d(sA + stage * kABytes, &tmap, full_bar(stage), kb * BK, it.rb * BM, it.img_m);
          tma_load_3d(sB + stage * kBBytes, &tmap, full_bar(stage), kb * BK, ct * BN, it.img_n);
          tma_load_3d(sB + stage * kBBytes + kABytes, &tmap, full_bar(stage), kb * BK,
                      ct * BN + 128, it.img_n);
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
      const Item it = decode_item(p, item);
      if (!it.valid) continue;
      const int ntiles = (it.n_n + BN - 1) / BN;
      for (int ct = 0; ct < ntiles; ++ct, ++tile) {
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
          for (int k = 0; k < BK / 16; ++k) {
            // +32 bytes per K=16 step inside the 128-byte swizzle row (address field is >>4)
            tc_mma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          }
          tc_commit(empty_bar(stage));  // frees the smem slot when these MMAs retire
          if (kb == nkb - 1) tc_commit(tfull_bar(as));
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue: running top-2
    const int q = warp - 4;  // TMEM lane quarter this warp may access (warp % 4)
    uint32_t tile = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const Item it = decode_item(p, item);
      if (!it.valid) continue;
      const int ntiles = (it.n_n + BN - 1) / BN;
      const int row = it.rb * BM + q * 32 + lane;
      // sift.cc:741-743 start best = second = 0, idx = -1.  p.prune (guided match lists only, 0 otherwise) is the
      // bound below which a score can change no accept / reject decision (vc_api.cu, match_impl): starting
      // there instead of at 0 gives the same match list and spares the geometric gate for every such score.
      float best = kGuided ? p.prune : 0.f, second = best;
      int bidx = -1;
      // guided matching: this row's keypoint, the pair's geometry, the other image's keypoint rows
      GuidedGeometry geom;
      float rx = 0.f, ry = 0.f;
      const float2* kp_n = nullptr;
      GuidedA row_a{};   // this row's keypoint as image-1 point (direction 0) ...
      GuidedB row_b{};   // ... or as image-2 point (direction 1): its products with H / F
      bool skip_h = false;
      float4* feat = feat_all + (warp - 4) * 64;  // this warp's staging: 32 columns x 2 float4
      if (kGuided) {
        geom = p.geom[it.pair];
        const int km = it.img_m >= p.kp_images ? it.img_m - p.kp_images : it.img_m;
        const int kn = it.img_n >= p.kp_images ? it.img_n - p.kp_images : it.img_n;
        const float2* kp_m = reinterpret_cast<const float2*>(p.kp) + (size_t)km * p.kmax;
        kp_n = reinterpret_cast<const float2*>(p.kp) + (size_t)kn * p.kmax;
        if (row < it.n_m) {
          const float2 r = __ldg(kp_m + row);
          rx = r.x;
          ry = r.y;
        }
        skip_h = geom.mode == 0 && geom.max_h >= 1e19f;
        if (it.dir == 0) row_a = guided_features_a(geom, rx, ry);
        else row_b = guided_features_b(geom, rx, ry);
      }
      for (int ct = 0; ct < ntiles; ++ct, ++tile) {
        const uint32_t as = tile & 1, aph = (tile >> 1) & 1;
        mbar_wait(tfull_bar(as), aph);
        tc_fence_after();
        const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          const int colbase = ct * BN + c * 32;
          if (colbase >= it.n_n) break;  // warp-uniform
          uint32_t vr[32];
          tmem_ld_32x32(t_addr + c * 32, vr);
          tmem_wait_ld();
          float v[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(vr[e]);
          if (colbase + 32 > it.n_n) {  // ragged last chunk: padding rows never match
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (colbase + e >= it.n_n) v[e] = 0.f;
          }
          if (kGuided) {
            // Geometric gate of guided matching.  Candidates that violate the pair's geometry score
            // `masked_score` (SIFTGPU/ProgramCU.cu:1573-1603 drives them to <= 0, COLMAP/feature/sift.cc:945-951
            // to distance 512^2).  Lane e holds the keypoint of column colbase + e and stages that point's
            // products with H / F in shared memory; every lane then reads entry e as a broadcast load and
            // decides all 32 candidates of its row branch-free (guided_fast); the few candidates whose
            // residual lies within rounding of the bound are re-decided by the exactly specified form.
            // Skipped for a chunk in which no score, gated or not, can enter any row's top-2.
            // Direction 1 has the roles swapped: the row is the image-2 keypoint.
            float cmx = v[0];
#pragma unroll
            for (int e = 1; e < 32; ++e) cmx = fmaxf(cmx, v[e]);
            if (__any_sync(0xffffffffu, fmaxf(cmx, p.masked_score) > second)) {
              const int cj = colbase + lane;
              const float2 mine = cj < it.n_n ? __ldg(kp_n + cj) : make_float2(0.f, 0.f);
              float4* slot = feat + 2 * lane;
              if (it.dir == 0) {
                const GuidedB cb = guided_features_b(geom, mine.x, mine.y);
                slot[0] = make_float4(cb.x, cb.y, cb.n2, 0.f);
              } else {
                const GuidedA ca = guided_features_a(geom, mine.x, mine.y);
                slot[0] = make_float4(ca.a, ca.b, ca.c, ca.hb);
                slot[1] = make_float4(ca.f0, ca.f1, ca.f2, ca.n1);
              }
              __syncwarp();
              // which tests the pair's mode asks for (warp-uniform): 2 directions x {H, F, H and F}
              const bool do_h = geom.mode != 1 && !skip_h, do_f = geom.mode != 2;
              const int n_live = it.n_n - colbase;
              unsigned undecided;
#define VC_GATE(D, H, F) guided_gate_chunk<D, H, F>(feat, row_a, row_b, geom.max_f, p.masked_score, n_live, v)
              if (it.dir == 0)
                undecided = do_h ? (do_f ? VC_GATE(false, true, true) : VC_GATE(false, true, false)) : VC_GATE(false, false, true);
              else
                undecided = do_h ? (do_f ? VC_GATE(true, true, true) : VC_GATE(true, true, false)) : VC_GATE(true, false, true);
#undef VC_GATE
              const unsigned any_und = __reduce_or_sync(0xffffffffu, undecided);
              if (any_und) {
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                  if (any_und & (1u << e)) {  // warp-uniform
                    const float cx = __shfl_sync(0xffffffffu, mine.x, e), cy = __shfl_sync(0xffffffffu, mine.y, e);
                    if (undecided & (1u << e)) {
                      const bool keep = it.dir == 0 ? guided_keep(geom, rx, ry, cx, cy) : guided_keep(geom, cx, cy, rx, ry);
                      if (!keep) v[e] = p.masked_score;
                    }
                  }
                }
              }
              __syncwarp();  // the staging slots are rewritten by the next chunk
            }
          }
          if (kDebug) {
            if (p.dbg_scores != nullptr && it.pair == 0 && it.dir == 0 && row < it.n_m) {
#pragma unroll
              for (int e = 0; e < 32; ++e)
                if (colbase + e < it.n_n)
                  p.dbg_scores[(size_t)row * p.dbg_ld + colbase + e] = v[e];
            }
          }
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            float m = v[8 * g];
#pragma unroll
            for (int e = 1; e < 8; ++e) m = fmaxf(m, v[8 * g + e]);
            // Nothing in this group can change (best, second, idx) unless it beats `second`.
            if (__any_sync(0xffffffffu, m > second)) {
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const float s = v[8 * g + e];
                if (s > best) {  // strict '>' keeps the lowest index on ties (sift.cc:747-751)
                  second = best;
                  best = s;
                  bidx = colbase + 8 * g + e;
                } else if (s > second) {
                  second = s;
                }
              }
            }
          }
        }
        tc_fence_before();
        mbar_arrive(tempty_bar(as));
      }
      if (row < it.n_m) {
        const size_t o = ((size_t)it.pair * 2 + it.dir) * p.kmax + row;
        p.m_out[o] = accept_scores(best, second, bidx, p);
        if (p.top_best != nullptr) {
          p.top_best[o] = best;
          p.top_second[o] = second;
          p.top_idx[o] = bidx;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------------------------------------
// Cross check + ordered emit (COLMAP/feature/sift.cc:803-823; replaces the host loop
// SIFTGPU/SiftMatchCU.cpp:183-191).  One CTA per pair; a block-wide exclusive scan keeps the
// output sorted by idx1.
__global__ void __launch_bounds__(1024, 1) emit_matches_kernel(const MatchParams p, uint32_t* matches,
                                                               int32_t* n_matches) {
  __shared__ int warp_tot[32];
  __shared__ int running;
  const int pair = blockIdx.x;
  const int i0 = p.pairs[2 * pair], i1 = p.pairs[2 * pair + 1];
  const int n1 = min(p.counts[i0], p.kmax), n2 = min(p.counts[i1], p.kmax);
  const int32_t* m12 = p.m_out + ((size_t)pair * 2) * p.kmax;
  const int32_t* m21 = m12 + p.kmax;
  uint32_t* out = matches + (size_t)pair * p.kmax * 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) running = 0;
  __syncthreads();
  if (n2 > 0) {
    for (int base = 0; base < n1; base += blockDim.x) {
      const int i = base + threadIdx.x;
      int j = -1;
      if (i < n1) {
        j = m12[i];
        if (j >= 0 && p.cross_check && m21[j] != i) j = -1;
      }
      const unsigned bal = __ballot_sync(0xffffffffu, j >= 0);
      if (lane == 0) warp_tot[warp] = __popc(bal);
      __syncthreads();
      int prefix = running;
      for (int w = 0; w < warp; ++w) prefix += warp_tot[w];
      if (j >= 0) {
        const int pos = prefix + __popc(bal & ((1u << lane) - 1u));
        out[2 * pos] = (uint32_t)i;
        out[2 * pos + 1] = (uint32_t)j;
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        int tot = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += warp_tot[w];
        running += tot;
      }
      __syncthreads();
    }
  }
  if (threadIdx.x == 0) n_matches[pair] = running;
}

__global__ void u8_to_bf16_kernel(const uint8_t* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                  size_t n) {
  // 16 elements per thread: one 16-byte load, two 16-byte stores
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (i + 16 <= n) {
    const uint4 in = *reinterpret_cast<const uint4*>(src + i);
    const uint32_t w[4] = {in.x, in.y, in.z, in.w};
    __align__(16) __nv_bfloat16 o[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) o[k] = __float2bfloat16_rn((float)((w[k >> 2] >> (8 * (k & 3))) & 0xFF));
    *reinterpret_cast<uint4*>(dst + i) = *reinterpret_cast<const uint4*>(o);
    *reinterpret_cast<uint4*>(dst + i + 8) = *reinterpret_cast<const uint4*>(o + 8);
  } else {
    for (size_t k = i; k < n; ++k) dst[k] = __float2bfloat16_rn((float)src[k]);
  }
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                   size_t n) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i + 8 <= n) {
    const float4 a = *reinterpret_cast<const float4*>(src + i);
    const float4 b = *reinterpret_cast<const float4*>(src + i + 4);
    __align__(16) __nv_bfloat16 o[8] = {
        __float2bfloat16_rn(a.x), __float2bfloat16_rn(a.y), __float2bfloat16_rn(a.z),
        __float2bfloat16_rn(a.w), __float2bfloat16_rn(b.x), __float2bfloat16_rn(b.y),
        __float2bfloat16_rn(b.z), __float2bfloat16_rn(b.w)};
    *reinterpret_cast<uint4*>(dst + i) = *reinterpret_cast<const uint4*>(o);
  } else {
    for (size_t k = i; k < n; ++k) dst[k] = __float2bfloat16_rn(src[k]);
  }
}

// Squared-L2 matching through the dot-product kernels (rule VC_RULE_L2): rows are augmented by 8
// columns so that (image-1 row) . (image-2 row) = K - |x - y|^2 exactly,
//   role 1: [2x | D0 D1 D2 | 1 1 1 1]              |x|^2 = D0 + 256 D1 + 65536 D2
//   role 2: [ y | -1 -256 -65536 | -E0 -256 E1 -65536 E2 | K]   |y|^2 = E0 + 256 E1 + 65536 E2
// every entry an integer with at most 8 significant bits (exact in bf16), every partial sum < 2^24.
__global__ void l2_max_norm2_kernel(const uint8_t* __restrict__ src, size_t rows, int d, int* __restrict__ out) {
  const size_t row = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const uint8_t* r = src + row * d;
  int acc = 0;
  for (int k = lane; k < d; k += 32) acc += (int)r[k] * (int)r[k];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0 && acc > 0) atomicMax(out, acc);
}

__global__ void l2_augment_kernel(const uint8_t* __restrict__ src, size_t rows, int d, float offset_k,
                                  __nv_bfloat16* __restrict__ dst) {
  const size_t row = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const uint8_t* r = src + row * d;
  __nv_bfloat16* o1 = dst + row * (d + 8);
  __nv_bfloat16* o2 = dst + (rows + row) * (d + 8);
  int acc = 0;
  for (int k = lane; k < d; k += 32) {
    const int v = r[k];
    acc += v * v;
    o1[k] = __float2bfloat16_rn((float)(2 * v));
    o2[k] = __float2bfloat16_rn((float)v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane < 8) {
    const float dig[3] = {(float)(acc & 255), (float)((acc >> 8) & 255), (float)(acc >> 16)};
    const float w[3] = {1.f, 256.f, 65536.f};
    float a, b;
    if (lane < 3) {
      a = dig[lane];
      b = -w[lane];
    } else if (lane < 6) {
      a = 1.f;
      b = -dig[lane - 3] * w[lane - 3];
    } else if (lane == 6) {
      a = 1.f;
      b = offset_k;
    } else {
      a = 0.f;  // 8th column: padding to a multiple of 8 (16-byte rows)
      b = 0.f;
    }
    o1[d + lane] = __float2bfloat16_rn(a);
    o2[d + lane] = __float2bfloat16_rn(b);
  }
}

// ---------------------------------------------------------------------------------------------
// Host launchers

int match_smem_bytes() { return kSmemBytes; }

cudaError_t launch_match_top2(const CUtensorMap& tmap, const MatchParams& p, int num_sms, bool debug,
                              cudaStream_t stream) {
  static SmemOptIn opt_in[2];
  auto kern = debug ? match_top2_kernel<true, false> : match_top2_kernel<false, false>;
  if (cudaError_t e = ensure_dynamic_smem(kern, kSmemBytes, opt_in[debug]); e != cudaSuccess) return e;
  if (p.n_items <= 0) return cudaSuccess;
  const int grid = p.n_items < num_sms ? p.n_items : num_sms;
  kern<<<grid, kThreads, kSmemBytes, stream>>>(tmap, p);
  return cudaGetLastError();
}

cudaError_t launch_match_top2_guided(const CUtensorMap& tmap, const MatchParams& p, int num_sms,
                                     cudaStream_t stream) {
  static SmemOptIn opt_in;
  if (cudaError_t e = ensure_dynamic_smem(match_top2_kernel<false, true>, kSmemBytes, opt_in); e != cudaSuccess)
    return e;
  if (p.n_items <= 0) return cudaSuccess;
  const int grid = p.n_items < num_sms ? p.n_items : num_sms;
  match_top2_kernel<false, true><<<grid, kThreads, kSmemBytes, stream>>>(tmap, p);
  return cudaGetLastError();
}

cudaError_t launch_emit_matches(const MatchParams& p, uint32_t* matches, int32_t* n_matches,
                                cudaStream_t stream) {
  if (p.n_pairs <= 0) return cudaSuccess;
  emit_matches_kernel<<<p.n_pairs, 1024, 0, stream>>>(p, matches, n_matches);
  return cudaGetLastError();
}

namespace {

// single CTA: exclusive scan of counts[0..n) into offsets[0..n]
__global__ void __launch_bounds__(1024) scan_counts_kernel(const int32_t* __restrict__ counts, int n,
                                                           int32_t* __restrict__ offsets) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < n ? counts[i] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += t;
    }
    if (lane == 31) warp_tot[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      warp_tot[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int before = carry + (warp ? warp_tot[warp - 1] : 0) + x - v;
    if (i < n) offsets[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) offsets[n] = carry;
}

__global__ void __launch_bounds__(256) pack_matches_kernel(const uint32_t* __restrict__ matches,
                                                           const int32_t* __restrict__ n_matches,
                                                           const int32_t* __restrict__ offsets, int kmax,
                                                           uint32_t* __restrict__ packed) {
  const int pair = blockIdx.x;
  const int n = n_matches[pair];
  const uint2* src = reinterpret_cast<const uint2*>(matches) + (size_t)pair * kmax;
  uint2* dst = reinterpret_cast<uint2*>(packed) + offsets[pair];
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}

}  // namespace

cudaError_t launch_compact_matches(const uint32_t* matches, const int32_t* n_matches, int n_pairs, int kmax,
                                   int32_t* offsets, uint32_t* packed, cudaStream_t stream) {
  if (n_pairs <= 0) return cudaSuccess;
  scan_counts_kernel<<<1, 1024, 0, stream>>>(n_matches, n_pairs, offsets);
  pack_matches_kernel<<<n_pairs, 256, 0, stream>>>(matches, n_matches, offsets, kmax, packed);
  return cudaGetLastError();
}

cudaError_t launch_u8_to_bf16(const uint8_t* src, void* dst, size_t n, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const size_t threads = (n + 15) / 16;
  u8_to_bf16_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  return cudaGetLastError();
}

cudaError_t launch_f32_to_bf16(const float* src, void* dst, size_t n, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const size_t threads = (n + 7) / 8;
  f32_to_bf16_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  return cudaGetLastError();
}

cudaError_t launch_l2_max_norm2(const uint8_t* src, size_t rows, int d, int* out, cudaStream_t stream) {
  cudaError_t e = cudaMemsetAsync(out, 0, sizeof(int), stream);
  if (e != cudaSuccess || rows == 0) return e;
  l2_max_norm2_kernel<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, stream>>>(src, rows, d, out);
  return cudaGetLastError();
}

cudaError_t launch_l2_augment(const uint8_t* src, size_t rows, int d, float offset_k, void* dst,
                              cudaStream_t stream) {
  if (rows == 0) return cudaSuccess;
  l2_augment_kernel<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, stream>>>(
      src, rows, d, offset_k, reinterpret_cast<__nv_bfloat16*>(dst));
  return cudaGetLastError();
}

}  // namespace vc
