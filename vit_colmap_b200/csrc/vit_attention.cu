// Multi-head self-attention of the ViT blocks (head_dim 64) on tcgen05: softmax(Q K^T / 8) V per
// (image, head), straight from the packed QKV GEMM output, without materialising the T x T scores.
//
// Work item = (image, head, 128-row query tile).  A persistent CTA (one per SM) is two independent
// halves of 320 threads, each walking its own list of items with its own shared memory, barriers and
// TMEM columns.  Roles inside a half:
//   warp 0  TMA: the Q tile of the item, then a 2-stage ring of K_j and a 2-stage ring of V_j tiles
//           (128 keys x 64, 128B swizzle).  K and V get separate rings because their lifetimes differ:
//           K_j is free as soon as Q K_j^T has completed, V_j only after P V_j.
//   warp 1  tcgen05.mma issuer: S = Q K_j^T (M128 N128 K64) into TMEM as soon as the softmax threads
//           have pulled the previous S into registers (so Q K^T of tile j+1 -- or of the first key
//           tile of the next item -- overlaps the softmax of tile j), and, once a softmax half has
//           published its P (bf16, written to shared memory in the K-major swizzled operand layout),
//           O_h += P_h V_{j,h} (M128 N64 K64; V is consumed as an MN-major operand exactly as TMA
//           delivered it, so no transposed copy of V exists).  TMEM per half: S (128 columns),
//           O_0, O_1 (2 x 64 columns).
//   warps 2-9  softmax warpgroup, 256 threads: thread = one query row x one 64-key half of the
//           tile (warp%4 = TMEM lane quarter, (warp-2)/4 = key half).  The two key halves of a row
//           are independent split-K streams: each keeps its own lazily updated reference maximum,
//           row sum and O accumulator (rescaled only when the maximum grew by more than 2^8, so the
//           TMEM read-modify-write of O is rare) and never synchronises with its sibling until the
//           end of the item, where the two partial results are merged as in split-K flash
//           attention: O = (a0 O_0 + a1 O_1) / (a0 l_0 + a1 l_1), a_h = 2^(m_h - m).
// Two halves per SM put four softmax warps on every SM sub-partition (MUFU / TMEM latencies hidden)
// and are started a few key tiles apart, so the Q load, the split-K merge and the output store of
// one half run under the steady state of the other; inside a half the loads, Q K^T of the next item
// and its first softmax are already in flight while the previous item's output leaves.
//
// Why one CTA and not two co-resident ones: measured on this toolchain (CUDA 12.9, driver 580) the
// hardware does run two TMEM-allocating CTAs on one SM, but cudaOccupancyMaxActiveBlocksPerMultiprocessor
// reports 1 for any kernel containing tcgen05.alloc, so the co-residency is not something to build on;
// two halves in one CTA are the same machine with the schedule under our control.
#include "vit_kernels.cuh"

#include <math.h>

#include "vc_ptx.cuh"
#include "vc_launch.cuh"

namespace vc {

namespace {

constexpr int kTile = 128, kHd = 64;
constexpr int kQBytes = kTile * kHd * 2;          // 16 KB
constexpr int kPBytes = kTile * kTile * 2;        // 32 KB
constexpr int kTmemCols = 512;

// K-major A (bits 15 = 0), MN-major B (bit 16 = 1): B = V tile [keys, 64] with the head dim contiguous
__host__ __device__ constexpr uint32_t idesc_pv() { return umma_idesc_bf16(128, 64) | (1u << 16); }

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Blackwell packed fp32 pairs (FFMA2 / FADD2 / FMUL2) and the 3-input maximum (FMNMX3)
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)),
        "l"(*reinterpret_cast<uint64_t*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}


// One 64-key half tile of the online softmax for one query row (thread = row x key half).  The 64
// scores are pulled into registers with one round of tcgen05.ld and never re-read.  kMasked is only
// instantiated for the last, partial key tile, so full tiles carry no per-element selects.
template <bool kMasked>
__device__ __forceinline__ void softmax_half(uint32_t t_s, uint32_t t_o, uint8_t* prow, int r, int valid,
                                             float sc, bool have_o, uint32_t bar_s_free, uint32_t bar_o_ready,
                                             uint32_t o_parity, float& m_ref, float& l_run) {
  uint32_t a[32], b[32];
  tmem_ld_32x32(t_s, a);
  tmem_ld_32x32(t_s + 32, b);
  tmem_wait_ld();
  // the scores now live in registers: hand S back so Q K^T of the next key tile overlaps this softmax
  tc_fence_before();
  mbar_arrive(bar_s_free);
  if (kMasked) {
#pragma unroll
    for (int e = 0; e < 32; ++e) {
      if (e >= valid) a[e] = 0xFF800000u;  // -inf
      if (32 + e >= valid) b[e] = 0xFF800000u;
    }
  }
  float mx[4] = {-1e30f, -1e30f, -1e30f, -1e30f};
#pragma unroll
  for (int e = 0; e < 32; e += 8) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      mx[q] = fmax3(mx[q], __uint_as_float(a[e + 2 * q]), __uint_as_float(a[e + 2 * q + 1]));
      mx[q] = fmax3(mx[q], __uint_as_float(b[e + 2 * q]), __uint_as_float(b[e + 2 * q + 1]));
    }
  }
  const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * sc;
  if (have_o) {
    // P V of the previous key tile must have completed before P is overwritten / O is rescaled
    mbar_wait_sleep(bar_o_ready, o_parity, 20);
    tc_fence_after();
  }
  // lazy reference maximum: move it (and rescale O, l) only when it is more than 2^8 behind
  if (__any_sync(0xffffffffu, m_tile > m_ref + 8.0f)) {
    const float m_new = fmaxf(m_ref, m_tile);
    const float alpha = ex2(m_ref - m_new);
    m_ref = m_new;
    l_run *= alpha;
    if (have_o) {
      const float2 a2 = make_float2(alpha, alpha);
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t o[16];
        tmem_ld_32x16(t_o + c * 16, o);
        tmem_wait_ld();
#pragma unroll
        for (int e = 0; e < 16; e += 2) {
          const float2 t = fmul2(make_float2(__uint_as_float(o[e]), __uint_as_float(o[e + 1])), a2);
          o[e] = __float_as_uint(t.x);
          o[e + 1] = __float_as_uint(t.y);
        }
        tmem_st_32x16(t_o + c * 16, o);
      }
      tmem_wait_st();
    }
  }
  const float2 sc2 = make_float2(sc, sc), nm2 = make_float2(-m_ref, -m_ref);
  float2 sum[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
  auto emit = [&](uint32_t (&x)[32], int c) {
    // 16 keys at a time keeps the live set small: 8 packed registers, two 16-byte stores
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      uint32_t pk[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int i0 = 16 * q + 2 * e;
        float2 y = ffma2(make_float2(__uint_as_float(x[i0]), __uint_as_float(x[i0 + 1])), sc2, nm2);
        y.x = ex2(y.x);  // masked keys: exp2(-inf) = 0
        y.y = ex2(y.y);
        sum[e & 1] = fadd2(sum[e & 1], y);
        pk[e] = pack_bf16(y.x, y.y);
      }
      // chunk c covers keys [32c, 32c+32) of this half: 16-byte chunks 4c .. 4c+3 of the swizzled row
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const int chunk = (c * 4 + 2 * q + g) ^ (r & 7);
        *reinterpret_cast<uint4*>(prow + chunk * 16) =
            make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
      }
    }
  };
  emit(a, 0);
  emit(b, 1);
  l_run += (sum[0].x + sum[0].y) + (sum[1].x + sum[1].y);
}

}  // namespace

namespace {
constexpr int kHalfThreads = 320;
constexpr int kRing = 2;
constexpr int kTileBytes = kTile * kHd * 2;                             // 16 KB
constexpr int kHalfBytes = kQBytes + 2 * kRing * kTileBytes + kPBytes;  // 112 KB
constexpr int kHalfBars = 24;
#ifndef VC_SKEW
#define VC_SKEW 3
#endif
constexpr int kSmemBytes = 2 * kHalfBytes + 512 + 1024;
static_assert(kSmemBytes <= 232448, "shared memory budget of the two halves");
}  // namespace

// barrier of the 256 softmax threads of one half (ids as immediates: a register id makes ptxas
// reserve all sixteen barriers)
__device__ __forceinline__ void half_sync(int half) {
  if (half == 0)
    asm volatile("bar.sync 1, 256;" ::: "memory");
  else
    asm volatile("bar.sync 2, 256;" ::: "memory");
}

__global__ void __launch_bounds__(2 * kHalfThreads, 1)
vit_attention_kernel(const __grid_constant__ CUtensorMap map_qkv, const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_addr);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int half = warp / 10, lw = warp % 10;
  const uint32_t sQ = base + half * kHalfBytes;
  const uint32_t sK = sQ + kQBytes;
  const uint32_t sV = sK + kRing * kTileBytes;
  const uint32_t sP = sV + kRing * kTileBytes;
  uint8_t* sP_gen = smem + half * kHalfBytes + kQBytes + 2 * kRing * kTileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * kHalfBytes);
  const uint32_t bar0 = smem_u32(bars) + 8u * kHalfBars * half;
  const uint32_t q_full = bar0, q_empty = bar0 + 8u;
  auto k_full = [&](int s) { return bar0 + 8u * (2 + s); };
  auto k_empty = [&](int s) { return bar0 + 8u * (4 + s); };
  auto v_full = [&](int s) { return bar0 + 8u * (6 + s); };
  auto v_empty = [&](int s) { return bar0 + 8u * (8 + s); };
  const uint32_t s_full = bar0 + 8u * 10, s_free = bar0 + 8u * 11, o_full = bar0 + 8u * 12;
  auto p_full = [&](int h) { return bar0 + 8u * (13 + h); };
  auto o_ready = [&](int h) { return bar0 + 8u * (15 + h); };
  const uint32_t skew = smem_u32(bars) + 8u * (2 * kHalfBars);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kHalfBars + 2);

  const int C = p.heads * kHd;
  const int q_tiles = (p.T + kTile - 1) / kTile;
  const int nkv = q_tiles;
  const int total = p.B * p.heads * q_tiles;
  const int first = 2 * blockIdx.x + half, stride = 2 * gridDim.x;
  const int n_items = first < total ? (total - first + stride - 1) / stride : 0;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_qkv);
    for (int hf = 0; hf < 2; ++hf) {
      const uint32_t b = smem_u32(bars) + 8u * kHalfBars * hf;
      for (int i = 0; i < 17; ++i) {
        const uint32_t count = i == 11 ? 256u : (i == 13 || i == 14) ? 128u : 1u;
        mbar_init(b + 8u * i, count);
      }
    }
    mbar_init(skew, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot + half * 256;
  griddep_wait();  // qkv is written by the previous kernel; the set-up above overlaps its tail

  if (lw == 0) {
    if (lane == 0 && n_items > 0) {
      // ---------------------------------------------------------- TMA producer of this half
      if (half == 1) mbar_wait_sleep(skew, 0, 256);
      uint32_t g = 0;
      for (int n = 0; n < n_items; ++n) {
        const int item = first + n * stride;
        const int qt = item % q_tiles, bh = item / q_tiles;
        const int head = bh % p.heads, img = bh / p.heads;
        mbar_wait_sleep(q_empty, (n & 1) ^ 1, 64);  // every Q K^T of the previous item has completed
        mbar_arrive_expect_tx(q_full, kQBytes);
        tma_load_3d(sQ, &map_qkv, q_full, head * kHd, qt * kTile, img);
        for (int j = 0; j < nkv; ++j, ++g) {
          const int s = g % kRing;
          const uint32_t ph = (g / kRing) & 1;
          mbar_wait_sleep(k_empty(s), ph ^ 1, 64);
          mbar_arrive_expect_tx(k_full(s), kTileBytes);
          tma_load_3d(sK + s * kTileBytes, &map_qkv, k_full(s), C + head * kHd, j * kTile, img);
          mbar_wait_sleep(v_empty(s), ph ^ 1, 64);
          mbar_arrive_expect_tx(v_full(s), kTileBytes);
          tma_load_3d(sV + s * kTileBytes, &map_qkv, v_full(s), 2 * C + head * kHd, j * kTile, img);
        }
      }
    }
  } else if (lw == 1) {
    if (lane == 0 && n_items > 0) {
      // ---------------------------------------------------------- MMA issuer of this half
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128);
      constexpr uint32_t idesc_o = idesc_pv();
      const uint64_t dq = umma_desc_k_sw128(sQ);
      auto issue_s = [&](uint32_t g, bool last_of_item) {
        const int s = g % kRing;
        const uint64_t db = umma_desc_k_sw128(sK + s * kTileBytes);
#pragma unroll
        for (int k = 0; k < kHd / 16; ++k) tc_mma_bf16(tmem_base, dq + 2 * k, db + 2 * k, idesc_s, k != 0);
        tc_commit(s_full);
        tc_commit(k_empty(s));
        if (last_of_item) tc_commit(q_empty);
      };
      mbar_wait(q_full, 0);
      mbar_wait(k_full(0), 0);
      tc_fence_after();
      issue_s(0, nkv == 1);
      const int skew_at = min(nkv / 2, VC_SKEW);
      uint32_t g = 0;
      for (int n = 0; n < n_items; ++n) {
        for (int j = 0; j < nkv; ++j, ++g) {
          const bool last_j = j + 1 == nkv;
          if (!last_j || n + 1 < n_items) {
            // Q K^T of the next key tile (or of the first key tile of the next item) as soon as the
            // softmax threads hold S(g) in registers
            const uint32_t g1 = g + 1;
            if (last_j) mbar_wait_sleep(q_full, (n + 1) & 1, 32);
            mbar_wait_sleep(k_full(g1 % kRing), (g1 / kRing) & 1, 32);
            mbar_wait_sleep(s_free, g & 1, 32);
            tc_fence_after();
            issue_s(g1, last_j ? nkv == 1 : j + 2 == nkv);
          }
          const int s = g % kRing;
          mbar_wait_sleep(v_full(s), (g / kRing) & 1, 32);
          const uint64_t dv = umma_desc_k_sw128(sV + s * kTileBytes);
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            mbar_wait_sleep(p_full(hh), g & 1, 32);
            tc_fence_after();
            const uint64_t dp = umma_desc_k_sw128(sP + hh * (kPBytes / 2));
#pragma unroll
            for (int k = 0; k < 4; ++k)
              tc_mma_bf16(tmem_base + 128 + hh * 64, dp + 2 * k, dv + 128 * (4 * hh + k), idesc_o, (j | k) != 0);
            tc_commit(o_ready(hh));
          }
          if (last_j) tc_commit(o_full);
          tc_commit(v_empty(s));
          if (half == 0 && n == 0 && j == skew_at) mbar_arrive(skew);  // start the other half now
        }
      }
    }
  } else {
    // ------------------------------------------------------------ softmax warpgroup of this half
    const int h = (lw - 2) >> 2;   // key half
    const int quarter = warp & 3;  // TMEM lane quarter a warp may touch = warp id % 4
    const int r = quarter * 32 + lane;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    const uint32_t t_s = lane_base + h * 64;
    const uint32_t t_o = lane_base + 128 + h * 64;
    const float sc = p.scale * 1.4426950408889634f;
    uint8_t* prow = sP_gen + h * (kPBytes / 2) + r * 128;
    uint32_t g = 0;
    for (int n = 0; n < n_items; ++n) {
      float m_ref = -1e30f, l_run = 0.f;
      for (int j = 0; j < nkv; ++j, ++g) {
        const int valid = min(kTile / 2, p.T - j * kTile - h * (kTile / 2));
        mbar_wait_sleep(s_full, g & 1, 32);
        tc_fence_after();
        if (valid == kTile / 2)
          softmax_half<false>(t_s, t_o, prow, r, valid, sc, j > 0, s_free, o_ready(h), (g - 1) & 1, m_ref, l_run);
        else
          softmax_half<true>(t_s, t_o, prow, r, valid, sc, j > 0, s_free, o_ready(h), (g - 1) & 1, m_ref, l_run);
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(p_full(h));
      }
      // merge the two key halves of the row (split-K) and store
      mbar_wait_sleep(o_full, n & 1, 32);
      tc_fence_after();
      float2* xch = reinterpret_cast<float2*>(sP_gen + kPBytes / 2);  // free once every P V has completed
      uint8_t* stg = sP_gen + quarter * 4096;
      xch[h * kTile + r] = make_float2(m_ref, l_run);
      half_sync(half);
      const float2 other = xch[(h ^ 1) * kTile + r];
      const float m_all = fmaxf(m_ref, other.x);
      const float a_me = ex2(m_ref - m_all), a_ot = ex2(other.x - m_all);
      const float inv = 1.0f / (l_run * a_me + other.y * a_ot);
      const float c0 = (h == 0 ? a_me : a_ot) * inv, c1 = (h == 0 ? a_ot : a_me) * inv;
      uint32_t o0[32], o1[32];
      tmem_ld_32x32(lane_base + 128 + h * 32, o0);
      tmem_ld_32x32(lane_base + 128 + 64 + h * 32, o1);
      tmem_wait_ld();
#pragma unroll
      for (int gq = 0; gq < 4; ++gq) {
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i0 = 8 * gq + 2 * e;
          pk[e] = pack_bf16(fmaf(__uint_as_float(o0[i0]), c0, __uint_as_float(o1[i0]) * c1),
                            fmaf(__uint_as_float(o0[i0 + 1]), c0, __uint_as_float(o1[i0 + 1]) * c1));
        }
        *reinterpret_cast<uint4*>(stg + lane * 128 + (((h * 4 + gq) ^ (lane & 7)) << 4)) =
            make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
      half_sync(half);
      {
        const int item = first + n * stride;
        const int qt = item % q_tiles, bh = item / q_tiles;
        const int head = bh % p.heads, img = bh / p.heads;
        const int tc = lane & 7;
        __nv_bfloat16* dst = p.out + (size_t)img * p.T * C + head * kHd + tc * 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int rr = 16 * h + (lane >> 3) + 4 * i;
          const int tt = qt * kTile + quarter * 32 + rr;
          if (tt < p.T) {
            const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 128 + ((tc ^ (rr & 7)) << 4));
            *reinterpret_cast<uint4*>(dst + (size_t)tt * C) = v;
          }
        }
      }
      // the staging tile and the exchange slots are P again from the next item's first key tile on
      half_sync(half);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(*tmem_slot, kTmemCols);
}

int attn_smem_bytes() { return kSmemBytes; }

static cudaError_t launch_items(const CUtensorMap& map_qkv, const AttnParams& p, int num_sms, cudaStream_t stream) {
  static SmemOptIn opt_in;
  if (cudaError_t e = ensure_dynamic_smem(vit_attention_kernel, kSmemBytes, opt_in); e != cudaSuccess) return e;
  const int q_tiles = (p.T + kTile - 1) / kTile;
  const int total = p.B * p.heads * q_tiles;
  const int grid = (total + 1) / 2 < num_sms ? (total + 1) / 2 : num_sms;
  return launch_pdl(vit_attention_kernel, dim3(grid), dim3(2 * kHalfThreads), kSmemBytes, stream, map_qkv, p);
}

cudaError_t launch_attention(const CUtensorMap& map_qkv, const AttnParams& p, cudaStream_t stream) {
  if (p.B <= 0 || p.T <= 0) return cudaSuccess;
  static int num_sms = 0;
  if (num_sms == 0) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
  }
  return launch_items(map_qkv, p, num_sms, stream);
}

}  // namespace vc
