// Multi-head self-attention of the ViT blocks (head_dim 64) on tcgen05: softmax(Q K^T / 8) V per
// (image, head), straight from the packed QKV GEMM output, without materialising the T x T scores.
//
// CTA = (image, head, pair of 128-row query tiles).  Roles:
//   warp 0  TMA: both Q tiles once, then a 3-stage ring of (K_j, V_j) 128-key tiles (128B swizzle)
//   warp 1  tcgen05.mma issuer: S_w = Q_w K_j^T (M128 N128 K64) into TMEM, and, once the softmax
//           warps have published P_w (bf16, written to shared memory in the K-major swizzled
//           operand layout), O_w^(j) = P_w V_j (M128 N64 K128; V is consumed as an MN-major operand
//           exactly as TMA delivered it, so no transposed copy of V exists anywhere)
//           and allocates TMEM (S_0, S_1: 2 x 128 columns, O_0, O_1: 2 x 64 columns).  O accumulates
//           in TMEM across the key tiles.
//   warps 2-5 / 6-9  softmax warpgroup of query tile 0 / 1: thread = one query row.  The 128 scores
//           of a tile are pulled into registers with one round of tcgen05.ld (TMEM latency is paid
//           once per tile), online softmax in the exp2 domain with a lazily updated reference
//           maximum: O and the row sum are only rescaled when the maximum grew by more than 2^8
//           (probabilities stay <= 256, exact in the final normalisation), so the TMEM
//           read-modify-write of O is rare.
// The two warpgroups share every K/V tile and keep the tensor core busy for each other.
#include "vit_kernels.cuh"

#include <math.h>
#include <stdlib.h>

#include "vc_ptx.cuh"

namespace vc {

namespace {

constexpr int kTile = 128, kHd = 64, kKvStages = 3;
constexpr int kQBytes = kTile * kHd * 2;          // 16 KB
constexpr int kKvBytes = 2 * kTile * kHd * 2;     // K + V of one stage, 32 KB
constexpr int kPBytes = kTile * kTile * 2;        // 32 KB per warpgroup
constexpr int kThreads = 320;  // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2-9 softmax
constexpr int kTmemCols = 512;
constexpr int kSmemBytes = 2 * kQBytes + kKvStages * kKvBytes + 2 * kPBytes + 1024 + 256;

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


// One 128-key tile of the online softmax for one query row.  kMasked is only instantiated for the
// last, partial key tile, so full tiles carry no per-element selects.
template <bool kMasked>
__device__ __forceinline__ void softmax_tile(uint32_t t_s, uint32_t t_o, uint8_t* prow, int r, int valid,
                                             float sc, bool have_o, float& m_ref, float& l_run) {
  // Software pipeline over TMEM: every tcgen05.ld is issued one step ahead of the arithmetic that
  // consumes it (tcgen05.wait::ld waits for all outstanding loads, so the order is ld-next,
  // compute-current, wait).  Pass 1 (maximum) reads the tile as two 64-column halves, pass 2
  // (probabilities) as four 32-column chunks.
  auto mask = [&](uint32_t (&x)[32], int col0) {
    if (kMasked) {
#pragma unroll
      for (int e = 0; e < 32; ++e)
        if (col0 + e >= valid) x[e] = 0xFF800000u;  // -inf
    }
  };
  float mx[4] = {-1e30f, -1e30f, -1e30f, -1e30f};
  auto fold_max = [&](const uint32_t (&x)[32]) {
#pragma unroll
    for (int e = 0; e < 32; e += 8) {
#pragma unroll
      for (int a = 0; a < 4; ++a)
        mx[a] = fmax3(mx[a], __uint_as_float(x[e + 2 * a]), __uint_as_float(x[e + 2 * a + 1]));
    }
  };
  uint32_t a0[32], a1[32], b0[32], b1[32];
  tmem_ld_32x32(t_s, a0);
  tmem_ld_32x32(t_s + 32, a1);
  tmem_wait_ld();
  tmem_ld_32x32(t_s + 64, b0);
  tmem_ld_32x32(t_s + 96, b1);
  mask(a0, 0);
  mask(a1, 32);
  fold_max(a0);
  fold_max(a1);
  tmem_wait_ld();
  tmem_ld_32x32(t_s, a0);  // first chunk of pass 2, in flight during the rest of pass 1
  mask(b0, 64);
  mask(b1, 96);
  fold_max(b0);
  fold_max(b1);
  const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * sc;
  // lazy reference maximum: move it (and rescale O, l) only when it is more than 2^8 behind
  if (__any_sync(0xffffffffu, m_tile > m_ref + 8.0f)) {
    const float m_new = fmaxf(m_ref, m_tile);
    const float alpha = ex2(m_ref - m_new);
    m_ref = m_new;
    l_run *= alpha;
    if (have_o) {
      const float2 a2 = make_float2(alpha, alpha);
      tmem_ld_32x32(t_o, b0);
      tmem_ld_32x32(t_o + 32, b1);
      tmem_wait_ld();
#pragma unroll
      for (int e = 0; e < 32; e += 2) {
        float2 t = fmul2(make_float2(__uint_as_float(b0[e]), __uint_as_float(b0[e + 1])), a2);
        b0[e] = __float_as_uint(t.x);
        b0[e + 1] = __float_as_uint(t.y);
        t = fmul2(make_float2(__uint_as_float(b1[e]), __uint_as_float(b1[e + 1])), a2);
        b1[e] = __float_as_uint(t.x);
        b1[e + 1] = __float_as_uint(t.y);
      }
      tmem_st_32x32(t_o, b0);
      tmem_st_32x32(t_o + 32, b1);
      tmem_wait_st();
    }
  }
  const float2 sc2 = make_float2(sc, sc), nm2 = make_float2(-m_ref, -m_ref);
  float2 sum[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
  auto emit = [&](uint32_t (&x)[32], int c) {
    mask(x, c * 32);
    uint32_t pk[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      float2 y = ffma2(make_float2(__uint_as_float(x[2 * e]), __uint_as_float(x[2 * e + 1])), sc2, nm2);
      y.x = ex2(y.x);  // masked keys: exp2(-inf) = 0
      y.y = ex2(y.y);
      sum[e & 1] = fadd2(sum[e & 1], y);
      pk[e] = pack_bf16(y.x, y.y);
    }
    // chunk c covers keys [32c, 32c+32): half c>>1, 16-byte chunks (c&1)*4 .. +3 of the swizzled row
    uint8_t* half = prow + (c >> 1) * (kPBytes / 2);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int chunk = ((c & 1) * 4 + g) ^ (r & 7);
      *reinterpret_cast<uint4*>(half + chunk * 16) =
          make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
    }
  };
  tmem_wait_ld();
  tmem_ld_32x32(t_s + 32, a1);
  emit(a0, 0);
  tmem_wait_ld();
  tmem_ld_32x32(t_s + 64, a0);
  emit(a1, 1);
  tmem_wait_ld();
  tmem_ld_32x32(t_s + 96, a1);
  emit(a0, 2);
  tmem_wait_ld();
  emit(a1, 3);
  l_run += (sum[0].x + sum[0].y) + (sum[1].x + sum[1].y);
}

}  // namespace

__global__ void __launch_bounds__(kThreads, 1)
vit_attention_kernel(const __grid_constant__ CUtensorMap map_qkv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_addr);
  const uint32_t sQ = base;                                    // 2 x 16 KB
  const uint32_t sKV = sQ + 2 * kQBytes;                       // stages x (K 16 KB, V 16 KB)
  const uint32_t sP = sKV + kKvStages * kKvBytes;              // 2 x 32 KB
  uint8_t* sP_gen = smem + 2 * kQBytes + kKvStages * kKvBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * kQBytes + kKvStages * kKvBytes + 2 * kPBytes);
  const uint32_t bar0 = smem_u32(bars);
  const uint32_t q_full = bar0;
  auto kv_full = [&](int s) { return bar0 + 8u * (1 + s); };
  auto kv_empty = [&](int s) { return bar0 + 8u * (1 + kKvStages + s); };
  auto s_full = [&](int w) { return bar0 + 8u * (1 + 2 * kKvStages + w); };
  auto p_full = [&](int w) { return bar0 + 8u * (3 + 2 * kKvStages + w); };
  auto o_full = [&](int w) { return bar0 + 8u * (5 + 2 * kKvStages + w); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 7 + 2 * kKvStages);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = p.heads * kHd;
  const int q_pairs = (p.T + 2 * kTile - 1) / (2 * kTile);
  const int qp = blockIdx.x % q_pairs;
  const int bh = blockIdx.x / q_pairs;
  const int head = bh % p.heads, img = bh / p.heads;
  const int q0 = qp * 2 * kTile;
  const int nw = (q0 + kTile < p.T) ? 2 : 1;
  const int nkv = (p.T + kTile - 1) / kTile;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_qkv);
    mbar_init(q_full, 1);
    for (int s = 0; s < kKvStages; ++s) {
      mbar_init(kv_full(s), 1);
      mbar_init(kv_empty(s), 1);
    }
    for (int w = 0; w < 2; ++w) {
      mbar_init(s_full(w), 1);
      mbar_init(p_full(w), 128);
      mbar_init(o_full(w), 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer
    mbar_arrive_expect_tx(q_full, 2 * kQBytes);
    tma_load_3d(sQ, &map_qkv, q_full, head * kHd, q0, img);
    tma_load_3d(sQ + kQBytes, &map_qkv, q_full, head * kHd, q0 + kTile, img);
    for (int j = 0; j < nkv; ++j) {
      const int s = j % kKvStages;
      const uint32_t ph = (j / kKvStages) & 1;
      mbar_wait_sleep(kv_empty(s), ph ^ 1, 256);
      mbar_arrive_expect_tx(kv_full(s), kKvBytes);
      tma_load_3d(sKV + s * kKvBytes, &map_qkv, kv_full(s), C + head * kHd, j * kTile, img);
      tma_load_3d(sKV + s * kKvBytes + kKvBytes / 2, &map_qkv, kv_full(s), 2 * C + head * kHd, j * kTile,
                  img);
    }
  } else if (warp == 1 && lane == 0) {
    // ------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128);
    constexpr uint32_t idesc_o = idesc_pv();
    auto issue_s = [&](int w, int s) {
      const uint64_t da = umma_desc_k_sw128(sQ + w * kQBytes);
      const uint64_t db = umma_desc_k_sw128(sKV + s * kKvBytes);
#pragma unroll
      for (int k = 0; k < kHd / 16; ++k)
        tc_mma_bf16(tmem_base + w * 128, da + 2 * k, db + 2 * k, idesc_s, k != 0);
      tc_commit(s_full(w));
    };
    mbar_wait(q_full, 0);
    mbar_wait(kv_full(0), 0);
    tc_fence_after();
    for (int w = 0; w < nw; ++w) issue_s(w, 0);
    for (int j = 0; j < nkv; ++j) {
      const int s = j % kKvStages;
      const bool more = j + 1 < nkv;
      const int s1 = (j + 1) % kKvStages;
      if (more) {
        mbar_wait(kv_full(s1), ((j + 1) / kKvStages) & 1);
        tc_fence_after();
      }
      for (int w = 0; w < nw; ++w) {
        mbar_wait_sleep(p_full(w), j & 1, 32);
        tc_fence_after();
        // O_w += P_w (128 x 128, two 64-key halves, K-major) * V_j (128 keys x 64, MN-major)
        const uint32_t pw = sP + w * kPBytes;
        const uint64_t dv = umma_desc_k_sw128(sKV + s * kKvBytes + kKvBytes / 2);
#pragma unroll
        for (int k = 0; k < kTile / 16; ++k) {
          const uint64_t dp = umma_desc_k_sw128(pw + (k >> 2) * (kPBytes / 2)) + 2 * (k & 3);
          tc_mma_bf16(tmem_base + 256 + w * 64, dp, dv + 128 * k, idesc_o, (j | k) != 0);
        }
        if (!more) tc_commit(o_full(w));
        if (more) issue_s(w, s1);   // its commit also covers the P V above: S_w ready => O_w up to date
      }
      tc_commit(kv_empty(s));
    }
  } else if (warp >= 2 && (warp - 2) / 4 < nw) {
    // ------------------------------------------------------------ softmax warpgroups
    const int w = (warp - 2) >> 2;
    const int quarter = warp & 3;       // TMEM lane quarter a warp may touch = warp id % 4
    const int r = quarter * 32 + lane;  // row inside the query tile == TMEM lane
    const uint32_t t_s = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + w * 128;
    const uint32_t t_o = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + 256 + w * 64;
    const float sc = p.scale * 1.4426950408889634f;  // scores are handled in the exp2 domain
    uint8_t* prow = sP_gen + w * kPBytes + r * 128;
    float m_ref = -1e30f, l_run = 0.f;
    for (int j = 0; j < nkv; ++j) {
      const int valid = min(kTile, p.T - j * kTile);  // keys of this tile inside the image
      mbar_wait_sleep(s_full(w), j & 1, 32);
      if (p.alternate && nw == 2) {
        // optional turn-taking: one warpgroup runs its softmax while the tensor core serves the other
        if (w == 1) mbar_wait_sleep(p_full(0), j & 1, 64);
        else if (j > 0) mbar_wait_sleep(p_full(1), (j - 1) & 1, 64);
      }
      tc_fence_after();
      if (valid == kTile)
        softmax_tile<false>(t_s, t_o, prow, r, valid, sc, j > 0, m_ref, l_run);
      else
        softmax_tile<true>(t_s, t_o, prow, r, valid, sc, j > 0, m_ref, l_run);
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_full(w));
    }
    mbar_wait_sleep(o_full(w), 0, 32);
    tc_fence_after();
    const int t = q0 + w * kTile + r;
    const float inv = 1.0f / l_run;
    uint4* dst = reinterpret_cast<uint4*>(p.out + ((size_t)img * p.T + t) * C + head * kHd);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t ov[32];
      tmem_ld_32x32(t_o + c * 32, ov);
      tmem_wait_ld();
      if (t < p.T) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 pk;
          pk.x = pack_bf16(__uint_as_float(ov[8 * g + 0]) * inv, __uint_as_float(ov[8 * g + 1]) * inv);
          pk.y = pack_bf16(__uint_as_float(ov[8 * g + 2]) * inv, __uint_as_float(ov[8 * g + 3]) * inv);
          pk.z = pack_bf16(__uint_as_float(ov[8 * g + 4]) * inv, __uint_as_float(ov[8 * g + 5]) * inv);
          pk.w = pack_bf16(__uint_as_float(ov[8 * g + 6]) * inv, __uint_as_float(ov[8 * g + 7]) * inv);
          dst[c * 4 + g] = pk;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

int attn_smem_bytes() { return kSmemBytes; }

cudaError_t launch_attention(const CUtensorMap& map_qkv, const AttnParams& p, cudaStream_t stream) {
  static bool configured = false;
  static int alternate = 0;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(vit_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kSmemBytes);
    if (e != cudaSuccess) return e;
    const char* env = getenv("VC_ATTN_ALTERNATE");  // tuning knob, see DESIGN.md
    if (env) alternate = atoi(env);
    configured = true;
  }
  if (p.B <= 0 || p.T <= 0) return cudaSuccess;
  AttnParams q = p;
  q.alternate = alternate;
  const int q_pairs = (p.T + 2 * kTile - 1) / (2 * kTile);
  vit_attention_kernel<<<p.B * p.heads * q_pairs, kThreads, kSmemBytes, stream>>>(map_qkv, q);
  return cudaGetLastError();
}

}  // namespace vc
