// Internal interface between the C ABI (vc_api.cu) and the matcher kernels (match_kernel.cu).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace vc {

// Two-view geometry of one pair for guided matching (device copy of vc_pair_geometry).
struct GuidedGeometry {
  float h[9];   // row-major homography, image 1 -> image 2
  float f[9];   // row-major fundamental matrix, x2^T F x1
  int mode;     // 0: SiftGPU rule (both tests, '<' keeps); 1: F only ('>' filters); 2: H only ('>' filters)
  float max_h, max_f;  // squared residual bounds
};

// Epilogue warps per TMEM lane quarter in the single-pass kernels (match_fused.cu, match_fused2.cu): each owns
// 8 / kMatchEpiGroups (rounded up) of a tile's eight 32-column chunks and writes its own row partial, so the
// row-partial buffers are sized with it (vc_api.cu part_bytes).  4 = two chunks per warp, no idle third group.
#ifndef VC_MATCH_EPI_GROUPS
#define VC_MATCH_EPI_GROUPS 4
#endif
constexpr int kMatchEpiGroups = VC_MATCH_EPI_GROUPS;
constexpr int kMatchChunksPerGroup = (8 + kMatchEpiGroups - 1) / kMatchEpiGroups;

struct MatchParams {
  const int32_t* counts;  // [n_images] valid descriptors per image
  const int32_t* pairs;   // [n_pairs, 2] image indices
  int n_pairs;
  int kmax;               // padded descriptors per image (tensor-map extent)
  int d;                  // descriptor dimension
  int rb_per_img;         // ceil(kmax / 128)
  int n_items;            // n_pairs * 2 * rb_per_img
  float inv_norm, max_ratio, max_distance;
  int rule;               // 0 angular (sift.cc:733-787), 1 squared-L2 (sift.cc:827-874)
  float l2_offset;        // rule 1: score = l2_offset - squared distance
  int cross_check;
  int32_t* m_out;         // [n_pairs, 2, kmax] accepted best index or -1
  float* top_best;        // optional diagnostics [n_pairs, 2, kmax]
  float* top_second;
  int32_t* top_idx;
  float* dbg_scores;      // optional dense scores of pair 0 (debug kernel only)
  int dbg_ld;
  // single-pass kernel (match_fused.cu) only
  int cb_per_img;         // ceil(kmax / 256)
  unsigned long long* col_key;  // two-SM kernel: [n_pairs, 2 (cta rank), kmax] column states
  unsigned int* col_sec;
  float prune;            // raw-score bound below which a score can change no accept / reject decision (0 = off)
  int dense_min;          // hit columns per 32x32 chunk from which the column merge takes the warp-reduction path
  int stagger;            // row-tile offset per column block (0 = every CTA of a pair sweeps in the same order)
  unsigned int* rowthr;   // [n_pairs, kmax] float bits: lower bound of each row's final second-best
  float* part_best;       // [n_pairs, 2*cb_per_img, kmax] row partials, one per 128 columns
  float* part_second;
  int32_t* part_idx;
  // guided matching (match_top2_kernel<.., true> only)
  const float* kp;              // [kp_images, kmax, 2] keypoint (x, y) per descriptor row
  int kp_images;                // image index modulo this = keypoint image (rule l2 doubles the image axis)
  const GuidedGeometry* geom;   // [n_pairs]
  float masked_score;           // score of a geometrically inconsistent candidate (0, or l2_offset - 512^2)
};

// Geometric consistency of a candidate (x1, y1) <-> (x2, y2).  Every operation is a separately rounded
// fp32 multiply / add / divide (no FMA contraction), so that the numpy oracle reproduces it bit for bit.
// mode 0: SIFTGPU/ProgramCU.cu:1573-1600 (homography test, then Sampson-style epipolar test, '<' keeps);
// mode 1 / 2: COLMAP/feature/sift.cc:1099-1122 (one test, '>' filters).
__device__ __forceinline__ bool guided_keep(const GuidedGeometry& g, float x1, float y1, float x2, float y2) {
  bool keep = true;
  if (g.mode != 1) {
    const float* H = g.h;
    const float a = __fadd_rn(__fadd_rn(__fmul_rn(H[0], x1), __fmul_rn(H[1], y1)), H[2]);
    const float b = __fadd_rn(__fadd_rn(__fmul_rn(H[3], x1), __fmul_rn(H[4], y1)), H[5]);
    const float c = __fadd_rn(__fadd_rn(__fmul_rn(H[6], x1), __fmul_rn(H[7], y1)), H[8]);
    const float dx = __fsub_rn(__fdiv_rn(a, c), x2), dy = __fsub_rn(__fdiv_rn(b, c), y2);
    const float hd = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    keep = g.mode == 0 ? hd < g.max_h : !(hd > g.max_h);
  }
  if (g.mode != 2 && keep) {
    const float* F = g.f;
    const float fx0 = __fadd_rn(__fadd_rn(__fmul_rn(F[0], x1), __fmul_rn(F[1], y1)), F[2]);
    const float fx1 = __fadd_rn(__fadd_rn(__fmul_rn(F[3], x1), __fmul_rn(F[4], y1)), F[5]);
    const float fx2 = __fadd_rn(__fadd_rn(__fmul_rn(F[6], x1), __fmul_rn(F[7], y1)), F[8]);
    const float ft0 = __fadd_rn(__fadd_rn(__fmul_rn(F[0], x2), __fmul_rn(F[3], y2)), F[6]);
    const float ft1 = __fadd_rn(__fadd_rn(__fmul_rn(F[1], x2), __fmul_rn(F[4], y2)), F[7]);
    const float e = __fadd_rn(__fadd_rn(__fmul_rn(x2, fx0), __fmul_rn(y2, fx1)), fx2);
    const float den = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(fx0, fx0), __fmul_rn(fx1, fx1)), __fmul_rn(ft0, ft0)),
                                __fmul_rn(ft1, ft1));
    const float se = __fdiv_rn(__fmul_rn(e, e), den);
    keep = g.mode == 0 ? se < g.max_f : !(se > g.max_f);
  }
  return keep;
}

// Acceptance tests on a row's (best, second) scores.  acos is evaluated in double and rounded once so
// the float result agrees with a correctly rounded host acosf.
__device__ __forceinline__ int accept_scores(float best, float second, int idx, const MatchParams& p) {
  if (idx < 0) return -1;
  if (p.rule == 1) {
    // COLMAP/feature/sift.cc:832,859-871: squared-L2 threshold, ratio test on the distances
    const float d1 = p.l2_offset - best, d2 = p.l2_offset - second;
    if (d1 > p.max_distance * p.max_distance / p.inv_norm) return -1;
    if (sqrtf(fmaxf(d1, 0.f)) >= p.max_ratio * sqrtf(fmaxf(d2, 0.f))) return -1;
    return idx;
  }
  // COLMAP/feature/sift.cc:759-783
  const float t1 = (float)acos((double)fminf(p.inv_norm * best, 1.0f));
  if (t1 > p.max_distance) return -1;
  const float t2 = (float)acos((double)fminf(p.inv_norm * second, 1.0f));
  if (t1 >= p.max_ratio * t2) return -1;
  return idx;
}

// Fast form of the same decision.  Per point, the products with H and F are formed once
// (image-1 point: H p1 and the bound max_h * c^2, F p1 and |F p1|_xy^2; image-2 point: |F^T p2|_xy^2), so a
// candidate costs a handful of FMAs and no division:
//   homography : (a - x2 c)^2 + (b - y2 c)^2   against  max_h * c^2
//   epipolar   : (x2 fx0 + y2 fx1 + fx2)^2     against  max_f * (|F p1|_xy^2 + |F^T p2|_xy^2)
// These differ from guided_keep's separately rounded operations by rounding only, so the fast form
// decides just the candidates whose residual is outside a +-1/64 band around the bound (far more than the
// rounding differences for pixel coordinates and max_error >= 0.25 px); it returns -1 inside the band
// (and for degenerate denominators) and the caller falls back to guided_keep.  Decisions are therefore
// those of guided_keep, at a fraction of the cost.
struct GuidedA {  // image-1 point
  float a, b, c, hb, f0, f1, f2, n1;
};
struct GuidedB {  // image-2 point
  float x, y, n2;
};
__device__ __forceinline__ GuidedA guided_features_a(const GuidedGeometry& g, float x1, float y1) {
  GuidedA A;
  A.a = fmaf(g.h[0], x1, fmaf(g.h[1], y1, g.h[2]));
  A.b = fmaf(g.h[3], x1, fmaf(g.h[4], y1, g.h[5]));
  A.c = fmaf(g.h[6], x1, fmaf(g.h[7], y1, g.h[8]));
  A.hb = g.max_h * A.c * A.c;
  A.f0 = fmaf(g.f[0], x1, fmaf(g.f[1], y1, g.f[2]));
  A.f1 = fmaf(g.f[3], x1, fmaf(g.f[4], y1, g.f[5]));
  A.f2 = fmaf(g.f[6], x1, fmaf(g.f[7], y1, g.f[8]));
  A.n1 = fmaf(A.f0, A.f0, A.f1 * A.f1);
  return A;
}
__device__ __forceinline__ GuidedB guided_features_b(const GuidedGeometry& g, float x2, float y2) {
  GuidedB B;
  B.x = x2;
  B.y = y2;
  const float t0 = fmaf(g.f[0], x2, fmaf(g.f[3], y2, g.f[6]));
  const float t1 = fmaf(g.f[1], x2, fmaf(g.f[4], y2, g.f[7]));
  B.n2 = fmaf(t0, t0, t1 * t1);
  return B;
}
// One candidate, branch-free: `drop` = one test certainly fails (dropped under every mode), `undecided` =
// no test certainly fails but one is inside the band.  kH / kF: which tests the pair's mode asks for
// (mode 0 with the homography absent -- identity, bound 1e20, SIFTGPU/SiftMatch.cpp:684 -- needs no H test:
// it passes for every finite keypoint pair).
template <bool kH, bool kF>
__device__ __forceinline__ void guided_fast(float max_f, const GuidedA& A, const GuidedB& B, bool& drop, bool& undecided) {
  constexpr float kLo = 1.f - 1.f / 64.f, kHi = 1.f + 1.f / 64.f;
  bool below = true, above = false;
  if (kH) {
    const float u = fmaf(-B.x, A.c, A.a), w = fmaf(-B.y, A.c, A.b);
    const float r = fmaf(u, u, w * w);
    below = r < A.hb * kLo;
    above = r > A.hb * kHi;
  }
  if (kF) {
    const float e = fmaf(B.x, A.f0, fmaf(B.y, A.f1, A.f2));
    const float r = e * e, bound = max_f * (A.n1 + B.n2);
    below = below & (r < bound * kLo);
    above = above | (r > bound * kHi);
  }
  drop = above;
  undecided = !above & !below;
}

// The gate over one 32-column chunk of a row: entry e of `feat` holds the staged products of column e's
// keypoint (direction 0: image-2 point {x, y, n2, -}; direction 1: image-1 point {a, b, c, hb}, {f0, f1, f2, n1}).
// Dropped candidates get `masked`; returns the bit mask of candidates the exact form has to decide.
template <bool kDir1, bool kH, bool kF>
__device__ __forceinline__ unsigned guided_gate_chunk(const float4* feat, const GuidedA& row_a, const GuidedB& row_b,
                                                      float max_f, float masked, int n_live, float (&v)[32]) {
  unsigned undecided = 0u;
#pragma unroll
  for (int e = 0; e < 32; ++e) {
    GuidedA A = row_a;
    GuidedB B = row_b;
    const float4 lo = feat[2 * e];
    if (!kDir1) {
      B.x = lo.x;
      B.y = lo.y;
      B.n2 = lo.z;
    } else {
      if (kH) {
        A.a = lo.x;
        A.b = lo.y;
        A.c = lo.z;
        A.hb = lo.w;
      }
      if (kF) {
        const float4 hi = feat[2 * e + 1];
        A.f0 = hi.x;
        A.f1 = hi.y;
        A.f2 = hi.z;
        A.n1 = hi.w;
      }
    }
    bool drop, und;
    guided_fast<kH, kF>(max_f, A, B, drop, und);
    const bool live = e < n_live;  // padding columns keep their score of 0
    v[e] = (live & drop) ? masked : v[e];
    undecided |= (live & und) ? (1u << e) : 0u;
  }
  return undecided;
}

int match_smem_bytes();
cudaError_t launch_match_top2(const CUtensorMap& tmap, const MatchParams& p, int num_sms, bool debug,
                              cudaStream_t stream);
// same kernel with the geometric gate of guided matching in the epilogue (p.kp, p.geom set)
cudaError_t launch_match_top2_guided(const CUtensorMap& tmap, const MatchParams& p, int num_sms,
                                     cudaStream_t stream);
int match_fused_smem_bytes();
// memset(rowthr) + fused kernel + row-partial merge; n_items must be n_pairs * cb_per_img
cudaError_t launch_match_fused(const CUtensorMap& tmap, const MatchParams& p, int num_sms,
                               cudaStream_t stream);
// two-SM (cta_group::2) variant, match_fused2.cu: + merge of the two CTAs' column states
cudaError_t launch_match_fused2(const CUtensorMap& tmap, const MatchParams& p, int num_sms,
                                cudaStream_t stream);
cudaError_t launch_emit_matches(const MatchParams& p, uint32_t* matches, int32_t* n_matches,
                                cudaStream_t stream);
// Packs the filled prefix of every pair's [kmax, 2] slot into one dense [total, 2] list (offsets =
// exclusive scan of n_matches, written to offsets[0..n_pairs]) so that the host entry point reads
// the matches of a pair group back with a single copy.
cudaError_t launch_compact_matches(const uint32_t* matches, const int32_t* n_matches, int n_pairs, int kmax,
                                   int32_t* offsets, uint32_t* packed, cudaStream_t stream);
cudaError_t launch_u8_to_bf16(const uint8_t* src, void* dst, size_t n, cudaStream_t stream);
cudaError_t launch_f32_to_bf16(const float* src, void* dst, size_t n, cudaStream_t stream);
// rule VC_RULE_L2: max_i |row_i|^2 -> *out (device int), and the augmented bf16 rows [2*rows, d+8]
cudaError_t launch_l2_max_norm2(const uint8_t* src, size_t rows, int d, int* out, cudaStream_t stream);
cudaError_t launch_l2_augment(const uint8_t* src, size_t rows, int d, float offset_k, void* dst,
                              cudaStream_t stream);

}  // namespace vc
