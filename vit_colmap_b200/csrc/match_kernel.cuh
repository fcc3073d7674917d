// Internal interface between the C ABI (vc_api.cu) and the matcher kernels (match_kernel.cu).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace vc {

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
  int stagger;            // row-tile offset per column block (0 = every CTA of a pair sweeps in the same order)
  unsigned int* rowthr;   // [n_pairs, kmax] float bits: lower bound of each row's final second-best
  float* part_best;       // [n_pairs, 2*cb_per_img, kmax] row partials, one per 128 columns
  float* part_second;
  int32_t* part_idx;
};

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

int match_smem_bytes();
cudaError_t launch_match_top2(const CUtensorMap& tmap, const MatchParams& p, int num_sms, bool debug,
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
