// C ABI of libvc_b200 (see include/vc_b200.h).  Host-side glue only: argument checks, TMA
// descriptor creation, workspace carving, launches.  No CPU compute path exists here: without an
// sm_100 device every entry point fails with VC_ERR_NO_DEVICE.
#include "../../include/vc_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "match_kernel.cuh"

#include "vc_internal.cuh"

namespace vc {
thread_local std::string g_err;
thread_local int64_t g_launches = 0;

int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
void add_launches(int n) { g_launches += n; }
}  // namespace vc

using vc::fail;
using vc::g_launches;
using vc::g_err;
using vc::align_up;

namespace {

int make_desc_map(vc_handle* h, const void* desc, int n_images, int kmax, int d) {
  if (h->map_ptr == desc && h->map_n == n_images && h->map_k == kmax && h->map_d == d) return VC_OK;
  if ((reinterpret_cast<uintptr_t>(desc) & 15) != 0)
    return fail(VC_ERR_INVALID, "descriptor base must be 16-byte aligned");
  if (d <= 0 || d % 8 != 0) return fail(VC_ERR_INVALID, "d=%d must be a positive multiple of 8", d);
  // dims fastest-first: (d, kmax, n_images); box = 64 bf16 (128 B, one swizzle row) x 128 rows
  cuuint64_t gdim[3] = {(cuuint64_t)d, (cuuint64_t)kmax, (cuuint64_t)n_images};
  cuuint64_t gstride[2] = {(cuuint64_t)d * 2, (cuuint64_t)d * 2 * (cuuint64_t)kmax};
  cuuint32_t box[3] = {64, 128, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = h->encode(&h->map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(desc), gdim,
                         gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    h->map_ptr = nullptr;
    return fail(VC_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  }
  h->map_ptr = desc;
  h->map_n = n_images;
  h->map_k = kmax;
  h->map_d = d;
  return VC_OK;
}

struct MatchWs {
  int32_t* m_out;
  size_t bytes;
};

size_t m_out_bytes(int n_pairs, int kmax) {
  return align_up((size_t)n_pairs * 2 * (size_t)kmax * sizeof(int32_t), 256);
}
size_t rowthr_bytes(int n_pairs, int kmax) { return align_up((size_t)n_pairs * kmax * 4, 256); }
size_t colstate_bytes(int n_pairs, int kmax) {  // two-SM kernel: u64 key + u32 second per (pair, rank, column)
  return align_up((size_t)n_pairs * 2 * kmax * 8, 256) + align_up((size_t)n_pairs * 2 * kmax * 4, 256);
}
size_t part_bytes(int n_pairs, int kmax) {
  // row partials: one per (column block, epilogue column group)
  return align_up((size_t)n_pairs * vc::kMatchEpiGroups * ((kmax + 255) / 256) * (size_t)kmax * 4, 256);
}

// Kernel choice: the single-pass kernel halves the tensor work; its heavier epilogue used to lose to
// the two-direction kernel below d = 256, but with the pruning bound (match_impl) it wins at every d
// (8192 x 8192, us/pair, dual vs fused: d=64 50.7 vs 23.4, d=128 51.3 vs 24.0, d=256 55.2 vs 28.9).
// VC_MATCH_KERNEL=dual|fused overrides (tests, comparisons).
bool use_fused_kernel(int d) {
  const char* env = getenv("VC_MATCH_KERNEL");
  if (env && strcmp(env, "dual") == 0) return false;
  if (env && strcmp(env, "fused") == 0) return true;
  (void)d;
  return true;
}

int check_match_args(vc_handle* h, const void* desc, const int32_t* counts, int n_images, int kmax,
                     int d, const int32_t* pairs, int n_pairs, const vc_match_options* opt) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!desc || !counts || !opt) return fail(VC_ERR_INVALID, "null pointer argument");
  if (n_pairs > 0 && !pairs) return fail(VC_ERR_INVALID, "null pairs");
  if (n_images <= 0 || kmax <= 0 || n_pairs < 0)
    return fail(VC_ERR_INVALID, "bad sizes n_images=%d kmax=%d n_pairs=%d", n_images, kmax, n_pairs);
  if (d <= 0 || d % 8 != 0) return fail(VC_ERR_INVALID, "d=%d must be a positive multiple of 8", d);
  return VC_OK;
}

// Exact pruning bound (angular rule, match lists only), as a raw score; 0 = no bound.
float prune_bound(const vc_match_options* opt) {
  if (opt->rule != 0 || !(opt->max_ratio > 0.f) || !(opt->inv_norm > 0.f) || getenv("VC_MATCH_NO_PRUNE")) return 0.f;
  const double a = std::max((double)opt->max_distance / (double)opt->max_ratio, (double)opt->max_distance);
  if (a >= 1.5707) return 0.f;
  const double c = cos(a) - 1e-3;
  return c > 0.0 ? (float)(c / (double)opt->inv_norm) : 0.f;
}

vc::MatchParams make_params(const int32_t* counts, const int32_t* pairs, int n_pairs, int kmax, int d,
                            const vc_match_options* opt, int32_t* m_out) {
  vc::MatchParams p{};
  p.counts = counts;
  p.pairs = pairs;
  p.n_pairs = n_pairs;
  p.kmax = kmax;
  p.d = d;
  p.rb_per_img = (kmax + 127) / 128;
  p.n_items = n_pairs * 2 * p.rb_per_img;
  p.inv_norm = opt->inv_norm;
  p.max_ratio = opt->max_ratio;
  p.max_distance = opt->max_distance;
  p.rule = opt->rule;
  p.l2_offset = opt->l2_offset;
  p.cross_check = opt->cross_check;
  p.m_out = m_out;
  return p;
}

}  // namespace

extern "C" {

const char* vc_last_error(void) { return g_err.c_str(); }
const char* vc_version(void) { return "vc_b200 0.1 (sm_100a)"; }

int64_t vc_launch_count(int reset) {
  const int64_t v = g_launches;
  if (reset) g_launches = 0;
  return v;
}

int vc_create(int device, vc_handle** out) {
  if (!out) return fail(VC_ERR_INVALID, "null out");
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n)
    return fail(VC_ERR_NO_DEVICE, "no CUDA device %d (found %d); this library has no CPU fallback",
                device, n);
  cudaDeviceProp prop;
  VC_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(VC_ERR_NO_DEVICE, "device %d is sm_%d%d; kernels are built for sm_100a only", device,
                prop.major, prop.minor);
  VC_CUDA(cudaSetDevice(device));
  vc_handle* h = new vc_handle();
  h->device = device;
  h->num_sms = prop.multiProcessorCount;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || fn == nullptr || qres != cudaDriverEntryPointSuccess) {
    delete h;
    return fail(VC_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  }
  h->encode = reinterpret_cast<vc::EncodeTiledFn>(fn);
  *out = h;
  return VC_OK;
}

int vc_destroy(vc_handle* h) {
  if (!h) return VC_OK;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  if (h->stage) cudaFree(h->stage);
  if (h->pin) cudaFreeHost(h->pin);
  if (h->ln_counters) cudaFree(h->ln_counters);
  for (auto st : h->host_streams)
    if (st) cudaStreamDestroy(st);
  for (auto ev : h->host_events)
    if (ev) cudaEventDestroy(ev);
  delete h;
  return VC_OK;
}

size_t vc_match_workspace_bytes(int n_pairs, int kmax) {
  if (n_pairs <= 0 || kmax <= 0) return 256;
  return m_out_bytes(n_pairs, kmax) + rowthr_bytes(n_pairs, kmax) + 3 * part_bytes(n_pairs, kmax) +
         colstate_bytes(n_pairs, kmax);
}

static int match_impl(vc_handle* h, const void* desc, const int32_t* counts, int n_images, int kmax,
                      int d, const int32_t* pairs, int n_pairs, const vc_match_options* opt,
                      uint32_t* matches, int32_t* n_matches, float* best, float* second,
                      int32_t* index, void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_match_args(h, desc, counts, n_images, kmax, d, pairs, n_pairs, opt);
  if (rc != VC_OK) return rc;
  if (n_pairs == 0) return VC_OK;
  if (!workspace || workspace_bytes < vc_match_workspace_bytes(n_pairs, kmax))
    return fail(VC_ERR_WORKSPACE, "workspace %zu B < required %zu B", workspace_bytes,
                vc_match_workspace_bytes(n_pairs, kmax));
  VC_CUDA(cudaSetDevice(h->device));
  rc = make_desc_map(h, desc, n_images, kmax, d);
  if (rc != VC_OK) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  vc::MatchParams p = make_params(counts, pairs, n_pairs, kmax, d, opt,
                                  reinterpret_cast<int32_t*>(workspace));
  p.top_best = best;
  p.top_second = second;
  p.top_idx = index;
  if (use_fused_kernel(d)) {
    uint8_t* w = static_cast<uint8_t*>(workspace) + m_out_bytes(n_pairs, kmax);
    p.rowthr = reinterpret_cast<unsigned int*>(w);
    w += rowthr_bytes(n_pairs, kmax);
    p.part_best = reinterpret_cast<float*>(w);
    w += part_bytes(n_pairs, kmax);
    p.part_second = reinterpret_cast<float*>(w);
    w += part_bytes(n_pairs, kmax);
    p.part_idx = reinterpret_cast<int32_t*>(w);
    p.cb_per_img = (kmax + 255) / 256;
    p.n_items = n_pairs * p.cb_per_img;
    // Exact pruning bound (angular rule, match lists only): a score s with acos(inv_norm * s) >
    // max_distance / max_ratio can neither be the best of an accepted match (distance test,
    // sift.cc:765-771) nor, as second-best, make the ratio test fail for one (t1 <= max_distance <
    // max_ratio * t2, sift.cc:773-780), so the epilogue never has to track it.  The diagnostics entry
    // (vc_match_pairs_top2) keeps the bound at 0 and returns exact top-2 values for every row.
    p.prune = best == nullptr ? prune_bound(opt) : 0.f;
    static const int stagger = [] {
      const char* e = getenv("VC_MATCH_STAGGER");
      return e ? atoi(e) : 2;
    }();
    p.stagger = stagger;
    // column merge: warp-reduction path from this many hit columns per 32x32 chunk (1 = always, 33 = never;
    // measured best at 6, tools/time_match_dims.py).  Read per call so that tests can force either path.
    {
      const char* e = getenv("VC_MATCH_DENSE_MIN");
      p.dense_min = e ? std::max(1, atoi(e)) : 6;
    }
    w += part_bytes(n_pairs, kmax);
    p.col_key = reinterpret_cast<unsigned long long*>(w);
    p.col_sec = reinterpret_cast<unsigned int*>(w + align_up((size_t)n_pairs * 2 * kmax * 8, 256));
    // two-SM kernel (cta_group::2): a third less L2->SM operand traffic, but its accumulator hand-over
    // spans two CTAs.  Measured on B200 at 8192 x 8192 with the pruning bound above
    // (tools/time_match_dims.py, us/pair, single-CTA vs two-SM): d=384 40.3 vs 41.5, d=768 83.4 vs 79.3,
    // d=1024 138.5 vs 102.2 -> used from d >= 768.  VC_MATCH_2CTA=0/1 forces a choice.
    static const int two_sm_env = [] {
      const char* e = getenv("VC_MATCH_2CTA");
      return e ? atoi(e) : -1;
    }();
    const bool two_sm = two_sm_env >= 0 ? two_sm_env != 0 : d >= 768;
    if (two_sm) {
      VC_CUDA(vc::launch_match_fused2(h->map, p, h->num_sms, s));
      g_launches += 3;
    } else {
      VC_CUDA(vc::launch_match_fused(h->map, p, h->num_sms, s));
      g_launches += 2;
    }
  } else {
    VC_CUDA(vc::launch_match_top2(h->map, p, h->num_sms, false, s));
    ++g_launches;
  }
  if (matches && n_matches) {
    VC_CUDA(vc::launch_emit_matches(p, matches, n_matches, s));
    ++g_launches;
  }
  return VC_OK;
}

int vc_match_pairs(vc_handle* h, const void* desc, const int32_t* counts, int n_images, int kmax,
                   int d, const int32_t* pairs, int n_pairs, const vc_match_options* opt,
                   uint32_t* matches, int32_t* n_matches, void* workspace, size_t workspace_bytes,
                   void* stream) {
  if ((matches == nullptr) != (n_matches == nullptr))
    return fail(VC_ERR_INVALID, "matches and n_matches must both be given or both be NULL");
  return match_impl(h, desc, counts, n_images, kmax, d, pairs, n_pairs, opt, matches, n_matches,
                    nullptr, nullptr, nullptr, workspace, workspace_bytes, stream);
}

int vc_match_pairs_top2(vc_handle* h, const void* desc, const int32_t* counts, int n_images,
                        int kmax, int d, const int32_t* pairs, int n_pairs,
                        const vc_match_options* opt, float* best, float* second, int32_t* index,
                        void* workspace, size_t workspace_bytes, void* stream) {
  if (!best || !second || !index) return fail(VC_ERR_INVALID, "null output pointer");
  return match_impl(h, desc, counts, n_images, kmax, d, pairs, n_pairs, opt, nullptr, nullptr, best,
                    second, index, workspace, workspace_bytes, stream);
}

int vc_match_pairs_guided(vc_handle* h, const void* desc, const float* keypoints, const int32_t* counts,
                          int n_images, int kmax, int d, const int32_t* pairs, const vc_pair_geometry* geometry,
                          int n_pairs, const vc_match_options* opt, uint32_t* matches, int32_t* n_matches,
                          void* workspace, size_t workspace_bytes, void* stream) {
  static_assert(sizeof(vc_pair_geometry) == sizeof(vc::GuidedGeometry), "vc_pair_geometry layout");
  int rc = check_match_args(h, desc, counts, n_images, kmax, d, pairs, n_pairs, opt);
  if (rc != VC_OK) return rc;
  if (!keypoints || !matches || !n_matches || (n_pairs > 0 && !geometry))
    return fail(VC_ERR_INVALID, "null pointer argument");
  if (n_pairs == 0) return VC_OK;
  if (!workspace || workspace_bytes < vc_match_workspace_bytes(n_pairs, kmax))
    return fail(VC_ERR_WORKSPACE, "workspace %zu B < required %zu B", workspace_bytes,
                vc_match_workspace_bytes(n_pairs, kmax));
  VC_CUDA(cudaSetDevice(h->device));
  rc = make_desc_map(h, desc, n_images, kmax, d);
  if (rc != VC_OK) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  vc::MatchParams p = make_params(counts, pairs, n_pairs, kmax, d, opt, reinterpret_cast<int32_t*>(workspace));
  p.kp = keypoints;
  // rule l2 doubles the image axis (vc_l2_augment): copy n + j of the descriptors is image j's keypoints
  p.kp_images = opt->rule == VC_RULE_L2 ? n_images / 2 : n_images;
  p.geom = reinterpret_cast<const vc::GuidedGeometry*>(geometry);
  // a gated candidate scores 0 under the angular rule (never a match: best starts at 0, strict '>') and
  // squared distance 512^2 under the l2 rule (COLMAP/feature/sift.cc:945-946)
  p.masked_score = opt->rule == VC_RULE_L2 ? opt->l2_offset - 262144.f : 0.f;
  p.prune = prune_bound(opt);
  VC_CUDA(vc::launch_match_top2_guided(h->map, p, h->num_sms, s));
  VC_CUDA(vc::launch_emit_matches(p, matches, n_matches, s));
  g_launches += 2;
  return VC_OK;
}

int vc_debug_scores(vc_handle* h, const void* desc, const int32_t* counts, int n_images, int kmax,
                    int d, int image1, int image2, float* scores, int ld, void* stream) {
  vc_match_options opt{0.8f, 0.7f, 1, 1.0f};
  int rc = check_match_args(h, desc, counts, n_images, kmax, d, nullptr, 0, &opt);
  if (rc != VC_OK) return rc;
  if (!scores || ld < kmax || image1 < 0 || image2 < 0 || image1 >= n_images || image2 >= n_images)
    return fail(VC_ERR_INVALID, "bad debug arguments");
  VC_CUDA(cudaSetDevice(h->device));
  rc = make_desc_map(h, desc, n_images, kmax, d);
  if (rc != VC_OK) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int32_t host_pair[2] = {image1, image2};
  int32_t* dev = nullptr;
  const size_t mb = m_out_bytes(1, kmax);
  VC_CUDA(cudaMalloc(&dev, 256 + mb));
  VC_CUDA(cudaMemcpyAsync(dev, host_pair, sizeof(host_pair), cudaMemcpyHostToDevice, s));
  vc::MatchParams p = make_params(counts, dev, 1, kmax, d, &opt, dev + 64);
  p.dbg_scores = scores;
  p.dbg_ld = ld;
  cudaError_t e = vc::launch_match_top2(h->map, p, h->num_sms, true, s);
  ++g_launches;
  cudaError_t e2 = cudaStreamSynchronize(s);
  cudaFree(dev);
  VC_CUDA(e);
  VC_CUDA(e2);
  return VC_OK;
}

int vc_convert_u8_to_bf16(const uint8_t* src, void* dst_bf16, size_t n, void* stream) {
  if (!src || !dst_bf16) return fail(VC_ERR_INVALID, "null pointer");
  VC_CUDA(vc::launch_u8_to_bf16(src, dst_bf16, n, reinterpret_cast<cudaStream_t>(stream)));
  ++g_launches;
  return VC_OK;
}

int vc_convert_f32_to_bf16(const float* src, void* dst_bf16, size_t n, void* stream) {
  if (!src || !dst_bf16) return fail(VC_ERR_INVALID, "null pointer");
  VC_CUDA(vc::launch_f32_to_bf16(src, dst_bf16, n, reinterpret_cast<cudaStream_t>(stream)));
  ++g_launches;
  return VC_OK;
}

}  // extern "C"

extern "C" int vc_l2_augment(vc_handle* h, const uint8_t* desc_u8, int n_images, int kmax, int d, void* out_bf16,
                             float* l2_offset, void* stream) {
  if (!h || !desc_u8 || !out_bf16 || !l2_offset) return fail(VC_ERR_INVALID, "null pointer argument");
  if (n_images <= 0 || kmax <= 0 || d <= 0 || d % 8 != 0) return fail(VC_ERR_INVALID, "bad sizes");
  VC_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const size_t rows = (size_t)n_images * kmax;
  // the first word of the output doubles as the reduction cell; the augment pass overwrites it
  int* cell = static_cast<int*>(out_bf16);
  VC_CUDA(vc::launch_l2_max_norm2(desc_u8, rows, d, cell, s));
  int max_norm2 = 0;
  VC_CUDA(cudaMemcpyAsync(&max_norm2, cell, sizeof(int), cudaMemcpyDeviceToHost, s));
  VC_CUDA(cudaStreamSynchronize(s));
  // K: power of two above 2 max|x|^2 >= max |x - y|^2; partial sums stay below 2^24 while K <= 2^23
  long long k = 2;
  while (k <= 2LL * max_norm2 + 1) k <<= 1;
  if (k > (1LL << 23))
    return fail(VC_ERR_INVALID, "descriptor norms too large for the exact l2 rule (need 2*max|x|^2 < 2^23, have %d)",
                max_norm2);
  VC_CUDA(vc::launch_l2_augment(desc_u8, rows, d, (float)k, out_bf16, s));
  g_launches += 2;
  *l2_offset = (float)k;
  return VC_OK;
}

// Host-buffer entry.  The descriptor upload is cut into up to four image chunks on an upload stream
// and the pairs are ordered by the last chunk they need, so group g is matched while chunk g+1 is
// still in flight; each group's match lists are packed on the device and read back with two copies
// on a download stream while the next group runs.  With `resident` the descriptors are already bf16
// in HBM (desc is a device pointer): nothing is uploaded and the pairs are cut into consecutive
// groups only so that the read-back of one overlaps the matching of the next.
static int match_host_impl(vc_handle* h, const void* desc, int dtype, bool resident, const int32_t* counts,
                           int n_images, int kmax, int d, const int32_t* pairs, int n_pairs,
                           const vc_match_options* opt, uint32_t* matches, int32_t* n_matches) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!desc || !counts || !opt || !matches || !n_matches || (n_pairs > 0 && !pairs))
    return fail(VC_ERR_INVALID, "null pointer argument");
  if (dtype != VC_DTYPE_U8 && dtype != VC_DTYPE_F32 && dtype != VC_DTYPE_BF16)
    return fail(VC_ERR_INVALID, "unknown dtype %d", dtype);
  if (n_images <= 0 || kmax <= 0 || n_pairs < 0 || d <= 0 || d % 8 != 0)
    return fail(VC_ERR_INVALID, "bad sizes");
  if (n_pairs == 0) return VC_OK;
  for (int i = 0; i < 2 * n_pairs; ++i)
    if (pairs[i] < 0 || pairs[i] >= n_images)
      return fail(VC_ERR_INVALID, "pair %d names image %d of %d", i / 2, pairs[i], n_images);
  VC_CUDA(cudaSetDevice(h->device));
  if (!h->host_streams[0]) {
    for (auto& st : h->host_streams) VC_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto& ev : h->host_events) VC_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  }
  cudaStream_t s = h->host_streams[0], s_up = h->host_streams[1], s_down = h->host_streams[2];

  const size_t img_el = (size_t)kmax * d;
  const size_t n_el = (size_t)n_images * img_el;
  const size_t el = dtype == VC_DTYPE_U8 ? 1 : (dtype == VC_DTYPE_F32 ? 4 : 2);
  // pair groups: chunk c holds images [c*per, (c+1)*per); a pair belongs to the chunk of its larger index
  const int n_groups = resident ? std::min(4, (n_pairs + 63) / 64)
                                : ((n_images >= 8 && n_el * el >= (size_t)(32u << 20)) ? 4 : 1);
  const int per = (n_images + n_groups - 1) / n_groups;
  std::vector<int> start(n_groups + 1, 0), order(n_pairs);
  if (resident) {
    for (int g = 0; g <= n_groups; ++g) start[g] = (int)((long long)n_pairs * g / n_groups);
    for (int i = 0; i < n_pairs; ++i) order[i] = i;
  } else {
    for (int i = 0; i < n_pairs; ++i) ++start[std::max(pairs[2 * i], pairs[2 * i + 1]) / per + 1];
    for (int g = 0; g < n_groups; ++g) start[g + 1] += start[g];
    std::vector<int> fill(start.begin(), start.end() - 1);
    for (int i = 0; i < n_pairs; ++i) order[fill[std::max(pairs[2 * i], pairs[2 * i + 1]) / per]++] = i;
  }
  int max_group = 0;
  for (int g = 0; g < n_groups; ++g) max_group = std::max(max_group, start[g + 1] - start[g]);

  const size_t src_bytes = (resident || dtype == VC_DTYPE_BF16) ? 0 : align_up(n_el * el, 256);
  const size_t bf_bytes = resident ? 0 : align_up(n_el * 2, 256);
  const size_t cnt_bytes = align_up((size_t)n_images * 4, 256);
  const size_t pair_bytes = align_up((size_t)n_pairs * 8, 256);
  const size_t out_bytes = align_up((size_t)n_pairs * kmax * 8, 256);
  const size_t nm_bytes = align_up((size_t)n_pairs * 4, 256);
  const size_t off_bytes = align_up((size_t)(n_pairs + n_groups) * 4, 256);
  const size_t pack_bytes = out_bytes;
  const size_t ws_bytes = vc_match_workspace_bytes(max_group, kmax);
  const size_t total =
      src_bytes + bf_bytes + cnt_bytes + pair_bytes + out_bytes + nm_bytes + off_bytes + pack_bytes + ws_bytes;
  if (h->stage_bytes < total) {
    VC_CUDA(cudaDeviceSynchronize());
    if (h->stage) cudaFree(h->stage);
    h->stage = nullptr;
    h->stage_bytes = 0;
    VC_CUDA(cudaMalloc(&h->stage, total));
    h->stage_bytes = total;
    h->map_ptr = nullptr;
  }
  // pinned scratch: sorted pair list, counts and offsets per sorted pair, packed match rows
  const size_t pin_pairs = align_up((size_t)n_pairs * 8, 256);
  const size_t pin_small = align_up((size_t)(n_pairs + n_groups) * 4, 256);
  const size_t pin_total = pin_pairs + 2 * pin_small + out_bytes;
  if (h->pin_bytes < pin_total) {
    VC_CUDA(cudaDeviceSynchronize());
    if (h->pin) cudaFreeHost(h->pin);
  if (h->ln_counters) cudaFree(h->ln_counters);
    h->pin = nullptr;
    h->pin_bytes = 0;
    VC_CUDA(cudaHostAlloc(&h->pin, pin_total, cudaHostAllocDefault));
    h->pin_bytes = pin_total;
  }
  uint8_t* pin = static_cast<uint8_t*>(h->pin);
  int32_t* p_pairs = reinterpret_cast<int32_t*>(pin);
  int32_t* p_nm = reinterpret_cast<int32_t*>(pin + pin_pairs);
  int32_t* p_off = reinterpret_cast<int32_t*>(pin + pin_pairs + pin_small);
  uint32_t* p_rows = reinterpret_cast<uint32_t*>(pin + pin_pairs + 2 * pin_small);
  for (int k = 0; k < n_pairs; ++k) {
    p_pairs[2 * k] = pairs[2 * order[k]];
    p_pairs[2 * k + 1] = pairs[2 * order[k] + 1];
  }

  uint8_t* base = static_cast<uint8_t*>(h->stage);
  uint8_t* d_src = base;
  uint8_t* d_bf = resident ? static_cast<uint8_t*>(const_cast<void*>(desc)) : d_src + src_bytes;
  int32_t* d_cnt = reinterpret_cast<int32_t*>(base + src_bytes + bf_bytes);
  int32_t* d_pairs = reinterpret_cast<int32_t*>(reinterpret_cast<uint8_t*>(d_cnt) + cnt_bytes);
  uint32_t* d_out = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(d_pairs) + pair_bytes);
  int32_t* d_nm = reinterpret_cast<int32_t*>(reinterpret_cast<uint8_t*>(d_out) + out_bytes);
  int32_t* d_off = reinterpret_cast<int32_t*>(reinterpret_cast<uint8_t*>(d_nm) + nm_bytes);
  uint32_t* d_pack = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(d_off) + off_bytes);
  void* d_ws = reinterpret_cast<uint8_t*>(d_pack) + pack_bytes;

  VC_CUDA(cudaMemcpyAsync(d_cnt, counts, (size_t)n_images * 4, cudaMemcpyHostToDevice, s));
  VC_CUDA(cudaMemcpyAsync(d_pairs, p_pairs, (size_t)n_pairs * 8, cudaMemcpyHostToDevice, s));
  for (int g = 0; g < n_groups; ++g) {
    const int i0 = std::min(n_images, g * per), i1 = std::min(n_images, (g + 1) * per);
    const size_t off = (size_t)i0 * img_el, cnt = resident ? 0 : (size_t)(i1 - i0) * img_el;
    if (cnt) {
      uint8_t* dst = dtype == VC_DTYPE_BF16 ? d_bf : d_src;
      VC_CUDA(cudaMemcpyAsync(dst + off * el, static_cast<const uint8_t*>(desc) + off * el, cnt * el,
                              cudaMemcpyHostToDevice, s_up));
    }
    VC_CUDA(cudaEventRecord(h->host_events[g], s_up));
    VC_CUDA(cudaStreamWaitEvent(s, h->host_events[g], 0));
    if (cnt && dtype == VC_DTYPE_U8) {
      VC_CUDA(vc::launch_u8_to_bf16(d_src + off, d_bf + off * 2, cnt, s));
      ++g_launches;
    } else if (cnt && dtype == VC_DTYPE_F32) {
      VC_CUDA(vc::launch_f32_to_bf16(reinterpret_cast<const float*>(d_src) + off, d_bf + off * 2, cnt, s));
      ++g_launches;
    }
    const int p0 = start[g], np = start[g + 1] - start[g];
    if (np > 0) {
      int rc = vc_match_pairs(h, d_bf, d_cnt, n_images, kmax, d, d_pairs + 2 * (size_t)p0, np, opt,
                              d_out + (size_t)p0 * kmax * 2, d_nm + p0, d_ws, ws_bytes, s);
      if (rc != VC_OK) {
        cudaDeviceSynchronize();
        return rc;
      }
      VC_CUDA(vc::launch_compact_matches(d_out + (size_t)p0 * kmax * 2, d_nm + p0, np, kmax, d_off + p0 + g,
                                         d_pack + (size_t)p0 * kmax * 2, s));
      g_launches += 2;
    }
    VC_CUDA(cudaEventRecord(h->host_events[8 + g], s));
  }
  // read-back, group by group, while later groups are still being matched
  for (int g = 0; g < n_groups; ++g) {
    const int p0 = start[g], np = start[g + 1] - start[g];
    if (np == 0) continue;
    VC_CUDA(cudaStreamWaitEvent(s_down, h->host_events[8 + g], 0));
    VC_CUDA(cudaMemcpyAsync(p_nm + p0, d_nm + p0, (size_t)np * 4, cudaMemcpyDeviceToHost, s_down));
    VC_CUDA(cudaMemcpyAsync(p_off + p0 + g, d_off + p0 + g, (size_t)(np + 1) * 4, cudaMemcpyDeviceToHost, s_down));
    VC_CUDA(cudaStreamSynchronize(s_down));
    const int32_t* off = p_off + p0 + g;
    const size_t rows = (size_t)off[np];
    uint32_t* host_rows = p_rows + (size_t)p0 * kmax * 2;
    if (rows) {
      VC_CUDA(cudaMemcpyAsync(host_rows, d_pack + (size_t)p0 * kmax * 2, rows * 8, cudaMemcpyDeviceToHost, s_down));
      VC_CUDA(cudaStreamSynchronize(s_down));
    }
    for (int k = 0; k < np; ++k) {
      const int orig = order[p0 + k];
      const int n = p_nm[p0 + k];
      n_matches[orig] = n;
      if (n > 0) std::memcpy(matches + (size_t)orig * kmax * 2, host_rows + (size_t)off[k] * 2, (size_t)n * 8);
    }
  }
  VC_CUDA(cudaStreamSynchronize(s));
  return VC_OK;
}

extern "C" {

int vc_match_pairs_host(vc_handle* h, const void* desc, int dtype, const int32_t* counts,
                        int n_images, int kmax, int d, const int32_t* pairs, int n_pairs,
                        const vc_match_options* opt, uint32_t* matches, int32_t* n_matches) {
  return match_host_impl(h, desc, dtype, false, counts, n_images, kmax, d, pairs, n_pairs, opt, matches,
                         n_matches);
}

int vc_match_pairs_resident(vc_handle* h, const void* desc_bf16, const int32_t* counts, int n_images,
                            int kmax, int d, const int32_t* pairs, int n_pairs,
                            const vc_match_options* opt, uint32_t* matches, int32_t* n_matches) {
  if (desc_bf16) {
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, desc_bf16) != cudaSuccess || attr.type != cudaMemoryTypeDevice) {
      cudaGetLastError();
      return fail(VC_ERR_INVALID, "vc_match_pairs_resident needs a device pointer to bf16 descriptors");
    }
  }
  return match_host_impl(h, desc_bf16, VC_DTYPE_BF16, true, counts, n_images, kmax, d, pairs, n_pairs, opt,
                         matches, n_matches);
}

}  // extern "C"
