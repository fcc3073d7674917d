/*
 * vc_b200.h -- C ABI of the B200-native vit-colmap hot path (libvc_b200 / _vc_b200.so).
 *
 * Plain pointers and sizes only; no torch / C++ types cross this boundary.  Every function
 * returns 0 on success or a negative vc_status; vc_last_error() gives the message of the last
 * failure on the calling thread.  No exception crosses the ABI.  All device work is enqueued on
 * the caller's stream (a cudaStream_t passed as void*); nothing synchronises unless stated.
 * A handle belongs to one device and one host thread at a time.
 *
 * Each entry point names the reference interface it replaces (paths under /root/reference,
 * COLMAP = third_party/colmap/src/colmap, SIFTGPU = third_party/colmap/src/thirdparty/SiftGPU).
 */
#ifndef VC_B200_H_
#define VC_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vc_handle vc_handle;

typedef enum {
  VC_OK = 0,
  VC_ERR_INVALID = -1,   /* bad argument */
  VC_ERR_CUDA = -2,      /* CUDA runtime / driver error */
  VC_ERR_NO_DEVICE = -3, /* no sm_100 device: there is NO CPU fallback */
  VC_ERR_WORKSPACE = -4  /* workspace too small */
} vc_status;

/* Matching thresholds: COLMAP/feature/sift.h:108-126 (SiftMatchingOptions: max_ratio 0.8,
 * max_distance 0.7, cross_check true), vit_colmap/utils/config.py:58-62. */
typedef enum { VC_DTYPE_U8 = 0, VC_DTYPE_F32 = 1, VC_DTYPE_BF16 = 2 } vc_dtype;

typedef struct {
  float max_ratio;
  float max_distance;
  int32_t cross_check;
  /* score scale applied before acos: 1/512^2 for uint8 COLMAP descriptors
   * (COLMAP/feature/sift.cc:63,739), 1.0 for unit-norm float descriptors. */
  float inv_norm;
  /* Acceptance rule applied to each row's / column's (best, second) scores:
   *   VC_RULE_ANGULAR (0): theta = acos(min(score * inv_norm, 1)); reject theta1 > max_distance or
   *     theta1 >= max_ratio * theta2 -- FindBestMatchesOneWayBruteForce, COLMAP/feature/sift.cc:733-787
   *     (the GPU / brute-force path).
   *   VC_RULE_L2 (1): scores are l2_offset - squared L2 distance (the caller augments the descriptors
   *     so that the dot product is exactly that, see vit_colmap_b200/matching.py); reject
   *     d1^2 > max_distance^2 / inv_norm or sqrt(d1^2) >= max_ratio * sqrt(d2^2) --
   *     FindBestMatchesOneWayIndex, COLMAP/feature/sift.cc:827-874 (the CPU index path, done exactly). */
  int32_t rule;
  float l2_offset;
} vc_match_options;
enum { VC_RULE_ANGULAR = 0, VC_RULE_L2 = 1 };

const char* vc_last_error(void);
const char* vc_version(void);

/* Opens a per-device context (workspaces, cached TMA descriptors).  Fails with
 * VC_ERR_NO_DEVICE unless the device is compute capability 10.x.
 * Replaces: SiftMatchGPU creation in COLMAP/feature/sift.cc:1174-1248 (one matcher per GPU,
 * COLMAP/controllers/feature_matching_utils.cc:66-84). */
int vc_create(int device, vc_handle** out);
int vc_destroy(vc_handle* h);

/* Bytes of device scratch vc_match_pairs needs for n_pairs pairs of up to kmax descriptors. */
size_t vc_match_workspace_bytes(int n_pairs, int kmax);

/*
 * Fused exhaustive matcher for a block of image pairs (descriptors resident in HBM).
 *   desc      : device, bf16 [n_images, kmax, d] row-major, 16-byte aligned, d % 8 == 0
 *   counts    : device, int32 [n_images], valid rows per image (rows >= count are ignored)
 *   pairs     : device, int32 [n_pairs, 2], indices into the image axis (image1, image2)
 *   matches   : device, uint32 [n_pairs, kmax, 2] -> (idx1, idx2) ascending in idx1
 *   n_matches : device, int32 [n_pairs]
 *   (matches == n_matches == NULL runs only the top-2 stage and leaves the accepted best
 *    indices, int32 [n_pairs, 2, kmax], at the start of `workspace`.)
 * Semantics: FindBestMatchesBruteForce, COLMAP/feature/sift.cc:733-825 (score matrix
 * sift.cc:922-971); replaces SiftMatchGPU::GetSiftMatch (SIFTGPU/SiftMatchCU.cpp:161-199) and
 * the kernels SIFTGPU/ProgramCU.cu:1411-1495,1696-1797.  The N x M score matrix stays in
 * tensor memory; only (best, second, index) per row/column reaches HBM.
 * Under the angular rule the epilogue does not track scores with acos(inv_norm * s) >
 * max_distance / max_ratio: they can change no accept / reject decision of sift.cc:765-780, so the
 * match lists are identical (vc_match_pairs_top2 returns exact values for every row regardless).
 * Environment overrides for comparisons: VC_MATCH_KERNEL=dual|fused, VC_MATCH_2CTA=0|1 (two-SM
 * cta_group::2 kernel, default from d >= 768), VC_MATCH_NO_PRUNE=1.
 */
int vc_match_pairs(vc_handle* h, const void* desc, const int32_t* counts, int n_images, int kmax,
                   int d, const int32_t* pairs, int n_pairs, const vc_match_options* opt,
                   uint32_t* matches, int32_t* n_matches, void* workspace, size_t workspace_bytes,
                   void* stream);

/* Diagnostics of the same launch (optional, may be skipped): per pair and direction
 * (0 = image1 rows, 1 = image2 rows) the raw best score, second-best score and best index.
 * All three device arrays are [n_pairs, 2, kmax].  Same arguments as vc_match_pairs. */
int vc_match_pairs_top2(vc_handle* h, const void* desc, const int32_t* counts, int n_images,
                        int kmax, int d, const int32_t* pairs, int n_pairs,
                        const vc_match_options* opt, float* best, float* second, int32_t* index,
                        void* workspace, size_t workspace_bytes, void* stream);

/*
 * Guided matching: the same exhaustive search with a geometric gate on every candidate
 * (i, j): keypoint i of image 1 and keypoint j of image 2 must agree with the pair's two-view geometry
 * within `max_error` pixels, otherwise the candidate does not take part (angular rule: score 0; l2 rule:
 * squared distance 512^2).  The gate is evaluated in the GEMM epilogue on the accumulator tile -- the
 * dot matrix is never written, where SIFTGPU's MultiplyDescriptorG_Kernel writes it gated to HBM.
 *   keypoints : device f32 [n_images (n_images / 2 under rule l2), kmax, 2] = (x, y) per descriptor row
 *   geometry  : device vc_pair_geometry [n_pairs]
 * mode VC_GUIDED_SIFTGPU: keep iff |hnorm(H p1) - p2|^2 < max_h AND (p2^T F p1)^2 / (|F p1|_xy^2 +
 *   |F^T p2|_xy^2) < max_f  (ProgramCU.cu:1573-1600; an absent H or F is passed as identity with bound
 *   1e20, SiftMatch.cpp:676-689);
 * mode VC_GUIDED_F / VC_GUIDED_H: COLMAP's CPU MatchGuided (feature/sift.cc:1093-1122): one test by
 *   TwoViewGeometry::config, a candidate is dropped iff its residual is > the bound.
 * Every residual operation is a separately rounded fp32 op (documented in match_kernel.cuh); candidates
 * whose residual is within rounding of the bound are the only place results may differ from a build of
 * the reference with other FMA contraction.
 * Replaces: SiftMatchGPU::GetGuidedSiftMatch (SIFTGPU/SiftMatchCU.cpp:148-159, ProgramCU.cu:1521-1663),
 * SiftCPUFeatureMatcher::MatchGuided (COLMAP/feature/sift.cc:1058-1157), as driven by
 * COLMAP/controllers/feature_matching_utils.cc:129-146.
 */
enum { VC_GUIDED_SIFTGPU = 0, VC_GUIDED_F = 1, VC_GUIDED_H = 2 };
typedef struct {
  float h[9];      /* row-major homography, image 1 -> image 2 (TwoViewGeometry::H) */
  float f[9];      /* row-major fundamental matrix, p2^T F p1 (TwoViewGeometry::F) */
  int32_t mode;
  float max_h;     /* max_error^2 (1e20 when H is absent under VC_GUIDED_SIFTGPU) */
  float max_f;
} vc_pair_geometry;
int vc_match_pairs_guided(vc_handle* h, const void* desc, const float* keypoints, const int32_t* counts,
                          int n_images, int kmax, int d, const int32_t* pairs, const vc_pair_geometry* geometry,
                          int n_pairs, const vc_match_options* opt, uint32_t* matches, int32_t* n_matches,
                          void* workspace, size_t workspace_bytes, void* stream);

/* Test hook: dense score matrix of ONE pair through the same TMA/tcgen05 main loop.
 * scores: device float [kmax, ld] (only [count1, count2] written). */
int vc_debug_scores(vc_handle* h, const void* desc, const int32_t* counts, int n_images, int kmax,
                    int d, int image1, int image2, float* scores, int ld, void* stream);

/*
 * Host-buffer entry (what a pycolmap-style caller binds): descriptors and results live in host
 * memory; the call uploads, matches every pair and downloads, synchronising before it returns.
 *   desc      : host [n_images, kmax, d] in `dtype`: VC_DTYPE_U8 (COLMAP FeatureDescriptors,
 *               feature/types.h:80-81), VC_DTYPE_F32 or VC_DTYPE_BF16 (unit-norm rows,
 *               north-star float mode).  Pinned memory makes the upload asynchronous.
 *   counts    : host int32 [n_images];  pairs: host int32 [n_pairs, 2]
 *   matches   : host uint32 [n_pairs, kmax, 2];  n_matches: host int32 [n_pairs]
 * Replaces: the matcher worker loop COLMAP/controllers/feature_matching_utils.cc:109-167
 * (GetDescriptors -> matcher->Match) as driven by pycolmap.match_exhaustive
 * (PYCOLMAP/pipeline/match_features.cc:28-51).
 */
int vc_match_pairs_host(vc_handle* h, const void* desc, int dtype, const int32_t* counts,
                        int n_images, int kmax, int d, const int32_t* pairs, int n_pairs,
                        const vc_match_options* opt, uint32_t* matches, int32_t* n_matches);

/* Same call for descriptors that are ALREADY resident in HBM (bf16 [n_images, kmax, d], device
 * pointer, written before the call on a stream the caller has synchronised): nothing but the pair
 * list is uploaded; match lists come back to host memory, the read-back of one pair group overlapping
 * the matching of the next.  This is what vit_colmap_b200.matching.match_exhaustive loops over pair
 * chunks after uploading the database's descriptors ONCE.
 * Replaces: the per-pair descriptor re-upload of SiftGPUFeatureMatcher::Match
 * (COLMAP/feature/sift.cc:1270-1284, SIFTGPU/SiftMatchCU.cpp:99-112). */
int vc_match_pairs_resident(vc_handle* h, const void* desc_bf16, const int32_t* counts, int n_images,
                            int kmax, int d, const int32_t* pairs, int n_pairs,
                            const vc_match_options* opt, uint32_t* matches, int32_t* n_matches);

/* Element-wise converters used by the host entry and the Python mirror (device pointers).
 * u8 -> bf16 is exact (0..255 fit bf16's 8-bit significand), so uint8 d<=258 matching through
 * the bf16 tensor-core path reproduces COLMAP's int32 dot products bit for bit. */
int vc_convert_u8_to_bf16(const uint8_t* src, void* dst_bf16, size_t n, void* stream);
int vc_convert_f32_to_bf16(const float* src, void* dst_bf16, size_t n, void* stream);

/* Rule VC_RULE_L2 preparation, on the device: uint8 descriptors [n_images, kmax, d] -> bf16
 * [2*n_images, kmax, d+8] whose rows are augmented so that (row of copy i) . (row of copy
 * n_images + j) = K - |x - y|^2 exactly (integers of at most 8 significant bits, partial sums below
 * 2^24: fp32 accumulation is exact, ties resolve as in COLMAP).  Match pair (i, j) as
 * (i, n_images + j) with opt.rule = VC_RULE_L2, opt.l2_offset = *l2_offset (= K, returned to the
 * host; the call synchronises `stream` once to read the largest squared norm).
 * Replaces: the squared-L2 search of COLMAP/feature/sift.cc:827-874 / feature/index.cc:41-131 (done
 * exactly instead of through FAISS IVF). */
int vc_l2_augment(vc_handle* h, const uint8_t* desc_u8, int n_images, int kmax, int d, void* out_bf16,
                  float* l2_offset, void* stream);

/* ------------------------------------------------------------------------------------------
 * Extraction post-processing (vit_colmap/features/vit_extractor.py ViTExtractor._dense_to_sparse
 * :168-252).  All buffers are device memory; every call is batched over n_images feature maps of
 * one patch grid (grid_h x grid_w <= 16384 cells, channels <= 1024 and a multiple of 4).
 * `tokens` is the backbone output in its native token-major layout [n_images, grid_h*grid_w,
 * channels] (x_norm_patchtokens, vit_extractor.py:135-146), f32 or bf16, image_stride elements
 * apart; the NCHW view of vit_extractor.py:150-156 is never materialised.
 * ------------------------------------------------------------------------------------------ */

/* Detection methods of ViTExtractor(detection_method=...), vit_extractor.py:254-279. */
typedef enum { VC_DETECT_HARRIS = 0, VC_DETECT_DOG = 1, VC_DETECT_COMBINED = 2 } vc_detect_method;

size_t vc_score_workspace_bytes(int n_images, int grid_h, int grid_w);

/* Distinctiveness map.  score: f32 [n_images, grid_h, grid_w], min-max normalised per image.
 * Replaces: _compute_distinctiveness / _harris_response / _dog_response,
 * vit_extractor.py:254-394 (forward-difference structure tensor averaged over channels, 3x3
 * Gaussian, 0.7*(det - 0.04 tr^2) + 0.3*sqrt(tr); |G7 - G11| of the channel mean). */
int vc_score_map(vc_handle* h, const void* tokens, int dtype, int n_images, int grid_h, int grid_w,
                 int channels, int64_t image_stride, int method, float* score, void* workspace,
                 size_t workspace_bytes, void* stream);

/* Allocation-free variant of vc_score_map: same arguments without the workspace, same kernels, same
 * result bit for bit.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Its scratch is a file-scope
 * `static __device__` array of VC_SCORE_STATIC_MAX_FLOATS floats (4 MiB) that every GPU loading the
 * library holds until the process exits; the call itself performs no cudaMalloc / cudaFree and
 * calls no allocating library routine.  Limit: n_images * 5 * grid_h * grid_w <=
 * VC_SCORE_STATIC_MAX_FLOATS (153 images of 37x37 cells, 21 of 85x114); larger requests fail with
 * VC_ERR_WORKSPACE and leave `score` untouched.  Every call on a device shares the one array, which
 * keeps the last call's contents (nothing in it has to start at zero: both regions are fully
 * written before they are read).  Calls that overlap on the GPU (two streams, or two handles on one
 * device) race on it: serialise them on one stream or with events. */
#define VC_SCORE_STATIC_MAX_FLOATS (1u << 20)
int vc_score_map_static(vc_handle* h, const void* tokens, int dtype, int n_images, int grid_h, int grid_w,
                        int channels, int64_t image_stride, int method, float* score, void* stream);

/* Keypoint selection.  coords: int32 [n_images, capacity, 2] as (y, x) grid cells, scores: f32
 * [n_images, capacity], counts: int32 [n_images]; rows come out in descending score order, ties
 * by the reference's concatenation order.  capacity >= min(num_keypoints, grid cells).
 * Replaces: _spatial_binning_selection (bin_size 16, per-bin topk, global topk)
 * vit_extractor.py:404-485 and the greedy _apply_nms (radius 1.5) :500-543. */
int vc_select_keypoints(vc_handle* h, const float* score, int n_images, int grid_h, int grid_w,
                        int num_keypoints, int bin_size, float nms_radius, int32_t* coords,
                        float* scores, int32_t* counts, int capacity, void* stream);

/* Descriptor quantisation rules of the byte blob written to the database. */
typedef enum {
  VC_QUANT_VIT = 0,       /* trunc(clip(512 v, 0, 255)), vit_extractor.py:250 */
  VC_QUANT_COLMAP = 1,    /* min(255, round(512 v)), COLMAP/feature/utils.cc:56-71 */
  VC_QUANT_TRAINABLE = 2  /* trunc(clip((v + 1) * 127.5, 0, 255)), trainable_vit_extractor.py:265-267 */
} vc_quant_mode;
/* OR-ed into quant_mode: RootSIFT normalisation (L1-normalise, sqrt(clamp(., 1e-8)), then the usual
 * L2 step) of the hybrid extractor, vit_colmap/features/hybrid_extractor.py:286-288. */
#define VC_NORM_ROOTSIFT 0x10

/* Sample, (project,) L2-normalise and pack descriptors at the selected keypoints.
 *   coords      : [n_images, capacity, 2] (y, x); int32 grid cells, or float32 grid units when
 *                 coords_are_float != 0 (bilinear between cells, border clamp)
 *   projection  : f32 [channels, d] applied before normalisation (PCA of
 *                 vit_extractor.py:588-653, fitted by the caller), or NULL with d == channels
 *   projection_bias : f32 [d] added after the projection, or NULL (the 1x1 descriptor convolution
 *                 of vit_feature_model.py:120 applied only at the selected keypoints)
 *   kp_scale    : host float[4] = {w_res/W, w_orig/w_res, h_res/H, h_orig/h_res} so that
 *                 x = (col + 0.5) * kp_scale[0] * kp_scale[1] (vit_extractor.py:229-236)
 *   keypoints   : f32 [n_images, capacity, 2] (x, y) pixel coordinates           (may be NULL)
 *   desc_bf16   : bf16 [n_images, capacity, d], unit rows = matcher input         (may be NULL)
 *   desc_u8     : uint8 [n_images, capacity, d], database blob rows              (may be NULL)
 *   desc_f32    : f32 [n_images, capacity, d], unit rows                         (may be NULL)
 * Rows >= counts[i] (up to capacity) are zero-filled in every output given, so the buffers may be
 * handed over uninitialised.
 * Replaces: _extract_descriptors_interp (F.grid_sample bilinear/border/align_corners=True)
 * vit_extractor.py:545-586, the projection apply :651, F.normalize :243 and the uint8
 * conversion :246-250. */
int vc_sample_pack(vc_handle* h, const void* tokens, int dtype, int n_images, int grid_h, int grid_w,
                   int channels, int64_t image_stride, const void* coords, int coords_are_float,
                   const int32_t* counts, int capacity, const float* projection,
                   const float* projection_bias, int d, int quant_mode, const float* kp_scale, float* keypoints, void* desc_bf16, uint8_t* desc_u8,
                   float* desc_f32, void* stream);

/* Keypoint selection of the `trainable_vit` extractor.
 *   kp_map    : device f32 [n_images, map_h*map_w, 4], channel-last head output at 1/4 resolution:
 *               score logit, dx, dy, orientation (raw when apply_tanh != 0: tanh(.)*pi is applied
 *               here as vit_colmap/model/vit_feature_model.py:222-223 does)
 *   kp_scale  : host float[4] = {w_orig/w_new, h_orig/h_new, w_orig-1, h_orig-1}
 *   coords    : int32 [n_images, capacity, 2] (y, x) cells, for vc_sample_pack(quant TRAINABLE)
 *   keypoints : f32 [n_images, capacity, 6] = x, y, 1, orientation, score, 0
 * Rows are in descending score order (ties: lower row-major cell first).
 * Replaces: sigmoid, _simple_nms (max_pool2d, ties kept), threshold, nonzero, topk and the
 * sub-pixel coordinate arithmetic of trainable_vit_extractor.py:114-138,183-254. */
size_t vc_trainable_workspace_bytes(int n_images, int map_h, int map_w);
int vc_trainable_select(vc_handle* h, const float* kp_map, int n_images, int map_h, int map_w,
                        int num_keypoints, float score_threshold, int nms_radius, int apply_tanh,
                        const float* kp_scale, int32_t* coords, float* keypoints, int32_t* counts,
                        int capacity, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * DINOv2 ViT-S/B/L/14 forward (the backbone the reference pulls through torch.hub,
 * vit_colmap/features/vit_extractor.py:95-98, and calls as
 * model.forward_features(x)["x_norm_patchtokens"], :135-146; preprocessing :117-132).
 * Linear weights are bf16 in PyTorch's [out, in] layout, everything else f32; all device memory.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const float *ln1_w, *ln1_b;
  const void* qkv_w;   /* bf16 [3C, C] */
  const float* qkv_b;  /* [3C] */
  const void* proj_w;  /* bf16 [C, C] */
  const float* proj_b;
  const float* ls1;    /* LayerScale gamma [C] */
  const float *ln2_w, *ln2_b;
  const void* fc1_w;   /* bf16 [4C, C] */
  const float* fc1_b;
  const void* fc2_w;   /* bf16 [C, 4C] */
  const float* fc2_b;
  const float* ls2;
  /* Optional (all six or none, NULL otherwise): LayerNorm folded into the GEMMs around it (VC_VIT_LN=fold).
   * qkv_wf = bf16(qkv_w * ln1_w) [3C, C], qkv_cs[n] = sum_k qkv_wf[n, k] (f32), qkv_bf = qkv_b + qkv_w ln1_b (f32);
   * fc1_* likewise with ln2.  Built by vit_colmap_b200/vit.py from the same state dict. */
  const void* qkv_wf;
  const float *qkv_cs, *qkv_bf;
  const void* fc1_wf;
  const float *fc1_cs, *fc1_bf;
} vc_vit_block;

typedef struct {
  int32_t embed_dim;   /* C: 384 / 768 / 1024 */
  int32_t depth;       /* 12 / 12 / 24 */
  int32_t num_heads;   /* C / 64 */
  int32_t patch_kpad;  /* 640: 3*14*14 = 588 patch values zero-padded to a multiple of 64 */
  const void* patch_w;     /* bf16 [C, patch_kpad], conv weight [C,3,14,14] flattened (c,dy,dx) */
  const float* patch_b;    /* [C] */
  const float* cls_token;  /* [C] */
  const float* pos_embed;  /* f32 [1 + grid_h*grid_w, C], interpolated to this grid by the caller */
  const float *norm_w, *norm_b;
  const vc_vit_block* blocks; /* HOST array [depth] of device pointers */
} vc_vit_weights;

size_t vc_vit_workspace_bytes(int n_images, int grid_h, int grid_w, int embed_dim);

/* images: device uint8 [n_images, 14*grid_h, 14*grid_w, 3] (HWC; channel order RGB, or BGR as
 * cv2.imread delivers when bgr != 0), already at a multiple-of-14 size.
 * tokens: device f32 [n_images, grid_h*grid_w, C] = x_norm_patchtokens (cls token dropped). */
int vc_vit_forward(vc_handle* h, const vc_vit_weights* w, const uint8_t* images, int bgr, int n_images,
                   int grid_h, int grid_w, float* tokens, void* workspace, size_t workspace_bytes,
                   void* stream);

/* Building blocks of the forward, exposed for the per-op parity tests.
 * vc_vit_linear: out = epilogue(A[M,K] bf16 * W[N,K]^T bf16 + bias); epilogue 0: bf16 out,
 * 1: bf16 GELU(erf) out, 2: f32 out += gamma * (.).  K % 64 == 0, N % 32 == 0. */
int vc_vit_linear(vc_handle* h, const void* a, const void* w, const float* bias, const float* gamma,
                  int m, int n, int k, int epilogue, void* out, void* stream);
/* vc_vit_linear with the residual epilogue (x += gamma * (A W^T + bias), fp32, [m, n]) followed IN THE SAME
 * KERNEL by LayerNorm(x; ln_w, ln_b, eps 1e-6) -> ln_out bf16 [m, n]: the CTA that stores the last column
 * tile of a 128-row block normalises that block out of L2.  n is the full row (<= 1024); ln_out must not
 * alias a.  Bit-identical to vc_vit_linear(epilogue 2) + vc_vit_layernorm; vc_vit_forward uses it only with
 * VC_VIT_FUSED_LN=1 (measured slower than the stand-alone LayerNorm launches). */
int vc_vit_linear_ln(vc_handle* h, const void* a, const void* w, const float* bias, const float* gamma,
                     const float* ln_w, const float* ln_b, int m, int n, int k, float* x, void* ln_out_bf16,
                     void* stream);
/* vc_vit_attention: qkv bf16 [n_images, tokens, 3*heads*64] -> out bf16 [n_images, tokens, heads*64] */
int vc_vit_attention(vc_handle* h, const void* qkv, int n_images, int tokens, int heads, void* out,
                     void* stream);
/* vc_vit_layernorm: f32 [rows, C] -> bf16 [rows, C], eps 1e-6 */
int vc_vit_layernorm(vc_handle* h, const float* x, const float* w, const float* b, int rows, int channels,
                     void* out_bf16, void* stream);

/* ------------------------------------------------------------------------------------------
 * Convolutional head of ViTFeatureModel (vit_colmap/model/vit_feature_model.py:12-29 UpsampleBlock,
 * :90-121 upsampler / trunk / heads, :173-295 forward / forward_from_backbone_features), eval mode.
 * Activations are channels-last bf16 on the device; convolution weights are bf16 repacked by the
 * caller (vit_colmap_b200/head.py) into GEMM rows, BatchNorm is folded into per-channel fp32
 * scale/shift applied in the epilogue:  y = gelu(conv_nobias * scale + shift),
 * scale = gamma / sqrt(var + 1e-5), shift = (conv_bias - mean) * scale + beta.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  int32_t backbone_dim;    /* C: 384 / 768 / 1024 (multiple of 64) */
  int32_t descriptor_dim;  /* d (multiple of 32) */
  /* upsampler.{0,1}: ConvTranspose2d(k4,s2,p1) as 4 sub-pixel phases, bf16 [4*512, 4*Cin]: row
   * (2*py+px)*512 + co, column (2*ty+tx)*Cin + ci = weight[ci, co, ky, kx] with (phase 0: tap 0 ->
   * k=3, offset -1; tap 1 -> k=1, offset 0; phase 1: tap 0 -> k=2, offset 0; tap 1 -> k=0, offset +1) */
  const void* deconv_w[2];
  const float* deconv_b[2];    /* [512] */
  /* Conv3x3 weights bf16 [Cout, 9*Cin], column (ky*3+kx)*Cin + ci */
  const void* conv_w[2];       /* [512, 9*512] */
  const float* conv_scale[2];  /* [512] */
  const float* conv_shift[2];
  const void* trunk_w;         /* [256, 9*512] */
  const float *trunk_scale, *trunk_shift;
  const void* heads_w;         /* [192, 9*256]: rows 0..63 keypoint_head.0, rows 64..191 descriptor_head.0 */
  const float *heads_scale, *heads_shift; /* [192] */
  const float* kp_w;           /* f32 [4, 64]  keypoint_head.3 */
  const float* kp_b;           /* f32 [4] */
  const void* desc_w;          /* bf16 [d, 128] descriptor_head.3 */
  const float* desc_b;         /* f32 [d] */
} vc_head_weights;

size_t vc_head_workspace_bytes(int n_images, int grid_h, int grid_w, int backbone_dim, int out_h,
                               int out_w);

/* tokens: device [n_images, grid_h*grid_w, C] patch tokens (f32 or bf16; = the NCHW feature map of
 *         vit_feature_model.py:165-169 in channels-last form).
 * out_h, out_w: target size, (14*grid_h)//4 x (14*grid_w)//4 in the reference (:195-197).
 * kp_map       : f32 [n_images, out_h, out_w, 4]  score logit, dx, dy, tanh(.)*pi        (required)
 * desc_map     : f32 [n_images, out_h, out_w, d]  L2-normalised (eps 1e-8)               (may be NULL)
 * features     : bf16 [n_images, out_h, out_w, 256] trunk output                         (may be NULL)
 * desc_hidden  : bf16 [n_images, out_h, out_w, 128] descriptor head before its 1x1 conv  (may be NULL;
 *                with desc_map == NULL the caller projects only the selected keypoints:
 *                vc_sample_pack(projection = desc_w^T, bias))
 * Replaces: ViTFeatureModel.forward_from_backbone_features, vit_feature_model.py:234-295. */
int vc_conv_head_forward(vc_handle* h, const vc_head_weights* w, const void* tokens, int dtype,
                         int n_images, int grid_h, int grid_w, int out_h, int out_w, float* kp_map,
                         float* desc_map, void* features, void* desc_hidden, void* workspace,
                         size_t workspace_bytes, void* stream);

/* Building block exposed for the per-op parity tests: one convolution through the implicit-GEMM
 * kernel.  kind 0: Conv3x3 pad 1 (w bf16 [Cout, 9*Cin]); 1: ConvTranspose2d k4 s2 p1 (w bf16
 * [4*Cout, 4*Cin] phase-packed as above); 2: Conv1x1 (w bf16 [Cout, Cin]).
 * epilogue 0: out bf16 = acc + shift; 1: out bf16 = gelu(acc*scale + shift); 2: out f32 = acc + shift.
 * in: bf16 [n, h, w, Cin]; out: [n, h*s, w*s, Cout] with s = 2 for kind 1. */
int vc_conv2d(vc_handle* h, const void* in, int n, int height, int width, int c_in, const void* w,
              int c_out, int kind, int epilogue, const float* scale, const float* shift, void* out,
              void* stream);
/* F.interpolate(bilinear, align_corners=False) on bf16 channels-last maps (vit_feature_model.py:205-212). */
int vc_resize_bilinear(vc_handle* h, const void* in, int n, int in_h, int in_w, int out_h, int out_w,
                       int channels, void* out, void* stream);

/* ------------------------------------------------------------------------------------------
 * Training-time similarity and hard-negative mining on patch tokens.
 * Reference: vit_colmap/dataloader/training_sampler.py TrainingSampler.select_invariant_points
 * :56-181, sample_features_at_coords :183-211, generate_hard_negatives :452-534;
 * vit_colmap/dataloader/homography_utils.py warp_patch_tokens :10-92, compute_feature_similarity
 * :219-253.  Feature maps are fp32, token-major [n, grid_h*grid_w, channels] (the backbone's own
 * layout), channels a multiple of 4 and <= 1024, at most 16384 cells; all buffers device memory.
 * Everything is computed in fp32: the outputs are index lists, which only equal the reference's
 * when the similarities agree to ~1e-6.
 * ------------------------------------------------------------------------------------------ */

/* [n, C, H, W] fp32 (the reference's layout) -> token-major [n, H*W, C].  Replaces nothing in the
 * reference (its permutes are views); exists so that callers holding NCHW maps can use the entries
 * below without a PyTorch copy. */
int vc_nchw_to_tokens(vc_handle* h, const float* nchw, int n_images, int channels, int grid_h, int grid_w,
                      float* tokens, void* stream);

/* similarity[b, y, x] = cosine(warp(tokens1[b])[y, x], tokens2[b][y, x]): tokens1 is resampled at
 * h_inv[b] * (patch centre (x+.5, y+.5) * patch_size) with F.grid_sample(bilinear, zeros,
 * align_corners=False) semantics after normalising by (img_w, img_h); both vectors are L2-normalised
 * with eps 1e-12.  h_inv: device f32 [n_pairs, 9] row-major = torch.linalg.inv(H) computed by the caller.
 * Replaces: homography_utils.compute_feature_similarity :219-253 (= warp_patch_tokens :10-92 +
 * F.normalize + channel sum), the first half of TrainingSampler.select_invariant_points :83-97. */
int vc_warp_similarity(vc_handle* h, const float* tokens1, const float* tokens2, int n_pairs, int grid_h,
                       int grid_w, int channels, const float* h_inv, float patch_size, float img_h, float img_w,
                       float* similarity, void* stream);

/* Per row of values [n_rows, n] (n <= 16384):
 *   order   int32 [n_rows, k]: indices by value descending, lower index first among equal values
 *           (torch.topk leaves tie order unspecified), -1 beyond n; may be NULL when k == 0
 *   above   int32 [n_rows, n]: ascending indices with value >= threshold (torch.nonzero order), or NULL
 *   n_above int32 [n_rows]: how many, or NULL
 * Replaces: the topk / threshold / hybrid selections of select_invariant_points :100-166 (the mode
 * logic on these three outputs stays on the host: vit_colmap_b200/dataloader/training_sampler.py). */
int vc_topk_rows(vc_handle* h, const float* values, int n_rows, int n, int k, float threshold, int32_t* order,
                 int32_t* above, int32_t* n_above, void* stream);

/* Hard negatives: for every positive k of image b, the n_hard (<= 32) cells with the highest cosine
 * similarity to pos_feat[b, k] among the cells at least `radius` (feature-map units, Euclidean) away
 * from pos_xy[b, k] = (x, y).  idx int32 [n_images, n_pos, n_hard] by similarity descending (lower cell
 * first on ties), -1 where fewer than n_hard cells qualify; counts int32 [n_images, n_pos].
 * The n_pos x cells similarity matrix is never stored.
 * Replaces: the double loop of TrainingSampler.generate_hard_negatives :489-512 (matmul per positive,
 * distance mask, topk); the gather that follows is vc_sample_bilinear. */
int vc_hard_negatives(vc_handle* h, const float* tokens, const float* pos_feat, const float* pos_xy, int n_images,
                      int grid_h, int grid_w, int channels, int n_pos, int n_hard, float radius, int32_t* idx,
                      int32_t* counts, void* stream);

/* out[b, i, :] = bilinear sample of tokens[b] at xy[b, i] = (x, y) feature-map units, exactly as the
 * reference does it: g = v / (size - 1) * 2 - 1 handed to F.grid_sample(bilinear, zeros,
 * align_corners=False), i.e. source position v * size / (size - 1) - 0.5.  A NaN coordinate gives a
 * zero row (padding entries).  out: f32 [n_images, n_points, channels].
 * Replaces: TrainingSampler.sample_features_at_coords :183-211. */
int vc_sample_bilinear(vc_handle* h, const float* tokens, int n_images, int grid_h, int grid_w, int channels,
                       const float* xy, int n_points, float* out, void* stream);

/* Number of kernels launched by this library on the calling thread since the last reset
 * (bench.py reports it as gpu_launches). */
int64_t vc_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* VC_B200_H_ */
