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
} vc_match_options;

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

/* Element-wise converters used by the host entry and the Python mirror (device pointers).
 * u8 -> bf16 is exact (0..255 fit bf16's 8-bit significand), so uint8 d<=258 matching through
 * the bf16 tensor-core path reproduces COLMAP's int32 dot products bit for bit. */
int vc_convert_u8_to_bf16(const uint8_t* src, void* dst_bf16, size_t n, void* stream);
int vc_convert_f32_to_bf16(const float* src, void* dst_bf16, size_t n, void* stream);

/* Number of kernels launched by this library on the calling thread since the last reset
 * (bench.py reports it as gpu_launches). */
int64_t vc_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* VC_B200_H_ */
