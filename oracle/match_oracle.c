/*
 * oracle/match_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the brute-force descriptor matcher that vit-colmap reaches
 * through pycolmap.match_exhaustive (vendored COLMAP 3.13.0.dev0).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product path (vit_colmap_b200/) never does.
 *
 * Parity pin: COLMAP itself cannot be built in this image (Eigen, glog, FAISS,
 * sqlite3.h absent; SiftGPU links OpenGL), so the restatement is pinned against
 * the known-answer tests of third_party/colmap/src/colmap/feature/sift_test.cc
 * (KA1..KA4, see tests/test_oracle_matching.py) and pairing_test.cc (KA5).
 *
 * Reference lines restated (paths under third_party/colmap/src/colmap/):
 *   feature/sift.cc:63        kSqSiftDescriptorNorm = 512*512
 *   feature/sift.cc:922-971   ComputeSiftDistanceMatrix(DOT_PRODUCT): int dot products
 *   feature/sift.cc:733-787   FindBestMatchesOneWayBruteForce
 *   feature/sift.cc:789-825   FindBestMatchesBruteForce (cross check, emit order)
 *   feature/sift.cc:985-1030  empty-input guards
 *   feature/pairing.cc:202-236 ExhaustivePairGenerator::Next
 *   util/types.h:188-208      ImagePairToPairId
 *
 * Two arithmetic modes share one control flow:
 *   u8  : uint8 rows, exact int32 dot products, scale 1/512^2     (COLMAP compat)
 *   f32 : float rows (unit norm), dot accumulated in double and rounded to
 *         float, scale 1                                          (north-star d=768 mode)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define VC_SQ_SIFT_NORM (512.0 * 512.0) /* feature/sift.cc:63 */

typedef struct {
  float best;   /* best score (raw dot product)          */
  float second; /* second-best score                     */
  int32_t idx;  /* index of best, -1 if no score > 0     */
} vc_top2;

/* feature/sift.cc:745-757 -- running best / second with init 0 and strict '>' */
static inline void top2_push(vc_top2* t, float s, int32_t j) {
  if (s > t->best) {
    t->idx = j;
    t->second = t->best;
    t->best = s;
  } else if (s > t->second) {
    t->second = s;
  }
}

/* feature/sift.cc:759-783 -- distance + ratio test on one (best, second) pair.
 * Returns idx or -1. */
static inline int32_t top2_accept(const vc_top2* t, float inv_norm, float max_ratio,
                                  float max_distance) {
  if (t->idx == -1) return -1;
  const float best_dist = acosf(fminf(inv_norm * t->best, 1.0f));
  if (best_dist > max_distance) return -1;
  const float second_dist = acosf(fminf(inv_norm * t->second, 1.0f));
  if (best_dist >= max_ratio * second_dist) return -1;
  return t->idx;
}

static float dot_u8(const uint8_t* a, const uint8_t* b, int d) {
  int32_t acc = 0; /* feature/sift.cc:939-942: cast<int>() then dot */
  for (int k = 0; k < d; ++k) acc += (int32_t)a[k] * (int32_t)b[k];
  return (float)acc; /* stored into RowMajorMatrixXf (sift.cc:945,961) */
}

static float dot_f32(const float* a, const float* b, int d) {
  double acc = 0.0;
  for (int k = 0; k < d; ++k) acc += (double)a[k] * (double)b[k];
  return (float)acc;
}

/* float accumulation, vectorisable: only used to TIME the CPU baseline (mode 2) */
static float dot_f32_fast(const float* a, const float* b, int d) {
  float acc = 0.f;
#pragma omp simd reduction(+ : acc)
  for (int k = 0; k < d; ++k) acc += a[k] * b[k];
  return acc;
}

/*
 * Core: fills row top-2 (n1 entries) and column top-2 (n2 entries).
 * Columns are scanned in ascending row order so tie-breaking (lowest index wins,
 * strict '>') is identical to running the one-way routine on the transpose
 * (feature/sift.cc:803-804).
 */
static void top2_both(const void* d1, int n1, const void* d2, int n2, int d, int is_u8,
                      vc_top2* rows, vc_top2* cols) {
  for (int j = 0; j < n2; ++j) {
    cols[j].best = 0.f;
    cols[j].second = 0.f;
    cols[j].idx = -1;
  }
  /* rows in parallel blocks; column state merged per thread in row order */
  int nthreads = 1;
#ifdef _OPENMP
  nthreads = omp_get_max_threads();
#endif
  if (nthreads > n1) nthreads = n1 > 0 ? n1 : 1;
  vc_top2* tcols = (vc_top2*)malloc((size_t)nthreads * (size_t)(n2 > 0 ? n2 : 1) * sizeof(vc_top2));
#pragma omp parallel num_threads(nthreads)
  {
    int t = 0;
#ifdef _OPENMP
    t = omp_get_thread_num();
#endif
    vc_top2* mycols = tcols + (size_t)t * (size_t)(n2 > 0 ? n2 : 1);
    for (int j = 0; j < n2; ++j) {
      mycols[j].best = 0.f;
      mycols[j].second = 0.f;
      mycols[j].idx = -1;
    }
    /* contiguous row chunk per thread keeps ascending-row order inside a thread */
    const int chunk = (n1 + nthreads - 1) / nthreads;
    const int lo = t * chunk;
    const int hi = lo + chunk < n1 ? lo + chunk : n1;
    for (int i = lo; i < hi; ++i) {
      vc_top2 r = {0.f, 0.f, -1};
      for (int j = 0; j < n2; ++j) {
        const float s =
            is_u8 == 1
                ? dot_u8((const uint8_t*)d1 + (size_t)i * d, (const uint8_t*)d2 + (size_t)j * d, d)
                : (is_u8 == 2 ? dot_f32_fast((const float*)d1 + (size_t)i * d,
                                             (const float*)d2 + (size_t)j * d, d)
                              : dot_f32((const float*)d1 + (size_t)i * d,
                                        (const float*)d2 + (size_t)j * d, d));
        top2_push(&r, s, j);
        top2_push(&mycols[j], s, i);
      }
      rows[i] = r;
    }
  }
  /* merge thread-local column states in ascending row-chunk order; equivalent to
   * one sequential scan because an earlier chunk wins ties (strict '>') */
  for (int t = 0; t < nthreads; ++t) {
    const vc_top2* mycols = tcols + (size_t)t * (size_t)(n2 > 0 ? n2 : 1);
    for (int j = 0; j < n2; ++j) {
      vc_top2* c = &cols[j];
      const vc_top2* m = &mycols[j];
      if (m->idx == -1) continue;
      if (m->best > c->best) {
        const float s2 = m->second > c->best ? m->second : c->best;
        c->second = s2;
        c->best = m->best;
        c->idx = m->idx;
      } else {
        if (m->best > c->second) c->second = m->best;
      }
    }
  }
  free(tcols);
}

/*
 * vc_oracle_match: full FindBestMatchesBruteForce (feature/sift.cc:789-825).
 *   mode_u8 == 1 : d1/d2 are uint8 [n,d] row-major, scores scaled by 1/512^2
 *   mode_u8 == 0 : d1/d2 are float [n,d] row-major, double accumulation, scale 1
 *   mode_u8 == 2 : as 0 with float (SIMD) accumulation -- CPU-baseline timing only
 * out_matches: caller buffer of at least 2*min(n1,n2) uint32 (pairs idx1,idx2),
 *              at least 2*n1 when cross_check == 0.
 * Optional diagnostics (may be NULL): row_top2 [n1], col_top2 [n2] as
 * (best, second, idx) triples -- used by tests to classify near-ties.
 * Returns number of matches.
 */
int vc_oracle_match(const void* d1, int n1, const void* d2, int n2, int d, int mode_u8,
                    float max_ratio, float max_distance, int cross_check,
                    uint32_t* out_matches, float* row_top2, float* col_top2) {
  if (n1 <= 0 || n2 <= 0) return 0; /* feature/sift.cc:1002-1005 */
  vc_top2* rows = (vc_top2*)malloc((size_t)n1 * sizeof(vc_top2));
  vc_top2* cols = (vc_top2*)malloc((size_t)n2 * sizeof(vc_top2));
  top2_both(d1, n1, d2, n2, d, mode_u8, rows, cols);
  const float inv_norm = mode_u8 == 1 ? (float)(1.0 / VC_SQ_SIFT_NORM) : 1.0f;

  int32_t* m12 = (int32_t*)malloc((size_t)n1 * sizeof(int32_t));
  int32_t* m21 = (int32_t*)malloc((size_t)n2 * sizeof(int32_t));
  for (int i = 0; i < n1; ++i) m12[i] = top2_accept(&rows[i], inv_norm, max_ratio, max_distance);
  for (int j = 0; j < n2; ++j) m21[j] = top2_accept(&cols[j], inv_norm, max_ratio, max_distance);

  int n = 0;
  for (int i = 0; i < n1; ++i) { /* ascending i1: feature/sift.cc:805,816 */
    const int32_t j = m12[i];
    if (j == -1) continue;
    if (cross_check && !(m21[j] != -1 && m21[j] == i)) continue; /* sift.cc:806-808 */
    out_matches[2 * n] = (uint32_t)i;
    out_matches[2 * n + 1] = (uint32_t)j;
    ++n;
  }
  if (row_top2) {
    for (int i = 0; i < n1; ++i) {
      row_top2[3 * i] = rows[i].best;
      row_top2[3 * i + 1] = rows[i].second;
      row_top2[3 * i + 2] = (float)rows[i].idx;
    }
  }
  if (col_top2) {
    for (int j = 0; j < n2; ++j) {
      col_top2[3 * j] = cols[j].best;
      col_top2[3 * j + 1] = cols[j].second;
      col_top2[3 * j + 2] = (float)cols[j].idx;
    }
  }
  free(rows);
  free(cols);
  free(m12);
  free(m21);
  return n;
}

/*
 * vc_oracle_exhaustive_pairs: ExhaustivePairGenerator (feature/pairing.cc:171-236),
 * all Next() calls concatenated.  ids: image ids in database order.
 * out_pairs: 2*n*(n-1)/2 uint32; block_sizes (optional, may be NULL): number of
 * pairs emitted by each Next() call, up to max_blocks entries; *num_blocks set.
 * Returns total number of pairs.
 */
int64_t vc_oracle_exhaustive_pairs(const uint32_t* ids, int n, int block_size,
                                   uint32_t* out_pairs, int32_t* block_sizes, int max_blocks,
                                   int* num_blocks) {
  int64_t total = 0;
  int nb = 0;
  size_t start1 = 0, start2 = 0;
  while (start1 < (size_t)n) { /* HasFinished: pairing.cc:198-200 */
    const size_t end1 = ((size_t)n < start1 + block_size ? (size_t)n : start1 + block_size) - 1;
    const size_t end2 = ((size_t)n < start2 + block_size ? (size_t)n : start2 + block_size) - 1;
    int32_t in_block = 0;
    for (size_t i1 = start1; i1 <= end1; ++i1) {
      for (size_t i2 = start2; i2 <= end2; ++i2) {
        const size_t b1 = i1 % block_size, b2 = i2 % block_size;
        if ((i1 > i2 && b1 <= b2) || (i1 < i2 && b1 < b2)) { /* pairing.cc:223-224 */
          out_pairs[2 * total] = ids[i1];
          out_pairs[2 * total + 1] = ids[i2];
          ++total;
          ++in_block;
        }
      }
    }
    if (block_sizes && nb < max_blocks) block_sizes[nb] = in_block;
    ++nb;
    start2 += block_size; /* pairing.cc:230-234 */
    if (start2 >= (size_t)n) {
      start2 = 0;
      start1 += block_size;
    }
  }
  if (num_blocks) *num_blocks = nb;
  return total;
}

/* util/types.h:100-101,200-208: kMaxNumImages = 2^31-1; pair id with id1<=id2 swap */
int64_t vc_oracle_pair_id(uint32_t id1, uint32_t id2) {
  const int64_t kMaxNumImages = 2147483647;
  if (id1 > id2) {
    const uint32_t t = id1;
    id1 = id2;
    id2 = t;
  }
  return kMaxNumImages * (int64_t)id1 + (int64_t)id2;
}

/* Launchers such as torchrun export OMP_NUM_THREADS=1; the timed CPU baseline asks for all cores. */
void vc_oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int vc_oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
