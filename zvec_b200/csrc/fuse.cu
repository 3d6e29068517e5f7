// K6 — multi-vector fusion: F per-field top-k lists of one query -> one fused top-n by document key.
//
// Replaces RrfReRanker.rerank and WeightedReRanker.rerank
// (python/zvec/extension/multi_vector_reranker.py:62-89 and :135-174), which the query executor
// applies to the per-field results of a multi-vector query (python/zvec/executor/
// query_executor.py:178-206,276-287):
//   RRF:      score(doc) = sum over the lists holding doc at rank r of 1 / (rank_constant + r + 1)
//   weighted: score(doc) = sum over the lists holding doc of normalize(doc.score) * weight[list],
//             normalize = 1 - 2 atan(s)/pi (L2) | 0.5 + atan(s)/pi (IP) | 1 - s/2 (cosine)
//   result:   heapq.nlargest(topn, by score) — descending, ties keep first-seen order (lists in order,
//             ranks in order).
// All arithmetic in fp64 like the Python code, sums in the same order (list by list), so RRF scores are
// bit-identical; the weighted scores differ only by CUDA's atan against libm's (<= 2 ulp).
// F*k is a few hundred entries per query (the 4 scans dominate): one CTA per query, the
// all-pairs form (n^2 key compares) is the cheapest thing that keeps the order semantics.
#include "internal.h"

namespace zvb {

namespace {

constexpr int kFuseThreads = 256;

struct FuseArgs {
  const uint64_t *keys;    // [F][q][k]
  const float *scores;     // [F][q][k] user-facing scores (weighted mode), or null (RRF)
  const uint32_t *counts;  // [F][q] valid entries per list
  const double *weights;   // [F] (weighted mode)
  uint32_t F, q, k, topn;
  uint32_t rank_constant;
  int mode;                // 0 RRF, 1 weighted L2, 2 weighted IP, 3 weighted cosine
  int negate;              // weighted: scores are internal (smaller is better) IP scores: user score = -s
  uint64_t *out_keys;      // [q][topn]
  double *out_scores;      // [q][topn]
  uint32_t *out_counts;    // [q]
};

__global__ void __launch_bounds__(kFuseThreads) fuse_kernel(const FuseArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  const uint32_t qi = blockIdx.x;
  const uint32_t n = a.F * a.k;
  uint64_t *key = reinterpret_cast<uint64_t *>(smem);
  double *contrib = reinterpret_cast<double *>(key + n);
  double *total = contrib + n;
  uint8_t *flag = reinterpret_cast<uint8_t *>(total + n);  // 0 absent, 1 first occurrence, 2 repeat
  __shared__ uint32_t ndocs;
  if (threadIdx.x == 0) ndocs = 0;
  const double pi = 3.141592653589793;  // math.pi
  for (uint32_t e = threadIdx.x; e < n; e += blockDim.x) {
    const uint32_t f = e / a.k, r = e % a.k;
    const bool live = r < a.counts[(size_t)f * a.q + qi];
    const size_t src = ((size_t)f * a.q + qi) * a.k + r;
    key[e] = live ? a.keys[src] : 0;
    flag[e] = live ? 1 : 0;
    double c = 0.0;
    if (live) {
      if (a.mode == 0) {
        c = 1.0 / (double)(a.rank_constant + r + 1);
      } else {
        double s = (double)a.scores[src];
        if (a.negate) s = -s;
        const double nz = a.mode == 1 ? 1.0 - 2 * atan(s) / pi : a.mode == 2 ? 0.5 + atan(s) / pi : 1.0 - s / 2.0;
        c = nz * a.weights[f];
      }
    }
    contrib[e] = c;
  }
  __syncthreads();
  // first occurrences sum their document's contributions in list order (dict insertion + "+=" order)
  for (uint32_t e = threadIdx.x; e < n; e += blockDim.x) {
    if (!flag[e]) continue;
    const uint64_t ke = key[e];
    bool first = true;
    for (uint32_t j = 0; j < e; ++j)
      if (flag[j] && key[j] == ke) { first = false; break; }
    if (!first) continue;  // flag stays 1 until every thread has looked; marked below
    double t = 0.0;
    for (uint32_t j = e; j < n; ++j)
      if (flag[j] && key[j] == ke) t += contrib[j];
    total[e] = t;
  }
  __syncthreads();
  for (uint32_t e = threadIdx.x; e < n; e += blockDim.x) {
    if (!flag[e]) continue;
    const uint64_t ke = key[e];
    for (uint32_t j = 0; j < e; ++j)
      if (flag[j] && key[j] == ke) { flag[e] = 2; break; }  // only ever reads flags as zero / non-zero
  }
  __syncthreads();
  // rank among documents: score descending, first seen first
  for (uint32_t e = threadIdx.x; e < n; e += blockDim.x) {
    if (flag[e] != 1) continue;
    atomicAdd(&ndocs, 1u);
    const double te = total[e];
    uint32_t rank = 0;
    for (uint32_t j = 0; j < n; ++j) {
      if (flag[j] != 1 || j == e) continue;
      const double tj = total[j];
      if (tj > te || (tj == te && j < e)) ++rank;
    }
    if (rank < a.topn) {
      a.out_keys[(size_t)qi * a.topn + rank] = key[e];
      a.out_scores[(size_t)qi * a.topn + rank] = te;
    }
  }
  __syncthreads();
  const uint32_t cnt = min(ndocs, a.topn);
  if (threadIdx.x == 0) a.out_counts[qi] = cnt;
  for (uint32_t i = cnt + threadIdx.x; i < a.topn; i += blockDim.x) {
    a.out_keys[(size_t)qi * a.topn + i] = ZVB200_INVALID_KEY;
    a.out_scores[(size_t)qi * a.topn + i] = __longlong_as_double(0x7ff8000000000000ll);
  }
}

}  // namespace

size_t fuse_max_entries() { return 1536; }  // F * k; 25 bytes of shared memory each

cudaError_t launch_fuse(const uint64_t *keys, const float *scores, const uint32_t *counts, const double *weights,
                        uint32_t F, uint32_t q, uint32_t k, uint32_t topn, uint32_t rank_constant, int mode, int negate,
                        uint64_t *out_keys, double *out_scores, uint32_t *out_counts, cudaStream_t st) {
  if (q == 0) return cudaSuccess;
  FuseArgs a;
  a.keys = keys; a.scores = scores; a.counts = counts; a.weights = weights;
  a.F = F; a.q = q; a.k = k; a.topn = topn; a.rank_constant = rank_constant; a.mode = mode; a.negate = negate;
  a.out_keys = out_keys; a.out_scores = out_scores; a.out_counts = out_counts;
  const size_t n = (size_t)F * k;
  const size_t smem = n * 25 + 16;
  fuse_kernel<<<q, kFuseThreads, smem, st>>>(a);
  count_launch();
  return cudaGetLastError();
}

}  // namespace zvb
