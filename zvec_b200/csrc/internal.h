// Host-side structures shared by the C ABI and the kernel launchers.
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include <stdint.h>

#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace zvb {

// What a scan kernel needs to know about the rows it streams (device pointers).
struct RowsView {
  const uint8_t *rows = nullptr;  // n rows, row_stride bytes apart, 16-B aligned
  uint64_t n = 0;
  uint32_t row_stride = 0;  // bytes, multiple of 16
  uint32_t chunks = 0;      // 16-byte chunks of a row that carry vector data
  uint32_t d = 0;           // elements that enter the dot product (tails stripped)
  int dtype = 0, metric = 0;
  const float4 *tails = nullptr;    // QI8: {a, b, s, s2} per row
  const float *norms = nullptr;     // |x|^2 per row (tensor-core L2 path)
  const uint32_t *drop = nullptr;   // bit set = deleted row or kInvalidKey hole
  bool query_first = false;         // QuantizedInteger -IP: evaluate as the reference's streamer does (common.cuh)
  // QuantizedInteger L2: bf16 image of every row (scaled codes + the row's share of the score's bilinear
  // terms), shadow_k halves per row — what the tcgen05 scan contracts instead of the int8 codes (scan_mma.cu)
  const uint8_t *shadow = nullptr;
  uint32_t shadow_k = 0;
  float shadow_mrow_max = 0.f;      // largest row share of the shadow filter's margin (the first-tile bootstrap's bound)
  // scores leave the scan multiplied by row_scale[row] (per posting list, expanded per row) or by score_scale: the
  // `distances[k] * norm_val` of IVFEntity::search (ivf_entity.cc:601,650; zvb200_store_set_score_scale).  Honoured by
  // the GEMV scan, the posting-list scans; the other kernels are not chosen for a scaled store
  const float *row_scale = nullptr;
  float score_scale = 1.f;
  bool scaled() const { return row_scale || score_scale != 1.f; }
};

// Queries in "compute layout", produced by prep_queries() on the device.
struct QueryView {
  const float *qf32 = nullptr;     // fp32/fp16 stores: q x (chunks*16/esz) floats, zero padded
  const uint8_t *qraw = nullptr;   // raw codes/halves: q x row_stride bytes, zero padded
  const float4 *qtails = nullptr;  // QI8 query tails
  const float *qnorms = nullptr;   // |q|^2
  const uint8_t *qshadow = nullptr;  // QuantizedInteger L2: bf16 image of the queries (the rows' counterpart)
  uint32_t q = 0;
  bool qpad_rows = false;          // qraw holds zero rows up to the next multiple of 128 queries
};

struct SearchParams {
  uint32_t k = 0;
  float threshold = 0.f;
  const uint32_t *filter = nullptr;  // device bitmap, bit set = drop
  bool no_shadow = false;            // QuantizedInteger L2: contract the int8 codes (kind::i8), not the bf16 shadow
};

// Partial top-k lists: part[query][nparts][k] packed keys, padded with kMaxKey.
struct PartView {
  u64 *part = nullptr;
  uint32_t nparts = 0;
};

// ---- launchers (each returns a cudaError_t; all asynchronous on `st`) ----
// K1: GEMV-style streaming scan, CUDA cores, best for q <= 4.
cudaError_t launch_scan_gemv(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                             uint64_t row_begin, uint64_t row_end, u64 *part, uint32_t *nparts,
                             int sm_count, cudaStream_t st, u64 *all_keys = nullptr /* [q][keys_stride]: every row's key
                             (kMaxKey = none) instead of per-CTA top-k lists */, uint32_t keys_stride = 0);
// the k smallest of `total` packed keys per query (keys[q][total], kMaxKey = none), sorted: one CTA per query, radix
// select in registers; total <= select_topk_max()
// plan != null: the k selected keys are a query's coarse ids — its posting-list plan (launch_ivf_plan's outputs) is
// written by the same kernel
struct IvfPlanArgs {
  const uint32_t *list_off = nullptr, *budget_sizes = nullptr;
  uint64_t max_scan_count = 0;
  uint32_t *seg_prefix = nullptr, *seg_row = nullptr, *seg_count = nullptr, *seg_list = nullptr;
};
cudaError_t launch_select_topk(const u64 *keys, uint32_t total, uint32_t q, uint32_t k, float *out_scores,
                               uint32_t *out_index, cudaStream_t st, const IvfPlanArgs *plan = nullptr);
uint32_t select_topk_max();
size_t scan_gemv_max_parts(int sm_count);
// K1s: bulk-async (cp.async.bulk + mbarrier ring) streaming scan, exact arithmetic, grid.y = query
cudaError_t launch_scan_stream(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                               uint64_t row_begin, uint64_t row_end, u64 *part, uint32_t *nparts,
                               int sm_count, cudaStream_t st);
size_t scan_stream_max_parts(int sm_count);
bool scan_stream_supports(const RowsView &rv, uint32_t k, uint32_t seg_cap);
// K2c: tiled CUDA-core contraction, exact (x-q)^2 for L2; any q.
cudaError_t launch_scan_tile(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                             uint64_t row_begin, uint64_t row_end, u64 *part, uint32_t *nparts,
                             int sm_count, cudaStream_t st);
size_t scan_tile_max_parts(int sm_count);
bool scan_tile_supports(uint32_t k);
// K2: tcgen05/TMEM contraction (fp16 stores, InnerProduct / Cosine), thread-per-query top-k
// epilogue; at most sm_count query tiles of 128 per launch.
cudaError_t launch_scan_mma(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                            uint64_t row_begin, uint64_t row_end, u64 *part, uint32_t *nparts,
                            uint32_t *pub /* qv.q * scan_mma_max_parts words of scratch */, int sm_count, cudaStream_t st,
                            int *kind_out = nullptr /* 0 f16, 1 tf32, 2 i8, 3 bf16 shadow */);
size_t scan_mma_max_parts(int sm_count);
bool scan_mma_supports(const RowsView &rv, uint32_t k);
// queries (stored layout, device) -> compute layout
cudaError_t launch_prep_queries(const RowsView &rv, const void *queries_dev, uint32_t q,
                                uint32_t elem_bytes, float *qf32, uint8_t *qraw, float4 *qtails,
                                float *qnorms, cudaStream_t st);
// part[q][nparts][k] -> final sorted (score,index) q*k, padded NaN / INVALID
cudaError_t launch_merge_parts(const u64 *part, uint32_t nparts, uint32_t q, uint32_t k,
                               float *out_scores, uint32_t *out_index, u64 *out_keys /* packed, may be null */,
                               cudaStream_t st);
// G sorted shard lists [g][q][k] -> q*k by (score, shard_base[g]+index)
cudaError_t launch_merge_shards(const float *scores, const uint32_t *index, uint32_t g,
                                const uint64_t *shard_base_dev, uint32_t q, uint32_t k,
                                float *out_scores, uint64_t *out_rows, cudaStream_t st);
// G sorted lists [g][q][k] with explicit rows of the whole index -> q*k by (score, row)
cudaError_t launch_merge_rows(const float *scores, const uint64_t *rows, uint32_t g, uint32_t q, uint32_t k,
                              float *out_scores, uint64_t *out_rows, cudaStream_t st);
// local rows of a list-grouped store -> rows of the whole index
cudaError_t launch_expand_list_scale(const uint32_t *list_off, uint32_t nlist, const float *scales, uint64_t n,
                                     float *row_scale, cudaStream_t st);
cudaError_t launch_ivf_global_rows(const uint32_t *index, uint64_t count, const uint32_t *list_off,
                                   const uint64_t *gbase, uint32_t nlist, uint64_t *out, cudaStream_t st);
// reference element layout (host order) -> store layout
cudaError_t launch_repack_rows(const uint8_t *src, uint32_t elem_bytes, uint64_t n, uint8_t *dst,
                               uint32_t row_stride, uint32_t vec_bytes, float4 *tails,
                               uint32_t tail_off, cudaStream_t st);
cudaError_t launch_row_norms(const RowsView &rv, uint64_t row_begin, uint64_t row_end,
                             float *norms, cudaStream_t st);
// QuantizedInteger L2 shadow operands (aux_kernels.cu): rows [row_begin, row_end) of the store, and a query batch
inline uint32_t qi8_shadow_k(uint32_t d) { return (d + 9u + 63u) & ~63u; }
cudaError_t launch_qi8_shadow_rows(const RowsView &rv, uint64_t row_begin, uint64_t row_end, uint8_t *shadow,
                                   uint32_t shadow_k, float *mrow_max /* device, running maximum */, cudaStream_t st);
cudaError_t launch_qi8_shadow_queries(const RowsView &rv, const uint8_t *qraw, const float4 *qtails, uint32_t q,
                                      uint8_t *out, uint32_t shadow_k, cudaStream_t st);
cudaError_t launch_distance_matrix(const RowsView &rv, const QueryView &qv, float *out,
                                   cudaStream_t st);

// dist[q][n] -> the `take` best (score, row) per query, ascending; scratch = q * p2 keys, p2 = pow2 >= n
cudaError_t launch_sort_rows(const float *dist, uint32_t q, uint32_t n, uint32_t p2, uint32_t take, u64 *scratch,
                             float *out_scores, uint32_t *out_index, cudaStream_t st);

// K4: IVF — list selection under the scan budget, then the posting-list scan (grid.y = query)
cudaError_t launch_ivf_plan(const uint32_t *coarse_index, uint32_t q, uint32_t nprobe, const uint32_t *list_off,
                            const uint32_t *budget_sizes, uint64_t max_scan_count, uint32_t *seg_prefix,
                            uint32_t *seg_row, uint32_t *seg_count, cudaStream_t st,
                            uint32_t *seg_list = nullptr /* [q][nprobe] list id of every visited segment */);
cudaError_t launch_scan_lists(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                              const uint32_t *seg_prefix, const uint32_t *seg_row, const uint32_t *seg_count,
                              uint32_t seg_cap, u64 *part, uint32_t *nparts, int sm_count, cudaStream_t st,
                              uint64_t rows_hint = 0 /* rows one query scans, if the caller knows: picks the tile size */);
size_t scan_lists_parts(uint32_t q, int sm_count);
cudaError_t launch_scan_lists_stream(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                                     const uint32_t *seg_prefix, const uint32_t *seg_row, const uint32_t *seg_count,
                                     uint32_t seg_cap, u64 *part, uint32_t *nparts, int sm_count, cudaStream_t st);
size_t scan_lists_stream_parts(uint32_t q, int sm_count);
// K4b: the same scan LIST-major for query batches (scan_lists_lm.cu): every probed list is streamed once per <= QB of
// the queries that probe it; part[q][cap][k], one partial list per (query, probe slot)
int scan_lists_lm_qb(const RowsView &rv, uint32_t k);  // queries per list pass; 0 = unsupported shape
size_t scan_lists_lm_scratch_words(uint32_t q, uint32_t cap, uint32_t nlist);
cudaError_t launch_scan_lists_lm(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                                 const uint32_t *seg_list, const uint32_t *seg_count, uint32_t cap,
                                 const uint32_t *list_off, uint32_t nlist, uint32_t *scratch, u64 *part,
                                 uint32_t *nparts, int sm_count, cudaStream_t st);

// Lanes of a warp that share one stored row in the streaming scans (log2).  Measured on B200
// (tools/scan_probe.py, 10M rows, batch 1): ~6 sixteen-byte chunks per lane is the sweet spot —
// fewer lanes per row saves shuffle-reduction steps and top-k pushes per byte, but below 2 lanes
// the per-lane strides defeat sector coalescing.  512-B rows: 32 lanes 3.4 TB/s, 4 lanes 7.0 TB/s.
inline int max_lpr_log2(uint32_t chunks) {
  int lg = 0;
  while ((2u << lg) <= chunks && lg < 5) ++lg;
  return lg;
}
inline int pick_lpr_log2(uint32_t chunks) {
  static const int force = [] { const char *e = getenv("ZVB200_LPR_LOG2"); return e ? atoi(e) : -1; }();
  const int top = max_lpr_log2(chunks);
  if (force >= 0) return force < top ? force : top;
  int lg = 1;
  while (lg < 5 && chunks >= 9u << lg) ++lg;  // chunks / lanes in [4.5, 9)
  return lg < top ? lg : top;
}

// K6: multi-vector fusion (RRF / weighted) of F per-field lists, one CTA per query (fuse.cu)
cudaError_t launch_fuse(const uint64_t *keys, const float *scores, const uint32_t *counts, const double *weights,
                        uint32_t F, uint32_t q, uint32_t k, uint32_t topn, uint32_t rank_constant, int mode, int negate,
                        uint64_t *out_keys, double *out_scores, uint32_t *out_counts, cudaStream_t st);
size_t fuse_max_entries();

// K7: query reformers (fp32 -> the store's query layout) and the refiner's candidate plan (refine.cu)
enum { RF_NONE = 0, RF_FP16 = 1, RF_COS_FP32 = 2, RF_COS_FP16 = 3, RF_QI8_L2 = 4, RF_QI8_IP = 5 };
int reform_kind(int dtype, int metric);  // -1: the store type has no fp32 reformer
cudaError_t launch_reform(const float *in, uint32_t q, uint32_t dim, int kind, uint8_t *out, uint32_t elem_bytes,
                          cudaStream_t st);
cudaError_t launch_rows_plan(const uint32_t *rows, uint32_t q, uint32_t m, uint64_t n, uint32_t *seg_prefix,
                             uint32_t *seg_row, uint32_t *seg_count, cudaStream_t st);

// K8: Lloyd update (kmeans.cu)
cudaError_t launch_kmeans_accumulate(const float *rows, uint64_t n, uint32_t dim, const uint32_t *labels,
                                     const float *dist, double *sums, unsigned long long *counts, double *scores,
                                     cudaStream_t st);
cudaError_t launch_kmeans_fix_labels(uint32_t *labels, float *dist, uint64_t n, cudaStream_t st);
cudaError_t launch_kmeans_finalize(const double *sums, const unsigned long long *counts, const double *scores,
                                   uint32_t k, uint32_t dim, float *centroids, unsigned long long *follows,
                                   double *sse, cudaStream_t st);

// cudaFuncAttributeMaxDynamicSharedMemorySize is per-FUNCTION state shared by every host thread: setting it to each
// launch's own size lets one thread lower it between another thread's set and its launch ("too many resources
// requested for launch", seen with eight contexts searching at once).  Every launcher sets this one constant.
constexpr int kMaxDynSmem = 232448;  // 227 KB, the per-CTA limit of sm_100

void count_launch(int n = 1);

}  // namespace zvb
