/* zvec_b200.h — C ABI of the B200-native exhaustive-scan path (distance + fused top-k)
 * that replaces zvec's Flat / IVF posting-list scan.  Plain pointers and sizes only.
 *
 * Score convention = the reference's INTERNAL one: smaller is better;
 * inner product is returned as -x.q, cosine as 1 - x^.q^; IndexMetric::normalize()
 * (src/core/metric/inner_product_metric.cc:377-379) is applied by the caller, exactly
 * as src/core/interface/index.cc:624-628 does after search_impl().
 *
 * Every call returns 0 or a NEGATIVE code with the reference's IndexError values
 * (src/core/framework/index_error.cc:19-70).  There is no CPU fallback: without a
 * CUDA device every compute entry point returns ZVB200_ERR_RUNTIME.
 */
#ifndef ZVEC_B200_H_
#define ZVEC_B200_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZVB200_OK 0
#define ZVB200_ERR_RUNTIME (-1)           /* IndexError_Runtime (CUDA / NCCL failure) */
#define ZVB200_ERR_UNSUPPORTED (-12)      /* IndexError_Unsupported */
#define ZVB200_ERR_NOMEMORY (-19)         /* IndexError_NoMemory */
#define ZVB200_ERR_NOREADY (-21)          /* IndexError_NoReady (store not sealed) */
#define ZVB200_ERR_MISMATCH (-24)         /* IndexError_Mismatch */
#define ZVB200_ERR_INVALID_ARGUMENT (-31) /* IndexError_InvalidArgument */

/* IndexMeta::DataType values (src/include/zvec/core/framework/index_meta.h:31-40). */
#define ZVB200_DT_FP16 1
#define ZVB200_DT_FP32 2
#define ZVB200_DT_INT8 4

/* Metric = reference metric name (+ origin metric for "QuantizedInteger").
 *  L2SQ    "SquaredEuclidean"  src/core/metric/euclidean_metric.cc:743-906
 *  L2      "Euclidean"         euclidean_metric.cc:910-1027
 *  IP      "InnerProduct"      inner_product_metric.cc:256-434        (score = -x.q)
 *  COSINE  "Cosine"            cosine_metric.cc:141-230; rows carry a trailing fp32 norm
 *                              slot (dim+1 fp32 / dim+2 fp16) that is not read
 *  QI8_*   "QuantizedInteger"  quantized_integer_metric.cc:30-318 with origin metric
 *                              SquaredEuclidean / InnerProduct / Cosine; rows = d int8
 *                              codes + 16 B tail (+4 B norm for cosine)
 *                              (quantized_integer_metric_matrix.h:36-60,94-112,167-174)
 */
#define ZVB200_METRIC_L2SQ 0
#define ZVB200_METRIC_L2 1
#define ZVB200_METRIC_IP 2
#define ZVB200_METRIC_COSINE 3
#define ZVB200_METRIC_QI8_L2SQ 4
#define ZVB200_METRIC_QI8_IP 5
#define ZVB200_METRIC_QI8_COSINE 6

#define ZVB200_INVALID_INDEX 0xffffffffu
#define ZVB200_INVALID_KEY 0xffffffffffffffffull /* kInvalidKey, index_helper / flat_streamer_entity.cc:269 */
#define ZVB200_MAX_TOPK 1024u
#define ZVB200_MAX_CANDIDATES 8192u /* candidate rows per query of zvb200_search_rows */

typedef struct zvb200_store zvb200_store; /* device-resident vector store: on one GPU, or row-sharded over several */
typedef struct zvb200_ctx zvb200_ctx;     /* per-thread search context (IndexContext) */
typedef struct zvb200_comm zvb200_comm;   /* one rank of a multi-process, one-GPU-per-process row-sharded search */

const char *zvb200_version(void);
/* Message of the last failing call on this thread. */
const char *zvb200_last_error(void);
/* Number of CUDA devices visible (0 when there is none / no driver). */
int zvb200_device_count(void);
/* Kernels launched by this library in this process so far (bench.py's gpu_launches). */
uint64_t zvb200_kernel_launches(void);

/* ---- store: replaces FlatSearcher::load / FlatStreamerEntity storage reads
 * (src/core/algorithm/flat/flat_searcher.cc:118-152, flat_streamer_entity.cc:900-991).
 * `dim` is the reference's STORED dimension (IndexMeta::dimension(), tails included);
 * one element is IndexMeta::element_size() bytes: dim*4 (fp32), dim*2 (fp16), dim (int8).
 */
int zvb200_store_create(int device, int dtype, int metric, uint32_t dim, zvb200_store **out);
/* The multi-GPU layer inside ONE process (what a zvec process gets with proxima.gpu.devices =
 * "0,1,2,3"): the same handle, rows sharded over `devices` as contiguous ranges — shard g holds rows
 * [g * S, (g + 1) * S), S = ceil(N / ndev) rounded up to 1024, N = the FIRST zvb200_store_reserve() (required before
 * the first append; appends beyond it fail).  Every store / context / search entry point below works on it unchanged
 * with GLOBAL row numbers; a search runs the scan on every device, each device's k best per query travel to
 * devices[0] over NVLink and are merged there by (score, global row) — the device form of
 * CombinedVectorColumnIndexer::Search's concat + sort + truncate (combined_vector_column_indexer.cc:98-232), bit
 * identical to the same rows in one store.  Not on a sharded store (ZVB200_ERR_UNSUPPORTED): zvb200_ivf_*,
 * zvb200_reform_queries*, zvb200_search_fp32, zvb200_search_refined.  ndev = 1 is zvb200_store_create.  A store
 * holds at most 2^32 - 2 rows either way (ZVB200_ERR_INVALID_ARGUMENT beyond). */
int zvb200_store_create_sharded(const int *devices, uint32_t ndev, int dtype, int metric, uint32_t dim,
                                zvb200_store **out);
int zvb200_store_destroy(zvb200_store *s);
/* Pre-size the device arrays for `rows` rows (optional; append grows geometrically). */
int zvb200_store_reserve(zvb200_store *s, uint64_t rows);
/* Append n rows in the reference's row-major element layout from HOST memory.
 * keys may be NULL (key = row index).  A key of ZVB200_INVALID_KEY marks a hole that the
 * scan skips (flat_streamer_entity.cc:269). */
int zvb200_store_append(zvb200_store *s, const void *rows, uint64_t n, const uint64_t *keys);
/* Same, rows already in DEVICE memory of the store's GPU (synthetic data, ingest kernels). */
int zvb200_store_append_device(zvb200_store *s, const void *rows_dev, uint64_t n);
/* Publish appended rows to searches (computes |x|^2 for the batched L2 path). */
int zvb200_store_seal(zvb200_store *s);
uint64_t zvb200_store_count(const zvb200_store *s);
/* Mark a row deleted / undeleted (FlatStreamerEntity DeletionMap, flat_index_format.h:106-126). */
int zvb200_store_set_deleted(zvb200_store *s, uint64_t row, int deleted);
/* Which operand the reference's caller hands its 1x1 distance function first.  FlatSearcher passes (feature, query)
 * (src/core/algorithm/flat/flat_distance_matrix.h:53-55): query_first = 0, the default.  FlatStreamer and the IVF list
 * scan pass (query, feature) (flat_streamer_entity.h:155-162, ivf_distance_calculator.h:174-183): query_first = 1.
 * Only the QuantizedInteger -IP / cosine float epilogue (quantized_integer_metric_matrix.h:94-112) can tell the
 * difference — it sums its two middle terms in operand order, which moves the last place of ~15 % of the scores;
 * every other distance, L2 included, is symmetric bit for bit. */
int zvb200_store_set_operand_order(zvb200_store *s, int query_first);
/* The same choice for the searches of ONE context: -1 = the store's (default), 0 / 1 as above.  FlatSearcher's
 * search_bf_by_p_keys_impl hands (query, feature) although its scans hand (feature, query)
 * (flat_searcher_context.h:1091-1093 vs flat_distance_matrix.h:53-55). */
int zvb200_ctx_set_operand_order(zvb200_ctx *c, int order);
/* Overwrite row `row` (< count) with a new vector and key: the in-place branch of
 * FlatStreamerEntity::add_vector_with_id (flat_streamer_entity.cc:969-985).  A key of
 * ZVB200_INVALID_KEY turns the row into a hole, any other key makes it live again. */
int zvb200_store_write(zvb200_store *s, uint64_t row, const void *vec, uint64_t key);
/* Copy one stored row back in the reference layout (IndexSearcher::get_vector). */
int zvb200_store_get_vector(const zvb200_store *s, uint64_t row, void *out, size_t cap);
/* Bytes of HBM one full scan must read (the roofline's algorithmic bytes). */
uint64_t zvb200_store_scan_bytes(const zvb200_store *s);

/* ---- context: replaces FlatSearcherContext / FlatStreamerContext
 * (flat_searcher_context.h:187-196, index_context.h:57-260): owns a CUDA stream, pinned
 * staging and device scratch.  One per calling thread. */
int zvb200_ctx_create(zvb200_store *s, zvb200_ctx **out);
int zvb200_ctx_destroy(zvb200_ctx *c);

/* ---- search: replaces FlatSearcher::search_impl(query, qmeta, count, ctx)
 * (flat_searcher.cc:186-211) + heap/sort/topk_to_result (index_document.h:250-261,
 * heap.h:103-115,173-175, flat_streamer_context.h:104-128).
 *  queries  : q consecutive elements in the STORED layout (post-reformer), HOST memory
 *  k        : 1..ZVB200_MAX_TOPK
 *  threshold: IndexContext::threshold() in internal score units; FLT_MAX = none.
 *             Rows with score > threshold (and NaN scores) are not returned.
 *  filter   : NULL, or a HOST bitmap of store_count bits, bit i set = drop row i
 *             (IndexFilter returning true; evaluated by the host adaptor).
 *  outputs  : q*k each, sorted ascending by (score, row index); unused slots hold
 *             key = ZVB200_INVALID_KEY, index = ZVB200_INVALID_INDEX, score = NaN.
 *             out_keys / out_index / out_counts may be NULL.
 * Tie-break (documented difference): among equal scores the LOWER row index
 * wins; the reference's choice at the k-th score is heap-order dependent.
 */
int zvb200_search(zvb200_ctx *c, const void *queries, uint32_t q, uint32_t k, float threshold,
                  const uint32_t *filter, uint64_t *out_keys, float *out_scores,
                  uint32_t *out_index, uint32_t *out_counts);

/* Same with DEVICE-resident queries / filter / outputs, asynchronous on the context's
 * stream; out_scores_dev and out_index_dev are q*k.  zvb200_ctx_sync() waits.
 * Searches of at most 8 queries that repeat on a context with the same arguments (for this form: the same device
 * pointers; the bytes behind them are read afresh) are recorded as one CUDA graph on the third call and replayed from
 * the fourth — zvb200_search, zvb200_ivf_search, zvb200_ivf_search_split and the _device forms.  Anything the store's
 * writers change, or a context buffer that has to grow, retires the recording; results never depend on it.
 * Environment ZVB200_GRAPH=0 turns it off. */
int zvb200_search_device(zvb200_ctx *c, const void *queries_dev, uint32_t q, uint32_t k,
                         float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                         uint32_t *out_index_dev);
int zvb200_ctx_sync(zvb200_ctx *c);
/* zvb200_search_device for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * Same arguments, same results bit for bit (prep_queries -> GEMV streaming scan -> merge_parts on the context's
 * stream), but every scratch buffer lives in one static device array of ZVB200_NOALLOC_SCRATCH_BYTES that each GPU
 * loading the library holds until exit; the call makes no cudaMalloc / cudaFree and no allocating library call.
 * Limits (ZVB200_ERR_INVALID_ARGUMENT beyond): q <= ZVB200_NOALLOC_MAX_QUERIES, k <= ZVB200_NOALLOC_MAX_TOPK,
 * and prepared queries + partial lists within the scratch (rows of up to ~16 KiB at the maximum q and k); plain
 * one-GPU stores only (ZVB200_ERR_UNSUPPORTED on a sharded one).  All calls share the array, which keeps the last
 * call's contents (the call re-initialises what it reads): calls that overlap on the GPU race — serialise them. */
#define ZVB200_NOALLOC_SCRATCH_BYTES 4194304
#define ZVB200_NOALLOC_MAX_QUERIES 8u
#define ZVB200_NOALLOC_MAX_TOPK 128u
int zvb200_search_device_noalloc(zvb200_ctx *c, const void *queries_dev, uint32_t q, uint32_t k,
                                 float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                                 uint32_t *out_index_dev);
/* Device time of the scan kernel(s) of the last search on this context, from CUDA events
 * recorded on its stream around them (the IndexContext::Profiler analogue,
 * src/include/zvec/core/framework/index_context.h:31-53).  Waits for that search. */
int zvb200_ctx_last_scan_ms(zvb200_ctx *c, float *ms);
/* The context's cudaStream_t (as void*) so callers can record events on it. */
void *zvb200_ctx_stream(zvb200_ctx *c);
/* Force a scan kernel: 0 auto, 1 GEMV scan (LDG), 2 tiled CUDA-core, 3 tcgen05, 4 bulk-async streaming scan,
 * 5 tcgen05 on the int8 codes themselves (kind::i8) where 3 would contract a QuantizedInteger L2 store's bf16
 * shadow operands. */
int zvb200_ctx_set_kernel(zvb200_ctx *c, int which);
/* Snapshot reads (FlatStreamer adds race with searches, tests/core/algorithm/flat/flat_streamer_test.cc:509-730):
 * the next searches on this context see rows [0, rows) only, and a filter bitmap handed to them covers exactly
 * those rows; 0 = every row (the default). */
int zvb200_ctx_set_row_limit(zvb200_ctx *c, uint64_t rows);
/* Name of the scan kernel the last search on this context used. */
const char *zvb200_ctx_last_kernel(const zvb200_ctx *c);
/* Operand kind of the last tcgen05 scan on this context: 0 kind::f16, 1 kind::tf32, 2 kind::i8,
 * 3 kind::f16 on the bf16 shadow operands of a QuantizedInteger L2 store; -1 none yet. */
int zvb200_ctx_last_mma_kind(const zvb200_ctx *c);

/* ---- distance functions: replaces ailego XxxDistanceMatrix<T,1,1>::Compute as handed out
 * by IndexMetric::distance() (src/core/metric/euclidean_metric.cc:785-851): out[j*n+i] =
 * dist(m[i], q[j]).  HOST pointers; computed on `device`. */
int zvb200_distance_matrix(int device, int metric, int dtype, uint32_t dim, const void *m,
                           uint64_t n, const void *q, uint32_t nq, float *out);

/* ---- multi-GPU shard merge: replaces the concat + std::sort + truncate of
 * CombinedVectorColumnIndexer::Search (src/db/index/column/vector_column/
 * combined_vector_column_indexer.cc:167-222).  `lists` G sorted per-shard results laid
 * out [g][q][k] (DEVICE memory, e.g. the NCCL all-gather output); index values are
 * shard-local and are offset by shard_base[g].  Output q*k sorted by (score, global row). */
int zvb200_merge_topk_device(zvb200_ctx *c, const float *scores_dev, const uint32_t *index_dev,
                             uint32_t g, const uint64_t *shard_base, uint32_t q, uint32_t k,
                             float *out_scores_dev, uint64_t *out_rows_dev);

/* ---- the multi-GPU layer across processes (one process per GPU: bench.py under torchrun, any multi-process
 * server).  Each rank holds a plain store with ITS contiguous row range and calls zvb200_search_sharded_device
 * collectively: scan -> exchange of the packed (score, row) keys -> merge by (score, global row), all on the
 * context's stream, every rank ends up with the merged result.  Exchange transport, agreed by all ranks at init:
 *   1  peer mailboxes: every rank stores its keys straight into a CUDA-IPC-mapped buffer of each peer over NVLink
 *      and raises a release flag; the merge kernel acquires the flags.  No collective launch (SURVEY §5's batch-1
 *      fast path); used whenever q * k <= 131072.
 *   0  one ncclAllGather of the u64 keys (CUDA IPC unavailable, ZVB200_COMM_MAILBOX=0, or a larger q * k).
 * NCCL is bound at run time (dlopen libnccl.so.2): the process's own copy if it has one (PyTorch's), else the
 * system's.  zvb200_comm_unique_id on one rank, the 128 bytes carried to the others by the caller (any side
 * channel), then zvb200_comm_init on every rank (collective).  Before zvb200_comm_destroy the caller makes sure no
 * peer is still inside a search (a barrier). */
#define ZVB200_COMM_ID_BYTES 128
int zvb200_comm_unique_id(void *id128);
int zvb200_comm_init(zvb200_ctx *c, int rank, int world, const void *id128, zvb200_comm **out);
int zvb200_comm_destroy(zvb200_comm *cm);
/* bases[r] = first global row of rank r's shard (ascending with r). */
int zvb200_comm_set_shard_bases(zvb200_comm *cm, const uint64_t *bases);
int zvb200_comm_transport(const zvb200_comm *cm);
/* ZVB200_ERR_RUNTIME if an exchange timed out (a peer never delivered) since init; synchronises the device. */
int zvb200_comm_status(zvb200_comm *cm);
/* queries / filter / outputs in DEVICE memory; out_scores_dev f32 [q][k], out_rows_dev u64 [q][k] (global rows,
 * ZVB200_INVALID_KEY / NaN padded).  zvb200_ctx_sync() waits. */
int zvb200_search_sharded_device(zvb200_ctx *c, zvb200_comm *cm, const void *queries_dev, uint32_t q, uint32_t k,
                                 float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                                 uint64_t *out_rows_dev);

/* ---- multi-vector fusion: replaces RrfReRanker.rerank / WeightedReRanker.rerank
 * (python/zvec/extension/multi_vector_reranker.py:62-89,135-174) applied by the query executor to
 * the per-field results of a multi-vector query (python/zvec/executor/query_executor.py:178-206,
 * 276-287).  HOST buffers.  keys [F][q][k] (document keys as returned by zvb200_search), counts
 * [F][q] valid entries per list, lists in field order.  Output per query: topn (key, fp64 score)
 * sorted like heapq.nlargest (score descending, ties in first-seen order), padded with
 * ZVB200_INVALID_KEY / NaN; out_counts[q] = documents returned.  F*k <= 1536.
 *   rrf:      score = sum 1 / (rank_constant + rank + 1)
 *   weighted: score = sum normalize(s) * weights[f]; fuse_metric ZVB200_METRIC_L2 | _L2SQ:
 *             1 - 2 atan(s)/pi, ZVB200_METRIC_IP: 0.5 + atan(s)/pi, ZVB200_METRIC_COSINE: 1 - s/2;
 *             s = the user-facing score (Doc.score): pass negate = 1 to hand in the internal
 *             InnerProduct scores of zvb200_search (IndexMetric::normalize flips their sign,
 *             src/core/metric/inner_product_metric.cc:377-379). */
int zvb200_fuse_rrf(zvb200_ctx *c, const uint64_t *keys, const uint32_t *counts, uint32_t nfields,
                    uint32_t q, uint32_t k, uint32_t topn, uint32_t rank_constant,
                    uint64_t *out_keys, double *out_scores, uint32_t *out_counts);
int zvb200_fuse_weighted(zvb200_ctx *c, const uint64_t *keys, const float *scores,
                         const uint32_t *counts, uint32_t nfields, uint32_t q, uint32_t k,
                         uint32_t topn, int fuse_metric, const double *weights, int negate,
                         uint64_t *out_keys, double *out_scores, uint32_t *out_counts);

/* ---- IVF: replaces IVFSearcher::search_impl (src/core/algorithm/ivf/ivf_searcher.cc:
 * 183-250): coarse search over centroids (IVFCentroidIndex::search, ivf_centroid_index.cc:
 * 273-295) + IVFEntity::search per list (ivf_entity.cc:662-716) under the scan budget of
 * IVFSearcherContext::update (ivf_searcher_context.h:61-79).
 * The store must hold the rows grouped by list: rows [list_offsets[l], list_offsets[l+1])
 * belong to list l.  centroids: nlist elements of the store's dtype/metric layout, HOST. */
int zvb200_ivf_attach(zvb200_store *s, const void *centroids, uint32_t nlist,
                      const uint64_t *list_offsets);
/* Posting lists an IVFBuilder quantised (proxima.ivf.builder.quantizer_class = Int8QuantizerConverter,
 * ivf_builder.cc:738-800): the lists are int8 rows while the centroid index keeps its own layout (fp32 rows, for
 * InnerProduct in the Mips-converted space of one more dimension block, ivf_centroid_index.cc:338-384).
 * zvb200_ivf_attach_typed = zvb200_ivf_attach with the centroids' own (dtype, metric, dim).  Such a store is searched
 * with zvb200_ivf_search_split: coarse_queries in the centroids' layout (what IVFCentroidIndex::search hands its
 * searcher after reformer_->transform, ivf_centroid_index.cc:273-295), list_queries in the lists' layout (what
 * IVFEntity::transform returns, ivf_searcher.cc:211-214); everything else as zvb200_ivf_search. */
int zvb200_ivf_attach_typed(zvb200_store *s, const void *centroids, uint32_t nlist, const uint64_t *list_offsets,
                            int c_dtype, int c_metric, uint32_t c_dim);
int zvb200_ivf_search_split(zvb200_ctx *c, const void *coarse_queries, const void *list_queries, uint32_t q, uint32_t k,
                            uint32_t nprobe, uint64_t max_scan_count, float threshold, const uint32_t *filter_bitmap,
                            uint64_t *out_keys, float *out_scores, uint32_t *out_index, uint32_t *out_counts);
/* Scores leave the scan multiplied by a constant — `distances[k] * norm_val` of IVFEntity::search
 * (ivf_entity.cc:601,650,675,715): norm_val is the quantiser's global 1/scale (IVFEntity::load, ivf_entity.cc:541-564) or
 * the scale of the row's posting list (quantize_by_centroid, IVFEntity::inverted_list_normalize_value, ivf_entity.h:
 * 204-224).  list_scales: nlist values, HOST, needs an attached IVF structure (NULL: `scale` for every row; 1 = off).
 * The product is what the threshold gates and the top-k compares, as in the reference's heap.  A scaled store is
 * scanned by the CUDA-core kernels (GEMV scan, posting-list scans) at any batch.  zvb200_ivf_attach* drops list scales;
 * a store that holds list scales refuses appends (ZVB200_ERR_UNSUPPORTED): they cover the rows that were attached. */
int zvb200_store_set_score_scale(zvb200_store *s, float scale, const float *list_scales);
/* Sharded IVF (each GPU holds a slice of EVERY posting list): list_sizes[l] = rows of list l in the
 * whole index.  The scan budget of zvb200_ivf_search then counts these, so every shard visits exactly the lists
 * the unsharded index would (IVFSearcher::search_impl's loop header, ivf_searcher.cc:216-219).  NULL = unsharded. */
int zvb200_ivf_set_budget_sizes(zvb200_store *s, const uint64_t *list_sizes);
/* nprobe = max(1, round(nlist*scan_ratio)); max_scan_count = max(bf_threshold,
 * ceil(N*scan_ratio)): lists are visited in coarse order while the rows scanned so far
 * are < max_scan_count. */
int zvb200_ivf_search(zvb200_ctx *c, const void *queries, uint32_t q, uint32_t k,
                      uint32_t nprobe, uint64_t max_scan_count, float threshold,
                      const uint32_t *filter_bitmap, uint64_t *out_keys, float *out_scores,
                      uint32_t *out_index, uint32_t *out_counts);
/* The same search with queries, filter and outputs on the device, asynchronous on the context's stream (the
 * multi-GPU layer: per-shard results stay on the GPU for the exchange). */
int zvb200_ivf_search_device(zvb200_ctx *c, const void *queries_dev, uint32_t q, uint32_t k,
                             uint32_t nprobe, uint64_t max_scan_count, float threshold,
                             const uint32_t *filter_dev, float *out_scores_dev, uint32_t *out_index_dev);
/* Sharded IVF, the merge on the device: list_global_base[l] = row of the WHOLE index (lists in order, each list's
 * rows in order) at which this store's slice of list l starts.  zvb200_ivf_global_rows_device maps local result
 * rows to rows of the whole index (ZVB200_INVALID_INDEX -> ZVB200_INVALID_KEY); after the all-gather,
 * zvb200_merge_topk_rows_device merges G lists [g][q][k] of (score, global row) by that pair — the result of the
 * unsharded zvb200_ivf_search, bit for bit (CombinedVectorColumnIndexer::Search's concat + sort + truncate,
 * combined_vector_column_indexer.cc:167-222, with the row-index tie-break kept global). */
int zvb200_ivf_set_global_bases(zvb200_store *s, const uint64_t *list_global_base);
int zvb200_ivf_global_rows_device(zvb200_ctx *c, const uint32_t *index_dev, uint64_t count,
                                  uint64_t *out_rows_dev);
int zvb200_merge_topk_rows_device(zvb200_ctx *c, const float *scores_dev, const uint64_t *rows_dev,
                                  uint32_t g, uint32_t q, uint32_t k, float *out_scores_dev,
                                  uint64_t *out_rows_dev);
/* threshold / filter_bitmap as in zvb200_search (bit i = drop stored row i, list order): a filtered list
 * still counts whole against the scan budget, like IVFEntity::search(list, query, filter, ...)
 * (ivf_entity.cc:587-659). */

/* ---- query reformers on the device: replace IndexReformer::transform, the per-query CPU work in front of a
 * search of a converted index (called from core_interface::Index::_prepare_for_search).  The reformer is the
 * one the store's (dtype, metric) implies:
 *   fp16                    HalfFloatReformer          src/core/quantizer/half_float_reformer.cc
 *   cosine fp32 / fp16      CosineReformer             src/core/quantizer/cosine_reformer.cc:66-112
 *   QuantizedInteger int8   IntegerStreamingReformer   src/core/quantizer/integer_quantizer_reformer.cc:308-370
 *                           (RecordQuantizer::quantize_record, record_quantizer.h:24-83)
 *   fp32 L2 / IP            none (copied)
 * queries_fp32: q x dim floats, dim = the ORIGINAL dimension (the store's dimension minus its tail).
 * out: q stored elements (the store's elem bytes each), i.e. what transform() returns.  int8 records are
 * bit-identical to the reference's; fp16 conversions too; the cosine norm may differ from the reference's
 * SIMD sum in the last place.  ZVB200_ERR_UNSUPPORTED for plain int8 stores (they take int8 queries). */
int zvb200_reform_queries(zvb200_ctx *c, const float *queries_fp32, uint32_t q, uint32_t dim, void *out);
int zvb200_reform_queries_device(zvb200_ctx *c, const float *queries_fp32_dev, uint32_t q, uint32_t dim,
                                 void *out_dev);
/* zvb200_search for ORIGINAL fp32 queries: reformer + scan in one call, nothing returns to the host in
 * between (core_interface::Index::Search = _prepare_for_search + _dense_search, index.cc:380-410). */
int zvb200_search_fp32(zvb200_ctx *c, const float *queries_fp32, uint32_t q, uint32_t dim, uint32_t k,
                       float threshold, const uint32_t *filter, uint64_t *out_keys, float *out_scores,
                       uint32_t *out_index, uint32_t *out_counts);

/* ---- brute force over given rows: replaces FlatStreamer::search_bf_by_p_keys_impl
 * (src/core/algorithm/flat/flat_streamer.cc:347-389) with stored-row indices for keys (the caller owns the
 * key -> row map).  rows: [q][m] HOST, ZVB200_INVALID_INDEX / out-of-range entries are skipped like keys the
 * entity does not hold; deleted and filtered rows are skipped; a row listed twice is scored twice, as there.
 * queries in the store's layout.  m <= ZVB200_MAX_CANDIDATES.  Outputs as zvb200_search. */
int zvb200_search_rows(zvb200_ctx *c, const void *queries, uint32_t q, const uint32_t *rows, uint32_t m,
                       uint32_t k, float threshold, const uint32_t *filter, uint64_t *out_keys,
                       float *out_scores, uint32_t *out_index, uint32_t *out_counts);

/* ---- refiner: replaces core_interface::Index::Search with a refiner_param (src/core/interface/index.cc:
 * 408-445): a coarse search for floor(k * scale_factor) rows (scale_factor 0 = 1, index.cc:794-801) on the
 * quantised index `coarse`, then the exact brute force over those candidates on the reference index `exact`
 * (a Flat index in the reference, index.cc:417-420), best k.  Both stores hold the same documents in the same
 * row order on one device; the candidates never leave the device.  queries_fp32 / dim as in
 * zvb200_reform_queries; filter applies to both stages (index.cc:439); results carry the exact store's keys. */
int zvb200_search_refined(zvb200_ctx *coarse, zvb200_ctx *exact, const float *queries_fp32, uint32_t q,
                          uint32_t dim, uint32_t k, float scale_factor, const uint32_t *filter,
                          uint64_t *out_keys, float *out_scores, uint32_t *out_index,
                          uint32_t *out_counts);

/* ---- k-means: replaces KmeansCluster::cluster + label (src/core/algorithm/cluster/
 * kmeans_cluster.cc:469-535, 571-610) as IVFBuilder trains its centroids with (ivf_builder.cc:442-485).
 * rows: n x dim fp32 (HOST for zvb200_kmeans, DEVICE for zvb200_kmeans_device).  centroids: k x dim fp32 HOST,
 * the initial centroids going in (the caller samples them, KmeansCluster::init_centroids is a reservoir sample
 * of the rows), the trained ones coming out.  Each iteration assigns every row to its nearest centroid under
 * `metric` (ZVB200_METRIC_L2SQ / L2 / IP; the strict '<' of LinearSeeker::seek, linear_seeker.cc:29-37 = the
 * lower centroid wins a tie), accumulates per-cluster sums in double and sets centroid = (float)(sum / count)
 * (vector_mean.h:204-226; an empty cluster becomes NaN and is never chosen again, as there).  Stops after
 * max_iterations or when |cost - previous cost| < epsilon (kmeans_cluster.cc:519-529).  Optional outputs:
 * labels[n] = nearest of the RETURNED centroids (label()), follows[k] = rows per cluster and *sse = cost of the
 * last assignment pass, *iterations = passes run.  Nothing is cut, but as PurgeCentroids does at the end of cluster()
 * (kmeans_cluster.cc:363-383, 533) the clusters nobody follows are swapped to the tail of centroids / follows. */
int zvb200_kmeans(int device, int metric, const float *rows, uint64_t n, uint32_t dim, float *centroids,
                  uint32_t k, uint32_t max_iterations, double epsilon, uint32_t *labels, uint64_t *follows,
                  double *sse, uint32_t *iterations);
int zvb200_kmeans_device(int device, int metric, const float *rows_dev, uint64_t n, uint32_t dim,
                         float *centroids, uint32_t k, uint32_t max_iterations, double epsilon,
                         uint32_t *labels, uint64_t *follows, double *sse, uint32_t *iterations);

/* ---- IVF build, after the centroids are trained (second half).
 * zvb200_ivf_assign replaces IVFBuilder::label -> IVFCentroidIndex::search_nearest_centroid
 * (src/core/algorithm/ivf/ivf_builder.h:243-262, ivf_centroid_index.cc:297-318: a top-1 search over the centroids per
 * row): labels[i] = nearest of the nlist centroids to row i under `metric` (the lower centroid wins a tie; a row no
 * centroid can score follows centroid 0).  rows / centroids: HOST, stored layout of (dtype, metric, dim).
 * zvb200_ivf_pack replaces the list-by-list walk of IVFBuilder::dump_index (ivf_builder.cc:672-683) over labels_:
 * order[] = row numbers posting list by posting list, ascending inside a list; list_offsets[l] = position in order[]
 * where list l starts (InvertedListMeta::id_offset, ivf_index_format.h:41-47), list_offsets[nlist] = n.  A host
 * adaptor hands order[] to the reference's IVFDumper (ivf_dumper.h:191-209) — integration/gpu_ivf_builder.cc — or
 * appends the rows in that order to a store and calls zvb200_ivf_attach. */
int zvb200_ivf_assign(int device, int dtype, int metric, uint32_t dim, const void *centroids, uint32_t nlist,
                      const void *rows, uint64_t n, uint32_t *labels);
int zvb200_ivf_pack(int device, const uint32_t *labels, uint64_t n, uint32_t nlist, uint32_t *order,
                    uint64_t *list_offsets);

#ifdef __cplusplus
}
#endif
#endif /* ZVEC_B200_H_ */
