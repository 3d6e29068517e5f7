// Host-side objects behind the C ABI handles (include/zvec_b200.h): the device-resident store, the search context,
// and the helpers api.cu and multi.cu share.  Not part of the ABI.
#pragma once
#include <float.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <shared_mutex>

#include "../../include/zvec_b200.h"
#include "internal.h"

int zvb_set_err(int code, const char *what, const char *detail = "");
int zvb_cuda_fail(cudaError_t e, const char *what);
#define set_err zvb_set_err
#define cuda_fail zvb_cuda_fail
#define CU(x)                                        \
  do {                                               \
    cudaError_t _e = (x);                            \
    if (_e != cudaSuccess) return cuda_fail(_e, #x); \
  } while (0)

using zvb::RowsView;
using zvb::QueryView;
using zvb::u64;

struct zvb200_ctx;

struct zvb200_store {
  int device = 0, dtype = 0, metric = 0, sm_count = 148;
  uint32_t dim = 0, elem_bytes = 0, d = 0, vec_bytes = 0, row_stride = 0, chunks = 0, tail_off = 0;
  bool has_tail = false, sealed = false;
  bool query_first = false;       // zvb200_store_set_operand_order
  uint8_t *rows = nullptr;
  float4 *tails = nullptr;
  float *norms = nullptr;
  uint8_t *shadow = nullptr;      // QuantizedInteger L2: bf16 operands of the tensor-core filter (built at seal)
  uint32_t shadow_k = 0;
  uint64_t shadowed = 0;
  float *shadow_mmax_dev = nullptr;  // running maximum of the rows' margin shares
  float shadow_mrow_max = 0.f;
  uint32_t *drop = nullptr;       // device bitmap (allocated lazily)
  uint64_t drop_words = 0;       // words allocated on the device
  uint64_t drop_synced = 0;      // words of drop_host the device copy holds
  uint64_t n = 0, cap = 0, normed = 0;
  std::vector<uint64_t> keys;     // empty = identity
  std::vector<uint32_t> drop_host;
  bool has_drop = false, drop_dirty = false;
  cudaStream_t stream = nullptr;
  uint8_t *staging = nullptr;     // device staging for host appends
  size_t staging_bytes = 0;
  // IVF
  zvb200_store *centroids = nullptr;
  std::vector<uint64_t> list_offsets;
  uint32_t *list_off_dev = nullptr;  // [nlist+1] first row of each posting list
  uint32_t *budget_dev = nullptr;    // [nlist] list sizes in the whole index (sharded IVF), or null
  uint64_t *gbase_dev = nullptr;     // [nlist] row of the whole index where this store's slice of each list starts
  uint32_t nlist = 0;
  uint64_t gen = 0;                  // bumped under the writers' lock: whatever a recorded search was built from may have changed
  float *row_scale = nullptr;        // [n] zvb200_store_set_score_scale with per-list values, expanded per row
  float score_scale = 1.f;
  bool coarse_typed = false;         // centroids in a layout of their own (zvb200_ivf_attach_typed): split queries only
  // writers (append / seal / write / set_deleted / attach) exclusive; the host-synchronous searches shared, from the
  // first look at the store to the stream sync that ends them: an append that re-allocates the rows (store_grow)
  // cannot free them under a search another thread has in flight (FlatStreamer adds race with searches,
  // tests/core/algorithm/flat/flat_streamer_test.cc:509-730)
  std::shared_mutex mu;
  struct Write {  // the writers' lock
    std::lock_guard<std::shared_mutex> g;
    explicit Write(zvb200_store *s) : g(s->mu) { ++s->gen; }
  };
  // Sharded parent (zvb200_store_create_sharded over several devices of this process): contiguous row
  // ranges, shard g holds rows [g * shard_rows, (g + 1) * shard_rows) of the store; the parent itself holds no rows,
  // only the layout, the row count and the keys.  shard_rows is fixed by the first zvb200_store_reserve.
  std::vector<zvb200_store *> shards;
  uint64_t shard_rows = 0;
  uint64_t *shard_base_dev = nullptr;  // [shards] on the root device

  RowsView view() const {
    RowsView v;
    v.rows = rows; v.n = n; v.row_stride = row_stride; v.chunks = chunks; v.d = d;
    v.dtype = dtype; v.metric = metric; v.tails = tails; v.norms = norms;
    v.drop = has_drop ? drop : nullptr;
    v.query_first = query_first;
    v.row_scale = row_scale; v.score_scale = score_scale;
    if (shadow && shadowed >= n) { v.shadow = shadow; v.shadow_k = shadow_k; v.shadow_mrow_max = shadow_mrow_max; }
    return v;
  }
};

struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = std::max(bytes, (size_t)4096);
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
struct HostBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    size_t want = std::max(bytes, (size_t)4096);
    cudaError_t e = cudaMallocHost(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

// A small host-side search recorded as ONE CUDA graph (H2D of the queries, every kernel of the search, D2H of the
// results): the second call of a context with the same arguments records it, later ones replay it with one
// cudaGraphLaunch.  The key holds everything the launches were built from; the store's generation (bumped by every
// writer) and a signature of the context's own buffers (pointer and size of each: a re-allocated buffer) retire a
// stale recording.  Recording costs a few hundred microseconds, so it waits for the THIRD call with one key, and a
// context whose recordings keep being retired before they are replayed (searches interleaved with ingest) stops
// recording.
struct GraphKey {  // compared with memcmp: 8-byte members first, an even number of 4-byte ones, no padding
  const void *store;
  const void *p0, *p1, *p2, *p3;  // the device-pointer forms: queries, filter, outputs
  uint64_t gen, epoch, rows, max_scan;
  uint32_t q, k, nprobe, thr_bits, mma_group_cap;
  int fn;                 // 1 zvb200_search, 2 zvb200_ivf_search, 3 _split, 4 / 5 the _device forms
  int force_kernel, operand_order;
};
static_assert(sizeof(GraphKey) == 5 * sizeof(void *) + 4 * 8 + 8 * 4, "GraphKey must not have padding");
struct ReplayGraph {
  GraphKey key;
  int seen = 0, churn = 0;
  uint32_t replays = 0;
  bool failed = false;
  cudaGraphExec_t exec = nullptr;
  uint32_t launches = 0;
  const char *last_kernel = "none";
  int last_mma_kind = -1;
  void reset() {  // another key: what was recorded (if anything) is dropped
    if (exec) {
      cudaGraphExecDestroy(exec);
      churn = replays >= 8 ? 0 : churn + 1;  // a recording that paid for itself does not count
      if (churn >= 16) failed = true;        // (for good: this context's searches do not repeat)
    }
    exec = nullptr; seen = 0; launches = 0; replays = 0;
  }
};

struct zvb200_ctx {
  zvb200_store *s = nullptr;
  int device = 0;  // copy: a context may outlive its store (destruction order is the caller's)
  cudaStream_t stream = nullptr;
  DevBuf qshadow;
  DevBuf qin, qin2, qf32, qraw, qtails, qnorms, part, out_scores, out_index, filter, misc, gtau;
  DevBuf coarse_scores, coarse_index, seg_prefix, seg_row, seg_count, seg_list, lm;  // IVF
  DevBuf qsrc, cand;  // fp32 queries before the reformer; candidate rows of the refiner
  QueryView last_qv;
  HostBuf h_q, h_q2, h_scores, h_index;
  int force_kernel = 0;
  uint32_t mma_group_cap = 0;  // queries per tcgen05 launch (0: ZVB200_MMA_GROUP or 1024)
  int last_mma_kind = -1;
  const char *last_kernel = "none";
  cudaEvent_t ev_scan0 = nullptr, ev_scan1 = nullptr;  // bracket the scan kernel(s) of the last search
  bool scan_timed = false;
  ReplayGraph rg;
  bool capturing = false;  // inside cudaStreamBeginCapture: nothing may allocate
  int operand_order = -1;  // zvb200_ctx_set_operand_order: -1 the store's, 0 feature first, 1 query first
  uint64_t row_limit = 0;  // zvb200_ctx_set_row_limit: searches see rows [0, row_limit) only (0 = every row)
  // a context of a sharded parent: one context per shard (on the shard's device), the exchange buffers on the root
  std::vector<zvb200_ctx *> sub;
  std::vector<cudaEvent_t> sub_done;
  cudaEvent_t ev_ready = nullptr;
  DevBuf xkeys, gather, out_rows;  // packed local keys; [G][q][k] gathered keys; merged global rows
  HostBuf h_rows;
  //! pointer and size of every buffer a recorded search may have been built from
  uint64_t buffers_signature() const {
    uint64_t h = 1469598103934665603ull;
    auto mix = [&h](const void *p, size_t cap) {
      h = (h ^ (uint64_t)(uintptr_t)p) * 1099511628211ull;
      h = (h ^ (uint64_t)cap) * 1099511628211ull;
    };
    const DevBuf *d[] = {&qshadow, &qin, &qin2, &qf32, &qraw, &qtails, &qnorms, &part, &out_scores, &out_index, &filter,
                         &misc, &gtau, &coarse_scores, &coarse_index, &seg_prefix, &seg_row, &seg_count, &seg_list, &lm,
                         &qsrc, &cand, &xkeys, &gather, &out_rows};
    for (const DevBuf *b : d) mix(b->p, b->cap);
    const HostBuf *hb[] = {&h_q, &h_q2, &h_scores, &h_index, &h_rows};
    for (const HostBuf *b : hb) mix(b->p, b->cap);
    return h;
  }
};

// the events that bracket a search's scan kernels; inside a recording they become event-record nodes of the graph
// (cudaEventRecordExternal), so zvb200_ctx_last_scan_ms keeps working for replayed searches
inline cudaError_t zvb_scan_event(zvb200_ctx *c, cudaEvent_t ev) {
  return c->capturing ? cudaEventRecordWithFlags(ev, c->stream, cudaEventRecordExternal) : cudaEventRecord(ev, c->stream);
}

// ---- sharded parents (multi.cu), called by the C ABI entry points of api.cu when store->shards is not empty ----
int zvb_sharded_destroy(zvb200_store *s);
int zvb_sharded_reserve(zvb200_store *s, uint64_t rows);
int zvb_sharded_append(zvb200_store *s, const void *rows, uint64_t n, const uint64_t *keys, bool on_device);
int zvb_sharded_seal(zvb200_store *s);
int zvb_sharded_set_deleted(zvb200_store *s, uint64_t row, int deleted);
int zvb_sharded_write(zvb200_store *s, uint64_t row, const void *vec, uint64_t key);
int zvb_sharded_get_vector(const zvb200_store *s, uint64_t row, void *out, size_t cap);
int zvb_sharded_ctx_create(zvb200_store *s, zvb200_ctx *c);
void zvb_sharded_ctx_destroy(zvb200_ctx *c);
// queries / filter / rows on the host (on_host) or on the root device; outputs on the root device
int zvb_sharded_search(zvb200_ctx *c, const void *queries, bool on_host, uint32_t q, uint32_t k, float threshold,
                       const uint32_t *filter, const uint32_t *rows /* search_rows: [q][m] global rows, host */,
                       uint32_t m, float *out_scores_dev, uint32_t *out_index_dev);


// ---- shared by api.cu (plain stores) and multi.cu (sharded stores, communicators) ----
int zvb_search_core(zvb200_ctx *c, zvb200_store *s, const void *queries_dev, uint32_t q, uint32_t k, float threshold,
                    const uint32_t *filter_dev, float *out_scores_dev, uint32_t *out_index_dev, u64 *out_keys_dev,
                    uint64_t row_begin, uint64_t row_end);
int zvb_rows_core(zvb200_ctx *c, zvb200_store *s, const void *queries_dev, uint32_t q, const uint32_t *rows_dev,
                  uint32_t m, uint32_t k, float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                  uint32_t *out_index_dev, u64 *out_keys_dev);
int zvb_check_search_args(zvb200_ctx *c, const void *queries, uint32_t q, uint32_t k);
int zvb_check_row_limit(uint64_t rows);
// the store's rows as this context's searches see them (operand order override)
inline RowsView zvb_view_for(const zvb200_ctx *c, const zvb200_store *s) {
  RowsView v = s->view();
  if (c->operand_order >= 0) v.query_first = c->operand_order != 0;
  return v;
}
