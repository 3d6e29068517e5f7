// zvec index plugin: GpuFlatSearcher / GpuFlatStreamer / GpuIVFSearcher / GpuIVFStreamer — the reference's
// IndexSearcher / IndexStreamer interfaces served by the B200 scan behind include/zvec_b200.h.
//
// Host adaptor (C++, ours).  It is compiled against the reference's PUBLIC framework headers
// (src/include/zvec/core/framework/*.h, used with -I where they lie, never copied) and
// registered with the reference's own macros INDEX_FACTORY_REGISTER_SEARCHER / _STREAMER
// (src/include/zvec/core/framework/index_factory.h:237-242), so
//   IndexFactory::CreateSearcher("GpuFlatSearcher")
// — and, after zvb200_install_overrides(), IndexFactory::CreateSearcher("FlatSearcher") —
// returns it; it is loaded like any zvec plugin: IndexPlugin::load(path) = dlopen
// (src/core/framework/index_plugin.cc:20-35).
//
// Division of labour (composition): every class embeds the STOCK class it stands in for, created under
// a name that is never overridden ("FlatSearcher32", "FlatStreamer32", IVFSearcher by type).  The stock object owns
// what is not arithmetic — file validation and its error codes, key maps, providers / get_vector / fetch_vector,
// persistence, group-by bookkeeping.  Every scan (search_impl, search_bf_impl, search_bf_by_p_keys_impl) runs on the
// GPU through the C ABI; the only searches the stock object still answers are group-by searches (SURVEY 8b allows
// delegating those) and IVF indexes whose layout the device store cannot hold, each logged once.
#include <zvec/core/framework/index_context.h>
#include <zvec/core/framework/index_error.h>
#include <zvec/core/framework/index_factory.h>
#include <zvec/core/framework/index_helper.h>
#include <zvec/core/framework/index_logger.h>
#include <zvec/core/framework/index_meta.h>
#include <zvec/core/framework/index_searcher.h>
#include <zvec/core/framework/index_streamer.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <limits>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <zvec/core/framework/index_segment_storage.h>
#include "core/algorithm/ivf/ivf_searcher.h"  // the stock IVF classes, used by type (composition)

#include "../include/zvec_b200.h"

namespace zvec {
namespace core {

namespace {

// segment ids of a Flat index file (src/core/algorithm/flat/flat_utility.h:32-34)
const std::string kSegKeys("flat.keys");
const std::string kSegFeatures("flat.features");
// QuantizedIntegerMetric's origin metric parameter (src/core/metric/quantized_integer_metric.cc:40-60)
const std::string kOriginMetric("proxima.quantized_integer.metric.origin_metric_name");
// proxima.gpu.* keys read through the same ailego::Params as every other searcher parameter
const std::string kParamDevice("proxima.gpu.device");    // one CUDA device (default 0)
const std::string kParamDevices("proxima.gpu.devices");  // "0,1,2,3": rows sharded over these devices (GpuFlatSearcher)

// IndexMeta (metric name [+ origin metric], data type) -> ZVB200_METRIC_* ; <0 = unsupported,
// like InnerProductMetric::init for types it has no kernel for (inner_product_metric.cc:259-281)
int metric_of(const IndexMeta &meta) {
  const std::string &m = meta.metric_name();
  if (m == "SquaredEuclidean") return ZVB200_METRIC_L2SQ;
  if (m == "Euclidean") return ZVB200_METRIC_L2;
  if (m == "InnerProduct") return ZVB200_METRIC_IP;
  if (m == "Cosine") return ZVB200_METRIC_COSINE;
  if (m == "QuantizedInteger") {
    std::string o;
    meta.metric_params().get(kOriginMetric, &o);
    if (o == "SquaredEuclidean" || o.empty()) return ZVB200_METRIC_QI8_L2SQ;
    if (o == "InnerProduct") return ZVB200_METRIC_QI8_IP;
    if (o == "Cosine") return ZVB200_METRIC_QI8_COSINE;
  }
  return -1;
}

int dtype_of(const IndexMeta &meta) {
  switch (meta.data_type()) {
    case IndexMeta::DataType::DT_FP16: return ZVB200_DT_FP16;
    case IndexMeta::DataType::DT_FP32: return ZVB200_DT_FP32;
    case IndexMeta::DataType::DT_INT8: return ZVB200_DT_INT8;
    default: return -1;
  }
}

std::vector<int> parse_devices(const ailego::Params &params) {
  std::vector<int> out;
  std::string s;
  if (params.get(kParamDevices, &s) && !s.empty()) {
    size_t pos = 0;
    while (pos <= s.size()) {
      size_t e = s.find(',', pos);
      if (e == std::string::npos) e = s.size();
      if (e > pos) out.push_back(std::atoi(s.substr(pos, e - pos).c_str()));
      pos = e + 1;
    }
  }
  if (out.empty()) {
    int d = 0;
    params.get(kParamDevice, &d);
    out.push_back(d);
  }
  return out;
}

void log_delegated_once(std::atomic<bool> *flag, const char *cls, const char *why) {
  bool expected = false;
  if (flag->compare_exchange_strong(expected, true))
    LOG_WARN("%s: %s is answered by the stock CPU class (logged once)", cls, why);
}

}  // namespace

/*! Search context shared by the four classes: the IndexContext contract (index_context.h:57-260 — topk, filter,
 *  threshold, group params, fetch_vector, result(i), group_result(i), magic) over a zvb200_ctx (CUDA stream + pinned
 *  staging + device scratch) and the stock class's own context for what the stock object answers.
 */
class GpuContext : public IndexContext {
 public:
  explicit GpuContext(IndexContext::Pointer inner) : inner_(std::move(inner)) {}
  ~GpuContext() override {
    if (ctx_) zvb200_ctx_destroy(ctx_);
  }
  void set_topk(uint32_t topk) override {
    topk_ = topk;
    if (inner_) inner_->set_topk(topk);
  }
  uint32_t topk() const override { return topk_; }
  //! FlatSearcherContext::set_group_params (flat_searcher_context.h:178-182): group_num > 0 = group-by search
  void set_group_params(uint32_t group_num, uint32_t group_topk) override {
    group_num_ = group_num;
    if (inner_) inner_->set_group_params(group_num, group_topk);
  }
  void set_fetch_vector(bool v) override {
    fetch_vector_ = v;
    if (inner_) inner_->set_fetch_vector(v);
  }
  bool fetch_vector() const override { return fetch_vector_; }
  const IndexDocumentList &result(void) const override { return results_[0]; }
  const IndexDocumentList &result(size_t i) const override { return results_[i]; }
  IndexDocumentList *mutable_result(size_t i) override { return &results_[i]; }
  const IndexGroupDocumentList &group_result(void) const override {
    return inner_ ? inner_->group_result() : IndexContext::group_result();
  }
  const IndexGroupDocumentList &group_result(size_t i) const override {
    return inner_ ? inner_->group_result(i) : IndexContext::group_result();
  }
  uint32_t magic(void) const override { return magic_; }

  bool group_by_search(void) const { return group_num_ > 0; }
  zvb200_ctx *handle() const { return ctx_; }
  IndexContext::Pointer &inner() { return inner_; }
  //! (re)bind to a device store; results are cleared like every reference context's reset(owner)
  void bind(zvb200_store *store, uint32_t magic) {
    if (ctx_) zvb200_ctx_destroy(ctx_);
    ctx_ = nullptr;
    magic_ = magic;
    if (store) zvb200_ctx_create(store, &ctx_);
    for (auto &r : results_) r.clear();
  }
  //! `count` empty result lists and cleared per-query statistics (reset_results of the reference contexts)
  void begin(size_t count) {
    if (results_.size() < count) results_.resize(count);
    for (size_t i = 0; i < count; ++i) results_[i].clear();
    stats_vec_.resize(count);
    for (auto &s : stats_vec_) s.clear();
    const size_t nres = count * (size_t)topk_;
    keys.resize(nres); scores.resize(nres); index.resize(nres); counts.resize(count);
  }
  //! per-query statistics (FlatStreamerContext::mutable_stats, flat_streamer_context.h:219-222): filled from the
  //! host after the kernel with what the reference's scan loops count (flat_streamer_entity.cc:246,306-314)
  Stats *mutable_stats(size_t idx = 0) { return &stats_vec_[idx]; }
  const Stats &stats(size_t idx = 0) const { return stats_vec_[idx]; }
  void fill_stats(size_t count, size_t calced, size_t filtered) {
    for (size_t i = 0; i < count && i < stats_vec_.size(); ++i) {
      stats_vec_[i].set_dist_calced_count(calced);
      stats_vec_[i].set_filtered_count(filtered);
    }
  }
  //! what the caller set on this context, mirrored onto the stock context before the stock object answers:
  //! a filter / group-by / threshold that was cleared here is cleared there too
  void sync_inner(void) {
    if (!inner_) return;
    inner_->set_topk(topk_);
    if (this->filter().is_valid()) inner_->set_filter(this->filter()); else inner_->reset_filter();
    if (this->group_by().is_valid()) inner_->set_group_by(this->group_by()); else inner_->reset_group_by();
    inner_->set_threshold(this->threshold());  // stock contexts carry no metric: the value is taken as it is
    inner_->set_fetch_vector(fetch_vector_);
  }
  void take_inner_results(size_t count) {
    if (results_.size() < count) results_.resize(count);
    for (size_t i = 0; i < count; ++i) results_[i] = inner_->result(i);
  }

  // host scratch reused across searches
  std::vector<uint64_t> keys;
  std::vector<float> scores;
  std::vector<uint32_t> index, counts, bitmap, rows;

 protected:
  uint32_t topk_{0};
  uint32_t magic_{0};
  uint32_t group_num_{0};
  bool fetch_vector_{false};
  zvb200_ctx *ctx_{nullptr};
  IndexContext::Pointer inner_{};
  std::vector<IndexDocumentList> results_{1};
  std::vector<Stats> stats_vec_{1};
};

namespace {

//! The host side of a filtered search: IndexFilter is an opaque std::function<bool(key)> (index_filter.h:23-62; true =
//! drop, flat_searcher_context.h single_enqueue_filter) with no identity to key a cache by, so every filtered search
//! evaluates it once per stored key — exactly as often as the reference calls it — into the bitmap the scan reads.
//! Above 256k keys the evaluation is spread over host threads, whole bitmap words each: the reference itself calls
//! one filter object from many searching threads at once (index.cc:24-45), so the callable is thread-safe by contract.
//! 10M keys: ~50 ms on one core, ~4 ms on 16 — next to a 2.4 ms scan.  Returns the rows dropped.
size_t build_filter_bitmap(const IndexFilter &filter, const uint64_t *keys, size_t n, bool skip_holes,
                           std::vector<uint32_t> *bitmap) {
  const size_t words = (n + 31) / 32;
  bitmap->assign(words, 0u);
  auto run = [&](size_t w0, size_t w1) -> size_t {
    size_t dropped = 0;
    for (size_t w = w0; w < w1; ++w) {
      uint32_t bits = 0;
      const size_t hi = std::min(n, (w + 1) * 32);
      for (size_t i = w * 32; i < hi; ++i)
        if (!(skip_holes && keys[i] == ZVB200_INVALID_KEY) && filter(keys[i])) { bits |= 1u << (i & 31); ++dropped; }
      (*bitmap)[w] = bits;
    }
    return dropped;
  };
  const unsigned hw = std::thread::hardware_concurrency();
  const size_t nt = n < ((size_t)1 << 18) ? 1 : std::min<size_t>({(size_t)(hw ? hw : 1), (size_t)16, words / 2048 + 1});
  if (nt <= 1) return run(0, words);
  std::vector<size_t> part(nt, 0);
  std::vector<std::thread> th;
  for (size_t t = 0; t < nt; ++t)
    th.emplace_back([&, t] { part[t] = run(words * t / nt, words * (t + 1) / nt); });
  size_t dropped = 0;
  for (size_t t = 0; t < nt; ++t) { th[t].join(); dropped += part[t]; }
  return dropped;
}

//! one search_bf_by_p_keys pass: each query's listed rows (ZVB200_INVALID_INDEX = skipped) in slices of
//! ZVB200_MAX_CANDIDATES, the slices' best-k merged by (score, row) on the host — k entries per slice, not a scan
int search_rows_sliced(GpuContext *ctx, const void *query, uint32_t count, const std::vector<std::vector<uint32_t>> &rows,
                       uint32_t k, std::vector<std::vector<std::pair<float, std::pair<uint32_t, uint64_t>>>> *out) {
  size_t longest = 0;
  for (auto &r : rows) longest = std::max(longest, r.size());
  out->assign(count, {});
  if (longest == 0) return 0;
  for (size_t s0 = 0; s0 < longest; s0 += ZVB200_MAX_CANDIDATES) {
    const uint32_t m = (uint32_t)std::min<size_t>(ZVB200_MAX_CANDIDATES, longest - s0);
    ctx->rows.assign((size_t)count * m, ZVB200_INVALID_INDEX);
    for (uint32_t i = 0; i < count; ++i)
      for (size_t j = s0; j < rows[i].size() && j < s0 + m; ++j) ctx->rows[(size_t)i * m + (j - s0)] = rows[i][j];
    const int rc = zvb200_search_rows(ctx->handle(), query, count, ctx->rows.data(), m, k, ctx->threshold(), nullptr,
                                      ctx->keys.data(), ctx->scores.data(), ctx->index.data(), ctx->counts.data());
    if (rc != 0) return rc;
    for (uint32_t i = 0; i < count; ++i)
      for (uint32_t j = 0; j < ctx->counts[i]; ++j) {
        const size_t o = (size_t)i * k + j;
        (*out)[i].push_back({ctx->scores[o], {ctx->index[o], ctx->keys[o]}});
      }
  }
  if (longest > ZVB200_MAX_CANDIDATES)
    for (auto &l : *out) {
      std::sort(l.begin(), l.end(), [](const auto &a, const auto &b) {
        return a.first < b.first || (a.first == b.first && a.second.first < b.second.first);
      });
      if (l.size() > k) l.resize(k);
    }
  return 0;
}

}  // namespace

/*! GPU Flat Searcher: IndexSearcher (index_searcher.h:29-54, index_runner.h:404-640) whose load() uploads the Flat
 *  index file to a device-resident store and whose searches are zvb200_search / zvb200_search_rows calls.
 *  Mirrors FlatSearcher<32> (src/core/algorithm/flat/flat_searcher.h:33-161, flat_searcher.cc:68-245): the embedded
 *  stock FlatSearcher<32> runs the same load() checks with the same error codes and keeps serving
 *  create_provider() / get_vector(key) / fetch_vector / group-by over the same mmap'ed file.
 */
class GpuFlatSearcher : public IndexSearcher {
 public:
  GpuFlatSearcher() = default;
  ~GpuFlatSearcher() override { this->unload(); }

  int init(const ailego::Params &params) override {
    params_ = params;
    devices_ = parse_devices(params);
    stock_ = IndexFactory::CreateSearcher("FlatSearcher32");  // never overridden (flat_searcher.cc:250)
    if (!stock_) return IndexError_NoExist;
    return stock_->init(params);
  }
  int cleanup(void) override {
    this->unload();
    return stock_ ? stock_->cleanup() : 0;
  }

  //! FlatSearcher::load (flat_searcher.cc:68-159) — by the stock object itself — then the upload
  int load(IndexStorage::Pointer cntr, IndexMetric::Pointer measure) override {
    if (!cntr) return IndexError_InvalidArgument;
    if (!stock_) {
      const int r = this->init(params_);
      if (r != 0) return r;
    }
    int rc = stock_->load(cntr, measure);
    if (rc != 0) return rc;
    meta_ = stock_->meta();
    const int metric = metric_of(meta_), dtype = dtype_of(meta_);
    if (metric < 0 || dtype < 0) return IndexError_Unsupported;
    auto keys_seg = cntr->get(kSegKeys);
    auto feat_seg = cntr->get(kSegFeatures);
    if (!keys_seg || !feat_seg) return IndexError_NoExist;
    if (keys_seg->data_size() % sizeof(uint64_t) != 0) return IndexError_InvalidLength;
    const size_t n = keys_seg->data_size() / sizeof(uint64_t);
    const size_t es = meta_.element_size();
    if (n * es != feat_seg->data_size()) return IndexError_Mismatch;
    const uint64_t *keys = nullptr;
    if (keys_seg->read(0, reinterpret_cast<const void **>(&keys), keys_seg->data_size()) != keys_seg->data_size())
      return IndexError_ReadData;
    keys_.assign(keys, keys + n);

    zvb200_store *st = nullptr;
    rc = zvb200_store_create_sharded(devices_.data(), (uint32_t)devices_.size(), dtype, metric, meta_.dimension(), &st);
    if (rc != 0) return rc;
    rc = zvb200_store_reserve(st, n);
    // upload in 64 MiB reads (the reference scans in 4 MiB reads, flat_searcher_context.h:187-196)
    const bool column = meta_.major_order() == IndexMeta::MO_COLUMN;
    const size_t rows_per_read = std::max<size_t>(32, ((size_t)64 << 20) / es / 32 * 32);
    std::vector<char> rowbuf;
    for (size_t r0 = 0; rc == 0 && r0 < n; r0 += rows_per_read) {
      const size_t m = std::min(rows_per_read, n - r0);
      const void *p = nullptr;
      if (feat_seg->read(r0 * es, &p, m * es) != m * es) { rc = IndexError_ReadData; break; }
      if (column) {
        // 32-row tiles interleaved by unit (flat_builder.cc:188-225, matrix_helper.h:24-33);
        // a final tile of < 32 rows is stored row-major.  Undo it on the host: the store is row-major.
        rowbuf.resize(m * es);
        const size_t us = IndexMeta::AlignSizeof(meta_.data_type()), units = es / us;  // flat_builder.h:77-85
        const char *src = (const char *)p;
        size_t i = 0;
        for (; i + 32 <= m; i += 32)
          for (size_t u = 0; u < units; ++u)
            for (size_t j = 0; j < 32; ++j)
              std::memcpy(&rowbuf[(i + j) * es + u * us], src + i * es + (u * 32 + j) * us, us);
        if (i < m) std::memcpy(&rowbuf[i * es], src + i * es, (m - i) * es);
        p = rowbuf.data();
      }
      rc = zvb200_store_append(st, p, m, keys_.data() + r0);
    }
    if (rc == 0) rc = zvb200_store_seal(st);
    if (rc != 0) { zvb200_store_destroy(st); return rc; }
    store_ = st;
    magic_ = IndexContext::GenerateMagic();
    stats_.set_loaded_count(n);
    return 0;
  }

  int unload(void) override {
    if (store_) zvb200_store_destroy(store_);
    store_ = nullptr;
    keys_.clear();
    {
      std::lock_guard<std::mutex> g(map_mu_);
      key_row_.clear();
      key_row_built_ = false;
    }
    return stock_ ? stock_->unload() : 0;
  }

  int search_impl(const void *query, const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return this->search_impl(query, qmeta, 1, context);
  }

  //! FlatSearcher::search_impl(query, qmeta, count, ctx) (flat_searcher.cc:186-211)
  int search_impl(const void *query, const IndexQueryMeta &qmeta, uint32_t count,
                  Context::Pointer &context) const override {
    if (!query || !count || !context || !store_) return IndexError_InvalidArgument;
    if (qmeta.element_size() != meta_.element_size()) return IndexError_Mismatch;
    auto *ctx = dynamic_cast<GpuContext *>(context.get());
    if (!ctx) return IndexError_InvalidArgument;
    if (ctx->magic() != magic_) this->rebind(ctx);
    if (ctx->group_by_search()) {  // flat_searcher.cc:178-179,205-206: bookkeeping per group id, the stock object's
      log_delegated_once(&logged_group_by_, "GpuFlatSearcher", "a group-by search");
      ctx->sync_inner();
      return stock_->search_impl(query, qmeta, count, ctx->inner());
    }
    const uint32_t k = ctx->topk();
    ctx->begin(count);
    if (k == 0) return 0;
    const size_t n = keys_.size();
    const uint32_t *bitmap = nullptr;
    size_t dropped = 0;
    if (ctx->filter().is_valid()) {
      // IndexFilter: true = drop (flat_searcher_context.h single_enqueue_filter); evaluated
      // once per row on the host, exactly as often as the reference calls it
      dropped = build_filter_bitmap(ctx->filter(), keys_.data(), n, false, &ctx->bitmap);
      bitmap = ctx->bitmap.data();
    }
    const int rc = zvb200_search(ctx->handle(), query, count, k, ctx->threshold(), bitmap, ctx->keys.data(),
                                 ctx->scores.data(), ctx->index.data(), ctx->counts.data());
    if (rc != 0) return rc;
    ctx->fill_stats(count, n - dropped, dropped);
    this->emit(ctx, count, k);
    return 0;
  }

  int search_bf_impl(const void *query, const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return this->search_impl(query, qmeta, 1, context);
  }
  int search_bf_impl(const void *query, const IndexQueryMeta &qmeta, uint32_t count,
                     Context::Pointer &context) const override {
    return this->search_impl(query, qmeta, count, context);
  }

  int search_bf_by_p_keys_impl(const void *query, const std::vector<std::vector<uint64_t>> &p_keys,
                               const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return this->search_bf_by_p_keys_impl(query, p_keys, qmeta, 1, context);
  }

  //! FlatSearcher::search_bf_by_p_keys_impl (flat_searcher.cc:213-237; the loop is
  //! FlatSearcherContext::search_bf_by_p_keys_impl, the streamer's twin is flat_streamer.cc:347-389): only the listed
  //! keys are scored — filtered keys and keys the index does not hold are skipped, a key listed twice is scored twice
  int search_bf_by_p_keys_impl(const void *query, const std::vector<std::vector<uint64_t>> &p_keys,
                               const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const override {
    if (!query || !count || !context || !store_) return IndexError_InvalidArgument;
    if (p_keys.size() != count) return IndexError_InvalidArgument;
    if (qmeta.element_size() != meta_.element_size()) return IndexError_Mismatch;
    auto *ctx = dynamic_cast<GpuContext *>(context.get());
    if (!ctx) return IndexError_InvalidArgument;
    if (ctx->magic() != magic_) this->rebind(ctx);
    if (ctx->group_by_search()) {
      log_delegated_once(&logged_group_by_, "GpuFlatSearcher", "a group-by search");
      ctx->sync_inner();
      return stock_->search_bf_by_p_keys_impl(query, p_keys, qmeta, count, ctx->inner());
    }
    const uint32_t k = ctx->topk();
    ctx->begin(count);
    if (k == 0) return 0;
    this->build_key_map();
    std::vector<std::vector<uint32_t>> rows(count);
    size_t calced = 0, filtered = 0;
    for (uint32_t i = 0; i < count; ++i) {
      rows[i].reserve(p_keys[i].size());
      for (uint64_t key : p_keys[i]) {
        if (ctx->filter().is_valid() && ctx->filter()(key)) { ++filtered; continue; }
        auto it = key_row_.find(key);
        if (it == key_row_.end()) continue;
        rows[i].push_back(it->second);
        ++calced;
      }
    }
    std::vector<std::vector<std::pair<float, std::pair<uint32_t, uint64_t>>>> lists;
    // FlatSearcherContext::search_bf_by_p_keys_impl hands the metric (query, feature) (flat_searcher_context.h:1091-1093)
    zvb200_ctx_set_operand_order(ctx->handle(), 1);
    const int rc = search_rows_sliced(ctx, query, count, rows, k, &lists);
    zvb200_ctx_set_operand_order(ctx->handle(), -1);
    if (rc != 0) return rc;
    for (uint32_t i = 0; i < count; ++i) {
      IndexDocumentList *out = ctx->mutable_result(i);
      for (auto &e : lists[i]) this->push_doc(ctx, out, e.second.second, e.first, e.second.first);
    }
    ctx->fill_stats(count, calced / count, filtered / count);
    return 0;
  }

  Context::Pointer create_context(void) const override {
    if (!stock_) return Context::Pointer();
    auto *c = new (std::nothrow) GpuContext(stock_->create_context());
    if (c) c->bind(store_, magic_);
    return Context::Pointer(c);
  }
  //! FlatSearcher::create_provider (flat_searcher.cc:26-65) / get_vector (flat_searcher.h:158-160): the stock
  //! object's, over the same mmap'ed file
  IndexProvider::Pointer create_provider(void) const override { return stock_ ? stock_->create_provider() : nullptr; }
  const void *get_vector(uint64_t key) const override { return stock_ ? stock_->get_vector(key) : nullptr; }
  const Stats &stats(void) const override { return stats_; }
  const IndexMeta &meta(void) const override { return meta_; }
  const ailego::Params &params(void) const override { return params_; }

  //! copy one stored row back from the device (reference layout)
  int get_vector_by_id(uint32_t id, void *out, size_t cap) const {
    return store_ ? zvb200_store_get_vector(store_, id, out, cap) : IndexError_NoReady;
  }
  uint32_t magic(void) const { return magic_; }
  zvb200_store *store(void) const { return store_; }

 private:
  void rebind(GpuContext *ctx) const { ctx->bind(store_, magic_); }
  void build_key_map(void) const {
    std::lock_guard<std::mutex> g(map_mu_);
    if (key_row_built_) return;
    key_row_.reserve(keys_.size());
    for (size_t i = 0; i < keys_.size(); ++i) key_row_.emplace(keys_[i], (uint32_t)i);  // first row wins, as get_id()
    key_row_built_ = true;
  }
  void push_doc(GpuContext *ctx, IndexDocumentList *out, uint64_t key, float score, uint32_t row) const {
    if (ctx->fetch_vector()) out->emplace_back(key, score, row, stock_->get_vector(key));  // flat_searcher_context.h:162
    else out->emplace_back(key, score, row);
  }
  void emit(GpuContext *ctx, uint32_t count, uint32_t k) const {
    for (uint32_t i = 0; i < count; ++i) {
      IndexDocumentList *out = ctx->mutable_result(i);
      for (uint32_t j = 0; j < ctx->counts[i]; ++j) {
        const size_t o = (size_t)i * k + j;
        this->push_doc(ctx, out, ctx->keys[o], ctx->scores[o], ctx->index[o]);
      }
    }
  }

  ailego::Params params_{};
  IndexMeta meta_{};
  Stats stats_{};
  IndexSearcher::Pointer stock_{};
  std::vector<uint64_t> keys_{};
  mutable std::mutex map_mu_{};
  mutable std::unordered_map<uint64_t, uint32_t> key_row_{};
  mutable bool key_row_built_{false};
  mutable std::atomic<bool> logged_group_by_{false};
  zvb200_store *store_{nullptr};
  uint32_t magic_{0};
  std::vector<int> devices_{0};
};

INDEX_FACTORY_REGISTER_SEARCHER(GpuFlatSearcher);

// ======================================================================================
// GpuFlatStreamer — FlatStreamer's interface (flat_streamer.h:28-160): rows arrive one by one
// between searches.  Composition: persistence (storage blocks, key maps, flush / dump / provider /
// get_vector, duplicate-key refusal) stays with an embedded stock FlatStreamer<32>, created under
// its never-overridden alias "FlatStreamer32" (flat_streamer.cc:486-489); every accepted add is
// mirrored into the device store, and every search is the GPU scan.
// ======================================================================================
class GpuFlatStreamer : public IndexStreamer {
 public:
  GpuFlatStreamer() = default;
  ~GpuFlatStreamer() override { this->cleanup(); }

  int init(const IndexMeta &meta, const ailego::Params &params) override {
    stock_ = IndexFactory::CreateStreamer("FlatStreamer32");
    if (!stock_) return IndexError_NoExist;
    int rc = stock_->init(meta, params);
    if (rc != 0) return rc;
    meta_ = stock_->meta();
    meta_.set_streamer("GpuFlatStreamer", 0U, params);
    params.get(kParamDevice, &device_);
    if (metric_of(meta_) < 0 || dtype_of(meta_) < 0) return IndexError_Unsupported;
    return 0;
  }

  int cleanup(void) override {
    this->close();
    if (stock_) stock_->cleanup();
    stock_.reset();
    return 0;
  }

  //! open = the stock open (it validates / recovers the storage), then every row it holds goes to the GPU
  int open(IndexStorage::Pointer stg) override {
    if (!stock_) return IndexError_NoReady;
    int rc = stock_->open(stg);
    if (rc != 0) return rc;
    std::unique_lock<std::shared_mutex> g(mu_);
    rc = zvb200_store_create(device_, dtype_of(meta_), metric_of(meta_), meta_.dimension(), &store_);
    if (rc != 0) return rc;
    // FlatStreamerEntity::row_major_distance hands the metric (query, feature) (flat_streamer_entity.h:155-162)
    if ((rc = zvb200_store_set_operand_order(store_, 1)) != 0) return rc;
    auto provider = stock_->create_provider();
    if (provider) {
      const size_t es = meta_.element_size();
      for (auto it = provider->create_iterator(); it && it->is_valid(); it->next()) {
        pend_keys_.push_back(it->key());
        const char *v = static_cast<const char *>(it->data());
        pend_rows_.insert(pend_rows_.end(), v, v + es);
        if (pend_rows_.size() >= ((size_t)64 << 20) && (rc = flush_pending()) != 0) return rc;
      }
    }
    magic_ = IndexContext::GenerateMagic();
    return flush_pending();
  }

  int close(void) override {
    {
      std::unique_lock<std::shared_mutex> g(mu_);
      if (store_) zvb200_store_destroy(store_);
      store_ = nullptr;
      pend_rows_.clear(); pend_keys_.clear(); all_keys_.clear(); key_row_.clear();
      identity_ = true; live_ = 0; pending_.store(0);
    }
    return stock_ ? stock_->close() : 0;
  }
  int flush(uint64_t checkpoint) override { return stock_ ? stock_->flush(checkpoint) : IndexError_NoReady; }
  int dump(const IndexDumper::Pointer &dumper) override { return stock_ ? stock_->dump(dumper) : IndexError_NoReady; }

  Context::Pointer create_context(void) const override {
    if (!stock_) return Context::Pointer();
    auto inner = stock_->create_context();
    if (!inner) return Context::Pointer();  // FlatStreamer::create_context: "open storage first" (flat_streamer.cc:197-205)
    auto *c = new (std::nothrow) GpuContext(std::move(inner));
    if (c) c->bind(store_, magic_);
    return Context::Pointer(c);
  }
  IndexProvider::Pointer create_provider(void) const override { return stock_ ? stock_->create_provider() : nullptr; }

  //! FlatStreamer::add_impl (flat_streamer.cc) -> FlatStreamerEntity::add (flat_streamer_entity.cc:149-210)
  int add_impl(uint64_t pkey, const void *vec, const IndexQueryMeta &qmeta, Context::Pointer &context) override {
    auto *ctx = dynamic_cast<GpuContext *>(context.get());
    if (!ctx || !stock_ || !store_) return IndexError_InvalidArgument;
    const int rc = stock_->add_impl(pkey, vec, qmeta, ctx->inner());
    if (rc != 0) return rc;  // duplicate key, size mismatch, index full: nothing to mirror
    std::unique_lock<std::shared_mutex> g(mu_);
    const char *v = static_cast<const char *>(vec);
    pend_rows_.insert(pend_rows_.end(), v, v + meta_.element_size());
    pend_keys_.push_back(pkey);
    pending_.fetch_add(1);
    return 0;
  }

  //! FlatStreamerEntity::add_vector_with_id (flat_streamer_entity.cc:900-991): rows up to `id` become
  //! kInvalidKey holes, row `id` gets key = id; an existing id is overwritten in place
  int add_with_id_impl(uint32_t id, const void *vec, const IndexQueryMeta &qmeta, Context::Pointer &context) override {
    auto *ctx = dynamic_cast<GpuContext *>(context.get());
    if (!ctx || !stock_ || !store_) return IndexError_InvalidArgument;
    int rc = stock_->add_with_id_impl(id, vec, qmeta, ctx->inner());
    if (rc != 0) return rc;
    std::unique_lock<std::shared_mutex> g(mu_);
    const size_t es = meta_.element_size();
    const uint64_t on_device = zvb200_store_count(store_);
    const uint64_t total = on_device + pend_keys_.size();
    const char *v = static_cast<const char *>(vec);
    if (id >= total) {
      for (uint64_t r = total; r < id; ++r) {
        pend_rows_.insert(pend_rows_.end(), es, 0);
        pend_keys_.push_back(ZVB200_INVALID_KEY);
      }
      pend_rows_.insert(pend_rows_.end(), v, v + es);
      pend_keys_.push_back(id);
      pending_.fetch_add(1);
    } else if (id >= on_device) {
      std::memcpy(&pend_rows_[(id - on_device) * es], v, es);
      pend_keys_[id - on_device] = id;
    } else {
      rc = zvb200_store_write(store_, id, vec, id);
      if (rc == 0) {
        if (all_keys_[id] == ZVB200_INVALID_KEY) ++live_;
        this->note_key(id, id);
        all_keys_[id] = id;
      }
    }
    return rc;
  }

  int search_impl(const void *query, const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return this->search_bf_impl(query, qmeta, 1, context);
  }
  int search_impl(const void *query, const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const override {
    return this->search_bf_impl(query, qmeta, count, context);
  }
  int search_bf_impl(const void *query, const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return this->search_bf_impl(query, qmeta, 1, context);
  }

  //! FlatStreamer::search_bf_impl -> FlatStreamerEntity::search / search_block (flat_streamer_entity.cc:212-316)
  int search_bf_impl(const void *query, const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const override {
    GpuContext *ctx = nullptr;
    int rc = this->enter(query, qmeta, count, context, &ctx);
    if (rc != 0) return rc;
    if (ctx->group_by_search()) {  // flat_streamer.cc:321-323
      log_delegated_once(&logged_group_by_, "GpuFlatStreamer", "a group-by search");
      ctx->sync_inner();
      return stock_->search_bf_impl(query, qmeta, count, ctx->inner());
    }
    const uint32_t k = ctx->topk();
    ctx->begin(count);
    if (k == 0) return 0;
    // The rows this search sees = the rows published when it starts.  The filter is evaluated over THAT many keys
    // under the shared lock (adds wait, other searches do not), and the GPU call runs outside every plugin lock
    // with the scan limited to the snapshot's rows: rows another thread publishes meanwhile are simply not seen.
    size_t n = 0, dropped = 0;
    const uint32_t *bitmap = nullptr;
    {
      std::shared_lock<std::shared_mutex> g(mu_);
      n = all_keys_.size();
      if (n == 0) return 0;
      dropped = n - live_;
      if (ctx->filter().is_valid()) {
        dropped += build_filter_bitmap(ctx->filter(), all_keys_.data(), n, true, &ctx->bitmap);
        bitmap = ctx->bitmap.data();
      }
    }
    zvb200_ctx_set_row_limit(ctx->handle(), n);
    rc = zvb200_search(ctx->handle(), query, count, k, ctx->threshold(), bitmap, ctx->keys.data(),
                       ctx->scores.data(), ctx->index.data(), ctx->counts.data());
    zvb200_ctx_set_row_limit(ctx->handle(), 0);
    if (rc != 0) return rc;
    // flat_streamer_entity.cc:246 (no filter, no deletion: every row of a block counts) / :306-314
    ctx->fill_stats(count, bitmap || dropped ? n - dropped : n, bitmap || dropped ? dropped : 0);
    for (uint32_t i = 0; i < count; ++i) {
      IndexDocumentList *out = ctx->mutable_result(i);
      for (uint32_t j = 0; j < ctx->counts[i]; ++j) {
        const size_t o = (size_t)i * k + j;
        this->push_doc(ctx, out, ctx->keys[o], ctx->scores[o]);
      }
    }
    return 0;
  }

  int search_bf_by_p_keys_impl(const void *query, const std::vector<std::vector<uint64_t>> &p_keys,
                               const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return this->search_bf_by_p_keys_impl(query, p_keys, qmeta, 1, context);
  }

  //! FlatStreamer::search_bf_by_p_keys_impl (flat_streamer.cc:347-389): what core_interface::Index::_dense_search
  //! calls whenever bf_pks is set (index.cc:603-610) — the second stage of the refiner flow (index.cc:408-445)
  int search_bf_by_p_keys_impl(const void *query, const std::vector<std::vector<uint64_t>> &p_keys,
                               const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const override {
    GpuContext *ctx = nullptr;
    int rc = this->enter(query, qmeta, count, context, &ctx);
    if (rc != 0) return rc;
    if (p_keys.size() < count) return IndexError_InvalidArgument;
    if (ctx->group_by_search()) {  // flat_streamer.cc:365-367
      log_delegated_once(&logged_group_by_, "GpuFlatStreamer", "a group-by search");
      ctx->sync_inner();
      return stock_->search_bf_by_p_keys_impl(query, p_keys, qmeta, count, ctx->inner());
    }
    const uint32_t k = ctx->topk();
    ctx->begin(count);
    if (k == 0) return 0;
    std::vector<std::vector<uint32_t>> rows(count);
    size_t calced = 0, filtered = 0;
    {
      std::shared_lock<std::shared_mutex> g(mu_);
      const size_t n = all_keys_.size();
      for (uint32_t i = 0; i < count; ++i) {
        rows[i].reserve(p_keys[i].size());
        for (uint64_t key : p_keys[i]) {
          if (ctx->filter().is_valid() && ctx->filter()(key)) { ++filtered; continue; }
          uint32_t row = ZVB200_INVALID_INDEX;
          if (identity_) {
            if (key < n && all_keys_[key] == key) row = (uint32_t)key;
          } else {
            auto it = key_row_.find(key);
            if (it != key_row_.end()) row = it->second;
          }
          if (row == ZVB200_INVALID_INDEX) continue;  // get_vector_by_key fails: skipped (flat_streamer.cc:379)
          rows[i].push_back(row);
          ++calced;
        }
      }
    }
    std::vector<std::vector<std::pair<float, std::pair<uint32_t, uint64_t>>>> lists;
    rc = search_rows_sliced(ctx, query, count, rows, k, &lists);
    if (rc != 0) return rc;
    for (uint32_t i = 0; i < count; ++i) {
      IndexDocumentList *out = ctx->mutable_result(i);
      for (auto &e : lists[i]) this->push_doc(ctx, out, e.second.second, e.first);
    }
    ctx->fill_stats(count, calced / count, filtered / count);
    return 0;
  }

  const Stats &stats(void) const override { return stock_->stats(); }
  const IndexMeta &meta(void) const override { return meta_; }
  const void *get_vector(uint64_t key) const override { return stock_ ? stock_->get_vector(key) : nullptr; }
  int get_vector(const uint64_t key, IndexStorage::MemoryBlock &block) const override {
    return stock_ ? stock_->get_vector(key, block) : IndexError_NoReady;
  }
  const void *get_vector_by_id(uint32_t id) const override { return stock_ ? stock_->get_vector_by_id(id) : nullptr; }
  int get_vector_by_id(const uint32_t id, IndexStorage::MemoryBlock &block) const override {
    return stock_ ? stock_->get_vector_by_id(id, block) : IndexError_NoReady;
  }
  uint32_t magic(void) const { return magic_; }
  zvb200_store *store(void) const { return store_; }

 private:
  //! argument checks, context rebinding and publication of pending rows shared by every search entry
  int enter(const void *query, const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context,
            GpuContext **out) const {
    if (!query || !count || !context || !store_) return IndexError_InvalidArgument;
    if (qmeta.element_size() != meta_.element_size()) return IndexError_Mismatch;
    auto *ctx = dynamic_cast<GpuContext *>(context.get());
    if (!ctx) return IndexError_InvalidArgument;
    if (ctx->magic() != magic_) ctx->bind(store_, magic_);
    if (pending_.load() != 0) {
      std::unique_lock<std::shared_mutex> g(mu_);
      const int rc = const_cast<GpuFlatStreamer *>(this)->flush_pending();
      if (rc != 0) return rc;
    }
    *out = ctx;
    return 0;
  }
  void push_doc(GpuContext *ctx, IndexDocumentList *out, uint64_t key, float score) const {
    if (ctx->fetch_vector()) {  // FlatStreamerContext::topk_to_result (flat_streamer_context.h:119-124)
      IndexStorage::MemoryBlock block;
      stock_->get_vector(key, block);
      out->emplace_back(key, score, key, block);
    } else {
      out->emplace_back(key, score, key);  // index() = key (flat_streamer_context.h:125)
    }
  }
  void note_key(uint64_t row, uint64_t key) {
    if (key == ZVB200_INVALID_KEY) return;
    if (identity_ && key != row) {  // first key that is not its row: from now on a real map
      identity_ = false;
      key_row_.reserve(all_keys_.size() + 1024);
      for (size_t i = 0; i < all_keys_.size(); ++i)
        if (all_keys_[i] != ZVB200_INVALID_KEY) key_row_.emplace(all_keys_[i], (uint32_t)i);
    }
    if (!identity_) key_row_[key] = (uint32_t)row;
  }
  //! pending host rows -> device store; caller holds mu_ exclusively
  int flush_pending(void) {
    if (!pend_keys_.empty()) {
      const int rc = zvb200_store_append(store_, pend_rows_.data(), pend_keys_.size(), pend_keys_.data());
      if (rc != 0) return rc;
      for (uint64_t key : pend_keys_) {
        this->note_key(all_keys_.size(), key);
        all_keys_.push_back(key);
        if (key != ZVB200_INVALID_KEY) ++live_;
      }
      pend_rows_.clear(); pend_keys_.clear();
      dirty_ = true;
    }
    pending_.store(0);
    if (dirty_) {
      const int rc = zvb200_store_seal(store_);
      if (rc != 0) return rc;
      dirty_ = false;
    }
    return 0;
  }

  IndexStreamer::Pointer stock_{};
  IndexMeta meta_{};
  mutable std::shared_mutex mu_{};
  zvb200_store *store_{nullptr};
  std::vector<char> pend_rows_{};
  std::vector<uint64_t> pend_keys_{}, all_keys_{};
  std::unordered_map<uint64_t, uint32_t> key_row_{};  // only once a key differs from its row (add_impl with own keys)
  bool identity_{true};
  size_t live_{0};
  std::atomic<size_t> pending_{0};
  mutable std::atomic<bool> logged_group_by_{false};
  bool dirty_{true};
  uint32_t magic_{0};
  int device_{0};
};

INDEX_FACTORY_REGISTER_STREAMER(GpuFlatStreamer);

// ======================================================================================
// GpuIVFSearcher — IVFSearcher's interface (ivf_searcher.h:28-118).  Composition: a stock
// IVFSearcher loads the index (centroid index, reformer wiring, group-by, and the index layouts the
// device store cannot hold); the posting lists it holds are exported ONCE through the reference's own
// IVFEntity reader into a device store grouped by list, the centroids of its nested Flat index
// become the coarse store, search_impl is zvb200_ivf_search and search_bf_impl (every list, and what
// search_impl does for an index at or below brute_force_threshold, ivf_searcher.cc:185-187) is zvb200_search
// over the same store.
// ======================================================================================
class GpuIVFSearcher;

class GpuIVFContext : public GpuContext {
 public:
  GpuIVFContext(const GpuIVFSearcher *owner, IndexContext::Pointer inner) : GpuContext(std::move(inner)) { this->reset(owner); }
  int update(const ailego::Params &params) override;
  void reset(const GpuIVFSearcher *owner);
  //! the reformer inside an IVFEntity keeps a scratch buffer: one clone per context, owned by the context
  //! (IVFSearcher::update_context does the same for its own contexts)
  IVFEntity *entity() { return entity_.get(); }
  uint32_t nprobe{1};
  uint64_t max_scan{0};

 private:
  const GpuIVFSearcher *owner_{nullptr};
  IVFEntity::Pointer entity_{};
};

class GpuIVFSearcher : public IndexSearcher {
 public:
  GpuIVFSearcher() = default;
  ~GpuIVFSearcher() override { this->unload(); }

  int init(const ailego::Params &params) override {
    params_ = params;
    params.get(kParamDevice, &device_);
    params.get(PARAM_IVF_SEARCHER_BRUTE_FORCE_THRESHOLD, &bf_threshold_);
    return stock_.init(params);
  }
  int cleanup(void) override {
    this->unload();
    return stock_.cleanup();
  }

  int load(IndexStorage::Pointer cntr, IndexMetric::Pointer measure) override {
    int rc = stock_.load(cntr, measure);
    if (rc != 0) return rc;
    meta_ = stock_.meta();
    magic_ = IndexContext::GenerateMagic();
    rc = entity_.load(cntr);
    if (rc != 0) return rc;
    // The posting lists have a meta of their own (IVFEntity::load_header, ivf_entity.cc:462-476): for an index built
    // with proxima.ivf.builder.quantizer_class = Int8QuantizerConverter they are int8 rows under the searcher's fp32
    // meta, and their scores carry the quantiser's global or per-list scale (IVFEntity::load, ivf_entity.cc:541-564)
    // and a per-query one (IVFReformerWrapper::normalize :325-353).  int4 lists (Int4QuantizerConverter: two signed
    // nibbles per byte, low nibble first, ivf_entity.cc:420-433) go to the device as int8 rows: the reference's int4
    // kernels sum exact integers over the sign-extended nibbles (euclidean_distance_matrix_int4.cc:35-37,
    // inner_product_matrix_int4.cc:49-51), so the int8 kernels over the unpacked values return the same numbers.
    lmeta_ = entity_.meta();
    const std::string &rf = lmeta_.reformer_name();
    int8_lists_ = rf == kInt8ReformerName && lmeta_.data_type() == IndexMeta::DataType::DT_INT8;
    int4_lists_ = rf == kInt4ReformerName && lmeta_.data_type() == IndexMeta::DataType::DT_INT4 &&
                  lmeta_.dimension() % 2 == 0 && lmeta_.element_size() == lmeta_.dimension() / 2;
    const bool plain = rf != kInt8ReformerName && rf != kInt4ReformerName && !cntr->get(IVF_INT8_QUANTIZED_PARAMS_SEG_ID) &&
                       !cntr->get(IVF_INT4_QUANTIZED_PARAMS_SEG_ID) && dtype_of(lmeta_) >= 0;
    accelerated_ = (plain || int8_lists_ || int4_lists_) && metric_of(lmeta_) >= 0 && entity_.vector_count() > 0;
    if (!accelerated_) {
      why_not_ = "an index with posting lists of an unsupported type";
      return 0;
    }
    // centroids = the nested Flat index in segment "ivf.centroid" (ivf_searcher.cc:62-86); row cid = centroid cid
    auto seg = cntr->get(IVF_CENTROID_SEG_ID, 0);
    if (!seg) return IndexError_InvalidFormat;
    IndexStorage::Pointer nested = std::make_shared<IndexSegmentStorage>(seg);
    if ((rc = nested->open(std::string(), false)) != 0) return rc;
    IndexMeta cmeta;
    if ((rc = IndexHelper::DeserializeFromStorage(nested.get(), &cmeta)) != 0) return rc;
    auto fseg = nested->get(kSegFeatures), kseg = nested->get(kSegKeys);
    const size_t nl = entity_.inverted_list_count(), n = entity_.vector_count(), es = lmeta_.element_size();
    const size_t des = int4_lists_ ? 2 * es : es;  // bytes of a row on the device
    const size_t ces = cmeta.element_size();
    if (!fseg || !kseg || metric_of(cmeta) < 0 || dtype_of(cmeta) < 0 || kseg->data_size() != nl * sizeof(uint64_t) ||
        fseg->data_size() != nl * ces) {
      accelerated_ = false;  // a centroid index of another shape (e.g. HNSW over centroids)
      why_not_ = "an index whose centroid index is not a Flat index of a supported type";
      return 0;
    }
    // the queries of the coarse search go through the centroid index's own reformer (IVFCentroidIndex::load / search,
    // ivf_centroid_index.cc:538-563,273-295): none for L2 lists, MipsReformer for quantised InnerProduct lists
    creformer_.reset();
    c_elem_ = ces;
    if (!cmeta.reformer_name().empty()) {
      const std::string &crf = cmeta.reformer_name();
      if (crf == kInt8ReformerName || crf == kInt4ReformerName) {
        accelerated_ = false;  // quantised centroids (Int8QuantizerReformer4IP and the like are private to the reference)
        why_not_ = "an index with quantised centroids";
        return 0;
      }
      creformer_ = IndexFactory::CreateReformer(crf);
      if (!creformer_) return IndexError_NoExist;
      if ((rc = creformer_->init(cmeta.reformer_params())) != 0) return rc;
    }
    const void *fp = nullptr;
    const uint64_t *kp = nullptr;
    if (fseg->read(0, &fp, fseg->data_size()) != fseg->data_size()) return IndexError_ReadData;
    if (kseg->read(0, reinterpret_cast<const void **>(&kp), kseg->data_size()) != kseg->data_size()) return IndexError_ReadData;
    std::vector<char> cent(nl * ces);
    const bool col = cmeta.major_order() == IndexMeta::MO_COLUMN;
    const size_t us = IndexMeta::AlignSizeof(cmeta.data_type()), units = ces / us;
    for (size_t i = 0; i < nl; ++i) {
      if (kp[i] >= nl) return IndexError_InvalidFormat;
      char *dst = &cent[(size_t)kp[i] * ces];
      const size_t blk = i / 32 * 32;
      if (col && blk + 32 <= nl) {
        const char *b = static_cast<const char *>(fp) + blk * ces;
        for (size_t u = 0; u < units; ++u) std::memcpy(dst + u * us, b + (u * 32 + (i - blk)) * us, us);
      } else {
        std::memcpy(dst, static_cast<const char *>(fp) + i * ces, ces);
      }
    }
    // posting lists in list order through IVFEntity (ivf_entity.h:89-170): rows + keys + offsets
    zvb200_store *st = nullptr;
    if ((rc = zvb200_store_create(device_, int4_lists_ ? ZVB200_DT_INT8 : dtype_of(lmeta_), metric_of(lmeta_),
                                  lmeta_.dimension(), &st)) != 0)
      return rc;
    // IVFDistanceCalculator::row_major_distance: (query, feature) (ivf_distance_calculator.h:174-183)
    rc = zvb200_store_set_operand_order(st, 1);
    if (rc == 0) rc = zvb200_store_reserve(st, n);
    std::vector<uint64_t> offs(nl + 1, 0);
    std::vector<char> rows;
    std::vector<uint64_t> keys;
    keys_.clear(); keys_.reserve(n);
    size_t pos = 0;
    auto flush = [&]() -> int {
      if (keys.empty()) return 0;
      const int r = zvb200_store_append(st, rows.data(), keys.size(), keys.data());
      rows.clear(); keys.clear();
      return r;
    };
    for (size_t l = 0; rc == 0 && l < nl; ++l) {
      auto lm = entity_.inverted_list_meta(l);
      if (!lm) { rc = IndexError_ReadData; break; }
      offs[l] = pos;
      auto lk = entity_.get_keys(lm->id_offset, lm->vector_count);
      if (!lk && lm->vector_count) { rc = IndexError_ReadData; break; }
      for (size_t j = 0; j < lm->vector_count; ++j, ++pos) {
        const void *v = entity_.get_vector(lm->id_offset + j);
        if (!v) { rc = IndexError_ReadData; break; }
        if (int4_lists_) {
          rows.resize(rows.size() + des);
          unpack_int4(static_cast<const uint8_t *>(v), es, reinterpret_cast<int8_t *>(&rows[rows.size() - des]));
        } else {
          rows.insert(rows.end(), static_cast<const char *>(v), static_cast<const char *>(v) + es);
        }
        keys.push_back(lk[j]);
        keys_.push_back(lk[j]);
      }
      if (rows.size() >= ((size_t)64 << 20)) rc = flush();
    }
    if (rc == 0) rc = flush();
    offs[nl] = pos;
    if (rc == 0 && pos != n) rc = IndexError_InvalidFormat;
    if (rc == 0) rc = zvb200_store_seal(st);
    if (rc == 0)
      rc = zvb200_ivf_attach_typed(st, cent.data(), (uint32_t)nl, offs.data(), dtype_of(cmeta), metric_of(cmeta),
                                   cmeta.dimension());
    if (rc == 0 && (int8_lists_ || int4_lists_)) {
      // distances[k] * norm_val (ivf_entity.cc:601,650): one value for the index, or one per posting list
      std::vector<float> nv(nl);
      bool same = true;
      for (size_t l = 0; l < nl; ++l) {
        nv[l] = entity_.inverted_list_normalize_value(l);
        same = same && nv[l] == nv[0];
      }
      rc = zvb200_store_set_score_scale(st, same ? nv[0] : 1.0f, same ? nullptr : nv.data());
    }
    if (rc != 0) { zvb200_store_destroy(st); return rc; }
    store_ = st;
    return 0;
  }

  int unload(void) override {
    if (store_) zvb200_store_destroy(store_);
    store_ = nullptr;
    keys_.clear();
    return 0;
  }

  int search_bf_impl(const void *query, const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return this->search_bf_impl(query, qmeta, 1, context);
  }
  //! IVFSearcher::search_bf_impl (ivf_searcher.cc:125-176): every posting list = one flat scan of the store
  int search_bf_impl(const void *query, const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const override {
    return this->run(true, query, qmeta, count, context);
  }
  int search_impl(const void *query, const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return this->search_impl(query, qmeta, 1, context);
  }
  //! IVFSearcher::search_impl (ivf_searcher.cc:183-250)
  int search_impl(const void *query, const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const override {
    // an index at or below the brute-force threshold is scanned whole (ivf_searcher.cc:185-187)
    return this->run(entity_.vector_count() <= bf_threshold_, query, qmeta, count, context);
  }

  Context::Pointer create_context(void) const override {
    auto inner = stock_.create_context();
    if (!inner) return Context::Pointer();
    auto *c = new (std::nothrow) GpuIVFContext(this, std::move(inner));
    return Context::Pointer(c);
  }
  IndexProvider::Pointer create_provider(void) const override { return stock_.create_provider(); }
  const Stats &stats(void) const override { return stock_.stats(); }
  const IndexMeta &meta(void) const override { return meta_; }
  const ailego::Params &params(void) const override { return params_; }
  uint32_t magic(void) const { return magic_; }
  zvb200_store *store(void) const { return store_; }
  size_t list_count(void) const { return entity_.inverted_list_count(); }
  size_t vector_count(void) const { return entity_.vector_count(); }
  uint32_t bf_threshold(void) const { return bf_threshold_; }
  IVFEntity::Pointer clone_entity(void) const { return entity_.clone(); }
  //! what IVFStreamer::get_vector_by_id reads (ivf_streamer.h:79-82)
  int vector_by_key(uint64_t key, IndexStorage::MemoryBlock &block) const { return entity_.get_vector_by_key(key, block); }
  //! the device store AND the stock searcher's index (IVFStreamer::unload, ivf_streamer.cc:119-129)
  int unload_all(void) {
    this->unload();
    return stock_.unload();
  }

 private:
  int run(bool brute_force, const void *query, const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const {
    auto *ctx = dynamic_cast<GpuIVFContext *>(context.get());
    if (!ctx || ctx->topk() == 0) return IndexError_InvalidArgument;
    if (!accelerated_ || !store_) {
      log_delegated_once(&logged_layout_, "GpuIVFSearcher", why_not_);
      return this->delegate(brute_force, query, qmeta, count, context);
    }
    if (ctx->group_by_search() || ctx->group_by().is_valid()) {
      log_delegated_once(&logged_group_by_, "GpuIVFSearcher", "a group-by search");
      return this->delegate(brute_force, query, qmeta, count, context);
    }
    if (!query || !count || qmeta.element_size() != meta_.element_size()) return IndexError_InvalidArgument;
    if (ctx->magic() != magic_) ctx->reset(this);
    // the coarse search sees the caller's queries through the centroid index's reformer (ivf_centroid_index.cc:273-295)
    const void *cq = query;
    IndexQueryMeta c_qmeta = qmeta;
    std::string cbuf;
    int rc = 0;
    if (creformer_ && !brute_force) {
      if ((rc = creformer_->transform(query, qmeta, count, &cbuf, &c_qmeta)) != 0) return rc;
      cq = cbuf.data();
    }
    if (!brute_force && c_qmeta.element_size() != c_elem_) return IndexError_Mismatch;
    // queries into the posting lists' format (entity->transform, ivf_searcher.cc:211-214)
    const void *tq = query;
    IndexQueryMeta iv_qmeta;
    rc = ctx->entity()->transform(query, qmeta, count, &tq, &iv_qmeta);
    if (rc != 0) return rc;
    if (iv_qmeta.element_size() != lmeta_.element_size()) return IndexError_Mismatch;
    std::vector<int8_t> unpacked;
    if (int4_lists_) {  // the queries as int8 too
      const size_t es = lmeta_.element_size();
      unpacked.resize((size_t)count * 2 * es);
      for (uint32_t i = 0; i < count; ++i) unpack_int4(static_cast<const uint8_t *>(tq) + i * es, es, &unpacked[(size_t)i * 2 * es]);
      tq = unpacked.data();
    }
    const uint32_t k = ctx->topk();
    ctx->begin(count);
    const size_t n = keys_.size();
    const uint32_t *bitmap = nullptr;
    size_t dropped = 0;
    if (ctx->filter().is_valid()) {
      dropped = build_filter_bitmap(ctx->filter(), keys_.data(), n, false, &ctx->bitmap);
      bitmap = ctx->bitmap.data();
    }
    if (brute_force)
      rc = zvb200_search(ctx->handle(), tq, count, k, ctx->threshold(), bitmap, ctx->keys.data(), ctx->scores.data(),
                         ctx->index.data(), ctx->counts.data());
    else if (cq == tq)
      rc = zvb200_ivf_search(ctx->handle(), tq, count, k, ctx->nprobe, ctx->max_scan, ctx->threshold(), bitmap,
                             ctx->keys.data(), ctx->scores.data(), ctx->index.data(), ctx->counts.data());
    else
      rc = zvb200_ivf_search_split(ctx->handle(), cq, tq, count, k, ctx->nprobe, ctx->max_scan, ctx->threshold(), bitmap,
                                   ctx->keys.data(), ctx->scores.data(), ctx->index.data(), ctx->counts.data());
    if (rc != 0) return rc;
    if (brute_force) ctx->fill_stats(count, n - dropped, dropped);
    for (uint32_t i = 0; i < count; ++i) {
      IndexDocumentList *out = ctx->mutable_result(i);
      if (int8_lists_ || int4_lists_) {
        // entity->normalize(q, &heap) (ivf_searcher.cc:240): one float factor per query — the quantiser's 1/scale, or
        // 1/(this query's own scale) for InnerProduct (ivf_entity.cc:325-353); 1.0f * factor is the factor itself
        IndexDocumentHeap probe(1);
        probe.emplace(0, 1.0f);
        ctx->entity()->normalize(i, &probe);
        const float factor = probe.begin()->score();
        for (uint32_t j = 0; j < ctx->counts[i]; ++j) ctx->scores[(size_t)i * k + j] *= factor;
      }
      for (uint32_t j = 0; j < ctx->counts[i]; ++j) {
        const size_t o = (size_t)i * k + j;
        if (ctx->fetch_vector()) {  // IVFSearcherContext::topk_to_result (ivf_searcher_context.h:190-197)
          IndexStorage::MemoryBlock block;
          entity_.get_vector_by_key(ctx->keys[o], block);
          out->emplace_back(ctx->keys[o], ctx->scores[o], ctx->keys[o], block);
        } else {
          out->emplace_back(ctx->keys[o], ctx->scores[o]);  // key and score only (ivf_searcher_context.h:196)
        }
      }
    }
    return 0;
  }
  //! the stock searcher with the caller's settings, results copied back
  int delegate(bool bf, const void *query, const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const {
    auto *ctx = dynamic_cast<GpuIVFContext *>(context.get());
    if (!ctx) return IndexError_InvalidArgument;
    ctx->sync_inner();
    auto &in = ctx->inner();
    const int rc = bf ? stock_.search_bf_impl(query, qmeta, count, in) : stock_.search_impl(query, qmeta, count, in);
    if (rc != 0) return rc;
    ctx->take_inner_results(count);
    return 0;
  }

  IVFSearcher stock_{};
  IVFEntity entity_{};
  ailego::Params params_{};
  IndexMeta meta_{};   // the searcher's meta: what the caller's queries look like
  IndexMeta lmeta_{};  // the posting lists' meta (IVFEntity::meta)
  IndexReformer::Pointer creformer_{};  // the centroid index's reformer, or null
  size_t c_elem_{0};
  bool int8_lists_{false}, int4_lists_{false};
  //! two signed nibbles per byte, low nibble = the even element (IVFReformerWrapper::transform, ivf_entity.cc:427-431)
  static void unpack_int4(const uint8_t *in, size_t bytes, int8_t *out) {
    for (size_t i = 0; i < bytes; ++i) {
      out[2 * i] = static_cast<int8_t>(static_cast<int8_t>(in[i] << 4) >> 4);
      out[2 * i + 1] = static_cast<int8_t>(static_cast<int8_t>(in[i]) >> 4);
    }
  }
  std::vector<uint64_t> keys_{};
  zvb200_store *store_{nullptr};
  const char *why_not_{"an index that was not loaded"};
  mutable std::atomic<bool> logged_group_by_{false}, logged_layout_{false};
  uint32_t bf_threshold_{1000u};
  uint32_t magic_{0};
  int device_{0};
  bool accelerated_{false};
};

//! IVFSearcherContext::update (ivf_searcher_context.h:61-79)
int GpuIVFContext::update(const ailego::Params &params) {
  if (!owner_) return IndexError_NoReady;
  uint32_t bf = owner_->bf_threshold();
  float ratio = 0.1f;  // kDefaultScanRatio (ivf_searcher_context.h:205)
  owner_->params().get(PARAM_IVF_SEARCHER_SCAN_RATIO, &ratio);
  owner_->params().get(PARAM_IVF_SEARCHER_BRUTE_FORCE_THRESHOLD, &bf);
  params.get(PARAM_IVF_SEARCHER_BRUTE_FORCE_THRESHOLD, &bf);
  params.get(PARAM_IVF_SEARCHER_SCAN_RATIO, &ratio);
  if (ratio <= 0.0f) return IndexError_InvalidArgument;
  nprobe = std::max(static_cast<uint32_t>(std::round(owner_->list_count() * ratio)), 1u);
  max_scan = std::max(bf, static_cast<uint32_t>(std::ceil(owner_->vector_count() * ratio)));
  return inner_ ? inner_->update(params) : 0;
}

void GpuIVFContext::reset(const GpuIVFSearcher *owner) {
  owner_ = owner;
  this->bind(owner->store(), owner->magic());
  entity_ = owner->clone_entity();
  this->update(owner->params());
}

INDEX_FACTORY_REGISTER_SEARCHER(GpuIVFSearcher);

// ======================================================================================
// GpuIVFStreamer — IVFStreamer's interface (ivf_streamer.h:27-120).  The reference's IVFStreamer
// is a read-only view of a built IVF index behind the streamer interface (open() loads what
// IVFSearcher::load loads, flush() is a no-op, there is no add_impl: ivf_streamer.cc:52-117), which
// is what core_interface::IVFIndex creates by name (indexes/ivf_index.cc:38-39).  Same composition
// here: the GPU searcher above does the work, this class maps the streamer life cycle onto it.
// ======================================================================================
class GpuIVFStreamer : public IndexStreamer {
 public:
  GpuIVFStreamer() = default;
  ~GpuIVFStreamer() override { this->unload(); }

  int init(const IndexMeta &meta, const ailego::Params &params) override {
    meta_ = meta;
    state_ = 1;
    return impl_.init(params);
  }
  int cleanup(void) override {
    this->unload();
    state_ = 0;
    return impl_.cleanup();
  }
  //! IVFStreamer::open (ivf_streamer.cc:52-117): meta comes from the storage, then centroids + lists
  int open(IndexStorage::Pointer storage) override {
    if (!storage) return IndexError_InvalidArgument;
    if (state_ != 1) return IndexError_Runtime;
    const int rc = impl_.load(storage, IndexMetric::Pointer());
    if (rc != 0) return rc;
    meta_ = impl_.meta();
    state_ = 2;
    return 0;
  }
  int flush(uint64_t /*check_point*/) override { return 0; }
  int close(void) override { return this->unload(); }
  int unload(void) override {
    if (state_ == 2) state_ = 1;
    return impl_.unload_all();
  }

  int search_bf_impl(const void *query, const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return impl_.search_bf_impl(query, qmeta, 1, context);
  }
  int search_bf_impl(const void *query, const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const override {
    return impl_.search_bf_impl(query, qmeta, count, context);
  }
  int search_impl(const void *query, const IndexQueryMeta &qmeta, Context::Pointer &context) const override {
    return impl_.search_impl(query, qmeta, 1, context);
  }
  int search_impl(const void *query, const IndexQueryMeta &qmeta, uint32_t count, Context::Pointer &context) const override {
    return impl_.search_impl(query, qmeta, count, context);
  }
  Context::Pointer create_context(void) const override {
    return state_ == 2 ? impl_.create_context() : Context::Pointer();
  }
  IndexProvider::Pointer create_provider(void) const override {
    return state_ == 2 ? impl_.create_provider() : IndexProvider::Pointer();
  }
  const Stats &stats(void) const override { return impl_.stats(); }
  const IndexMeta &meta(void) const override { return meta_; }
  int get_vector_by_id(const uint32_t id, IndexStorage::MemoryBlock &block) const override {
    return impl_.vector_by_key(id, block);
  }

 private:
  GpuIVFSearcher impl_{};
  IndexMeta meta_{};
  int state_{0};
};

INDEX_FACTORY_REGISTER_STREAMER(GpuIVFStreamer);

}  // namespace core
}  // namespace zvec

// Route the stock name to the GPU class: Factory::set() is map_[key] = func, so a later
// registration replaces the earlier one (src/include/zvec/ailego/pattern/factory.h:119-122).
// After this call core_interface / IVFCentroidIndex / index files whose meta says
// searcher = "FlatSearcher" reach the GPU path unchanged.
extern "C" int zvb200_install_overrides(void) {
  static zvec::ailego::Factory<zvec::core::IndexSearcher>::Register<zvec::core::GpuFlatSearcher> r("FlatSearcher");
  static zvec::ailego::Factory<zvec::core::IndexStreamer>::Register<zvec::core::GpuFlatStreamer> rs("FlatStreamer");
  static zvec::ailego::Factory<zvec::core::IndexSearcher>::Register<zvec::core::GpuIVFSearcher> ri("IVFSearcher");
  static zvec::ailego::Factory<zvec::core::IndexStreamer>::Register<zvec::core::GpuIVFStreamer> ris("IVFStreamer");
  (void)r; (void)rs; (void)ri; (void)ris;
  return 0;
}
