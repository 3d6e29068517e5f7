// TEST INFRASTRUCTURE ONLY — C ABI over the UNMODIFIED reference classes
// (compiled from /root/reference/src by oracle/build_ref.py into oracle/_ref/).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load it.  It mirrors what
// core_interface::Index::_dense_search does (src/core/interface/index.cc:581-652):
// reformer->transform(query) -> searcher->search_impl(query, qmeta, count, ctx)
// -> metric->normalize(score); and how tests build indexes
// (tests/core/algorithm/flat/flat_searcher_test.cpp:29-78).
#include <stdlib.h>
#include <unistd.h>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>
#include <zvec/core/framework/index_builder.h>
#include <zvec/core/framework/index_cluster.h>
#include <zvec/core/framework/index_features.h>
#include <zvec/core/framework/index_converter.h>
#include <zvec/core/framework/index_factory.h>
#include <zvec/core/framework/index_holder.h>
#include <zvec/core/framework/index_memory.h>
#include <zvec/core/framework/index_meta.h>
#include <zvec/core/framework/index_metric.h>
#include <zvec/core/framework/index_reformer.h>
#include <zvec/core/framework/index_plugin.h>
#include <zvec/core/framework/index_searcher.h>
#include <zvec/core/framework/index_segment_storage.h>
#include <zvec/core/framework/index_streamer.h>
#include <zvec/core/framework/index_trainer.h>
#include <zvec/core/interface/index.h>
#include <zvec/core/interface/index_factory.h>
#include "ivf/ivf_entity.h"
#include "ivf/ivf_index_format.h"

using namespace zvec::core;
using zvec::ailego::Params;

namespace {

// Zero-copy holder over caller memory: row i = base + i*elem, key = i.
class RawHolder : public IndexHolder {
 public:
  RawHolder(const void *p, size_t n, IndexMeta::DataType t, size_t dim, size_t elem)
      : p_((const char *)p), n_(n), t_(t), dim_(dim), elem_(elem) {}
  size_t count() const override { return n_; }
  size_t dimension() const override { return dim_; }
  IndexMeta::DataType data_type() const override { return t_; }
  size_t element_size() const override { return elem_; }
  bool multipass() const override { return true; }
  struct It : public IndexHolder::Iterator {
    const RawHolder *h; size_t i{0};
    explicit It(const RawHolder *hh) : h(hh) {}
    const void *data() const override { return h->p_ + i * h->elem_; }
    bool is_valid() const override { return i < h->n_; }
    uint64_t key() const override { return i; }
    void next() override { ++i; }
  };
  Iterator::Pointer create_iterator() override {
    return Iterator::Pointer(new It(this));
  }
 private:
  const char *p_; size_t n_; IndexMeta::DataType t_; size_t dim_, elem_;
};

struct RefIndex {
  IndexMeta in_meta;            // what the caller's rows / queries look like
  IndexMeta meta;               // what is stored (after converter)
  IndexConverter::Pointer converter;
  IndexReformer::Pointer reformer;
  IndexMetric::Pointer metric;
  IndexSearcher::Pointer searcher;
  IndexStreamer::Pointer streamer;  // set by zref_index_use_streamer: searches then go through it
  IndexStorage::Pointer streamer_storage;
  std::string path;
  std::string stored;           // converted rows, row-major, n * meta.element_size()
  size_t n{0};
  IndexRunner *runner() { return streamer ? static_cast<IndexRunner *>(streamer.get()) : searcher.get(); }
  ~RefIndex() {
    if (streamer) { streamer->close(); streamer->cleanup(); }
    if (searcher) { searcher->unload(); searcher->cleanup(); }
    if (!path.empty()) {
      ::unlink(path.c_str());
      IndexMemory::Instance()->remove(path);  // a ZREF_MEMORY_INDEX index
    }
  }
};

thread_local std::string g_err;
int fail(int code, const std::string &m) { g_err = m; return code < 0 ? code : -1; }

void parse_params(const char *kv, Params *p) {
  // "key=value;key=value" ; value typed by shape: int, float, bool, else string
  if (!kv) return;
  std::string s(kv);
  size_t pos = 0;
  while (pos < s.size()) {
    size_t e = s.find(';', pos);
    if (e == std::string::npos) e = s.size();
    std::string item = s.substr(pos, e - pos);
    pos = e + 1;
    size_t eq = item.find('=');
    if (eq == std::string::npos) continue;
    std::string k = item.substr(0, eq), v = item.substr(eq + 1);
    char *end = nullptr;
    long long iv = strtoll(v.c_str(), &end, 10);
    if (end && *end == 0 && !v.empty()) { p->set(k, (int64_t)iv); continue; }
    double dv = strtod(v.c_str(), &end);
    if (end && *end == 0 && !v.empty()) { p->set(k, (float)dv); continue; }
    if (v == "true" || v == "false") { p->set(k, v == "true"); continue; }
    p->set(k, v);
  }
}

}  // namespace

// An IndexTrainer that hands a builder centroids the caller already has (IVFBuilder::train(trainer),
// ivf_builder.cc:272-330): the way to build a stock IVF index and a GPU-built one from the SAME centroids.
class FixedCentroidTrainer : public IndexTrainer {
 public:
  FixedCentroidTrainer(const IndexMeta &meta, const void *cents, size_t count) : meta_(meta) {
    IndexCluster::CentroidList list;
    for (size_t i = 0; i < count; ++i)
      list.emplace_back((const char *)cents + i * meta.element_size(), meta.element_size());
    IndexCluster::Serialize(meta_, list, &bundle_);
  }
  int init(const IndexMeta &, const Params &) override { return 0; }
  int cleanup(void) override { return 0; }
  int train(IndexThreads::Pointer, IndexHolder::Pointer) override { return 0; }
  int load(IndexStorage::Pointer) override { return 0; }
  int dump(const IndexDumper::Pointer &) override { return 0; }
  const IndexMeta &meta(void) const override { return meta_; }
  const Stats &stats(void) const override { return stats_; }
  IndexBundle::Pointer indexes(void) const override { return bundle_; }

 private:
  IndexMeta meta_;
  Stats stats_;
  IndexBundle::Pointer bundle_;
};
static std::string g_fixed_centroids;   // zref_set_fixed_centroids: the next zref_index_create builds from these
static size_t g_fixed_count = 0;

extern "C" {

const char *zref_last_error(void) { return g_err.c_str(); }

// centroids (stored layout of the index to come, `count` of `elem` bytes) for the next zref_index_create calls:
// the builder is trained through train(trainer) instead of clustering; count = 0 clears
void zref_set_fixed_centroids(const void *cents, uint64_t count, uint32_t elem) {
  g_fixed_centroids.assign((const char *)cents, (size_t)count * elem);
  g_fixed_count = count;
}

// kind: "Flat" | "IVF".  dtype: IndexMeta::DataType of the CALLER's rows.
// converter: "" or a registered converter name (e.g. Int8StreamingConverter,
// HalfFloatConverter, CosineNormalizeConverter).  column_major: 0 row / 1 column.
// build_params / search_params: "k=v;k=v".
void *zref_index_create(const char *kind, const void *rows, uint64_t n, uint32_t dim,
                        int dtype, const char *metric, const char *converter,
                        int column_major, const char *build_params,
                        const char *search_params) {
  auto ix = std::make_unique<RefIndex>();
  auto dt = (IndexMeta::DataType)dtype;
  ix->in_meta.set_meta(dt, dim);
  ix->in_meta.set_metric(metric, 0, Params());
  ix->meta = ix->in_meta;
  ix->n = n;
  IndexHolder::Pointer holder =
      std::make_shared<RawHolder>(rows, n, dt, dim, ix->in_meta.element_size());
  if (converter && *converter) {
    ix->converter = IndexFactory::CreateConverter(converter);
    if (!ix->converter) { fail(-1, "no converter"); return nullptr; }
    if (ix->converter->init(ix->in_meta, Params()) != 0) { fail(-1, "converter init"); return nullptr; }
    if (IndexConverter::TrainAndTransform(ix->converter, holder) != 0) { fail(-1, "converter run"); return nullptr; }
    holder = ix->converter->result();
    ix->meta = ix->converter->meta();
    if (!ix->meta.reformer_name().empty()) {
      ix->reformer = IndexFactory::CreateReformer(ix->meta.reformer_name());
      if (!ix->reformer || ix->reformer->init(ix->meta.reformer_params()) != 0) { fail(-1, "reformer"); return nullptr; }
    }
  }
  ix->meta.set_major_order(column_major ? IndexMeta::MO_COLUMN : IndexMeta::MO_ROW);
  // keep the stored bytes (what a device store must ingest); ZREF_NO_STORED_COPY=1 skips the copy (full-size CPU arm)
  const char *nocopy = getenv("ZREF_NO_STORED_COPY");
  if (!(nocopy && atoi(nocopy) != 0)) {
    size_t es = ix->meta.element_size();
    ix->stored.resize(n * es);
    auto it = holder->create_iterator();
    size_t i = 0;
    for (; it->is_valid() && i < n; it->next(), ++i)
      std::memcpy(&ix->stored[i * es], it->data(), es);
  }
  ix->metric = IndexFactory::CreateMetric(ix->meta.metric_name());
  if (!ix->metric || ix->metric->init(ix->meta, ix->meta.metric_params()) != 0) { fail(-1, "metric"); return nullptr; }

  // kind = "Flat" | "IVF" | "<Builder kind>:<searcher class>" (e.g. "Flat:GpuFlatSearcher" after
  // zref_load_plugin: the index is built by the reference, searched by a plugin class)
  std::string k(kind), searcher_name;
  {
    size_t c = k.find(':');
    if (c != std::string::npos) { searcher_name = k.substr(c + 1); k = k.substr(0, c); }
    else searcher_name = k + "Searcher";
  }
  static std::atomic<int> seq{0};
  char buf[256];
  snprintf(buf, sizeof buf, "/tmp/zref_%d_%d.index", (int)getpid(), seq++);
  ix->path = buf;
  // ZREF_MEMORY_INDEX=1: MemoryDumper / MemoryReadStorage instead of a file under /tmp (the full-size CPU arm of
  // bench.py: 15 GB must not depend on the box's disk)
  const char *memenv = getenv("ZREF_MEMORY_INDEX");
  const bool in_memory = memenv && atoi(memenv) != 0;
  auto builder = IndexFactory::CreateBuilder(k + "Builder");
  auto dumper = IndexFactory::CreateDumper(in_memory ? "MemoryDumper" : "FileDumper");
  if (!builder || !dumper) { fail(-1, "no builder"); return nullptr; }
  Params bp; parse_params(build_params, &bp);
  int r = builder->init(ix->meta, bp);
  if (r != 0) { fail(r, "builder init"); return nullptr; }
  if (dumper->create(ix->path) != 0) { fail(-1, "dumper create"); return nullptr; }
  if (g_fixed_count) {
    IndexMeta tm = ix->meta;   // what IVFBuilder::train(trainer) compares with its converted meta
    tm.set_reformer(std::string(), 0, Params());
    tm.set_converter(std::string(), 0, Params());
    if (g_fixed_centroids.size() != g_fixed_count * tm.element_size()) { fail(-1, "fixed centroids: element size"); return nullptr; }
    auto trainer = std::make_shared<FixedCentroidTrainer>(tm, g_fixed_centroids.data(), g_fixed_count);
    r = builder->train(trainer);
    if (r == 0) r = builder->build(holder);
    if (r == 0) r = builder->dump(dumper);
  } else {
    r = IndexBuilder::TrainBuildAndDump(builder, holder, dumper);
  }
  if (r != 0) { fail(r, "build"); return nullptr; }
  dumper->close();
  ix->searcher = IndexFactory::CreateSearcher(searcher_name);
  auto storage = IndexFactory::CreateStorage(in_memory ? "MemoryReadStorage" : "MMapFileReadStorage");
  Params sp; parse_params(search_params, &sp);
  if (!ix->searcher || !storage) { fail(-1, "no searcher"); return nullptr; }
  if (ix->searcher->init(sp) != 0) { fail(-1, "searcher init"); return nullptr; }
  if (storage->open(ix->path, false) != 0) { fail(-1, "storage open"); return nullptr; }
  r = ix->searcher->load(storage, IndexMetric::Pointer());
  if (r != 0) { fail(r, "searcher load"); return nullptr; }
  return ix.release();
}

void zref_index_destroy(void *h) { delete (RefIndex *)h; }

// Replace the index's searcher by another registered class over the SAME file (an IVF build trains
// k-means with threads: two builds of the same rows are two different indexes).
int zref_index_use_searcher(void *h, const char *searcher_name, const char *search_params) {
  auto ix = (RefIndex *)h;
  auto sr = IndexFactory::CreateSearcher(searcher_name);
  auto storage = IndexFactory::CreateStorage("MMapFileReadStorage");
  if (!sr || !storage) return fail(-1, "no searcher");
  Params sp; parse_params(search_params, &sp);
  int r = sr->init(sp);
  if (r != 0) return fail(r, "searcher init");
  if (storage->open(ix->path, false) != 0) return fail(-1, "storage open");
  r = sr->load(storage, IndexMetric::Pointer());
  if (r != 0) return fail(r, "searcher load");
  if (ix->searcher) { ix->searcher->unload(); ix->searcher->cleanup(); }
  ix->searcher = sr;
  return 0;
}

// Search the index's file through a registered IndexStreamer class instead (IVFStreamer and the like are
// read-only streamers: init(meta, params) -> open(storage), as core_interface::IVFIndex does,
// src/core/interface/indexes/ivf_index.cc:38-39).  An empty class name goes back to the searcher.
int zref_index_use_streamer(void *h, const char *streamer_name, const char *search_params) {
  auto ix = (RefIndex *)h;
  if (ix->streamer) { ix->streamer->close(); ix->streamer->cleanup(); ix->streamer.reset(); ix->streamer_storage.reset(); }
  if (!streamer_name || !*streamer_name) return 0;
  auto st = IndexFactory::CreateStreamer(streamer_name);
  auto storage = IndexFactory::CreateStorage("MMapFileReadStorage");
  if (!st || !storage) return fail(-1, "no streamer");
  Params sp; parse_params(search_params, &sp);
  int r = st->init(ix->meta, sp);
  if (r != 0) return fail(r, "streamer init");
  if (storage->open(ix->path, false) != 0) return fail(-1, "storage open");
  r = st->open(storage);
  if (r != 0) return fail(r, "streamer open");
  ix->streamer = st;
  ix->streamer_storage = storage;
  return 0;
}
// 1 if IndexFactory has a streamer class of that name
int zref_has_streamer(const char *name) { return IndexFactory::HasStreamer(name) ? 1 : 0; }

// Load an index plugin with the reference's OWN loader (IndexPluginBroker -> IndexPlugin::load
// -> dlopen, src/core/framework/index_plugin.cc:20-56), as tools/core/bench.cc:796-803 does
// for its trailing argv.  Returns 0 or -1 (zref_last_error()).
int zref_load_plugin(const char *path) {
  static IndexPluginBroker broker;
  std::string err;
  if (!broker.emplace(std::string(path), &err)) return fail(-1, "plugin load: " + err);
  return 0;
}
// 1 if IndexFactory has a searcher class of that name
int zref_has_searcher(const char *name) { return IndexFactory::HasSearcher(name) ? 1 : 0; }

// Set a key filter on the next searches: drop keys with (key % mod) == rem (mod = 0: no filter).
// Exercises IndexContext::set_filter (index_context.h) through whichever searcher is loaded.
static thread_local uint64_t g_fmod = 0, g_frem = 0;
static thread_local float g_threshold = 0.f;
static thread_local int g_has_threshold = 0;
void zref_set_filter(uint64_t mod, uint64_t rem) { g_fmod = mod; g_frem = rem; }
// Group-by on the next streamer searches: group id = decimal string of (key % mod); group_num groups of group_topk
// documents (IndexContext::set_group_params / set_group_by).  mod = 0: off.
static thread_local uint64_t g_gmod = 0;
static thread_local uint32_t g_gnum = 0, g_gtopk = 0;
void zref_set_group_by(uint64_t mod, uint32_t group_num, uint32_t group_topk) { g_gmod = mod; g_gnum = group_num; g_gtopk = group_topk; }
void zref_set_threshold(int on, float v) { g_has_threshold = on; g_threshold = v; }

// Stored (post-converter) layout: dtype, dimension, element bytes, metric name.
int zref_index_info(void *h, int *dtype, uint32_t *dim, uint32_t *elem, uint64_t *n,
                    char *metric_name, uint32_t cap) {
  auto ix = (RefIndex *)h;
  *dtype = (int)ix->meta.data_type();
  *dim = ix->meta.dimension();
  *elem = ix->meta.element_size();
  *n = ix->n;
  if (metric_name && cap) snprintf(metric_name, cap, "%s", ix->meta.metric_name().c_str());
  return 0;
}

// Copy the stored rows (row-major, n*elem bytes) out.
int zref_index_stored(void *h, void *out, uint64_t cap) {
  auto ix = (RefIndex *)h;
  if (cap < ix->stored.size()) return fail(-31, "cap");
  std::memcpy(out, ix->stored.data(), ix->stored.size());
  return 0;
}

// Transform caller-format queries to the stored format (reformer->transform);
// identity copy when there is no reformer.  Returns bytes per query or <0.
int zref_index_transform_queries(void *h, const void *queries, uint32_t count,
                                 void *out, uint64_t cap) {
  auto ix = (RefIndex *)h;
  IndexQueryMeta qm(ix->in_meta.data_type(), ix->in_meta.dimension());
  size_t qes = ix->in_meta.element_size();
  size_t oes = ix->meta.element_size();
  if (cap < (uint64_t)count * oes) return fail(-31, "cap");
  for (uint32_t i = 0; i < count; ++i) {
    const char *q = (const char *)queries + i * qes;
    if (ix->reformer) {
      std::string o; IndexQueryMeta om;
      int r = ix->reformer->transform(q, qm, &o, &om);
      if (r != 0) return fail(r, "transform");
      if (o.size() != oes) return fail(-1, "transform size");
      std::memcpy((char *)out + i * oes, o.data(), oes);
    } else {
      std::memcpy((char *)out + i * oes, q, oes);
    }
  }
  return (int)oes;
}

// Search `count` caller-format queries, `batch` at a time through
// search_impl(q, qmeta, batch, ctx) on `threads` threads (own context each).
// out_keys/out_scores are count*topk (padded with key=~0, score=NaN);
// out_index = row index inside the index.  normalize!=0 applies
// metric->normalize like index.cc:624-628.
int zref_index_search(void *h, const void *queries, uint32_t count, uint32_t topk,
                      uint32_t batch, uint32_t threads, int normalize,
                      uint64_t *out_keys, float *out_scores, uint32_t *out_index,
                      uint32_t *out_counts) {
  auto ix = (RefIndex *)h;
  size_t oes = ix->meta.element_size();
  std::string tq((size_t)count * oes, '\0');
  int r = zref_index_transform_queries(h, queries, count, &tq[0], tq.size());
  if (r < 0) return r;
  IndexQueryMeta qm(ix->meta.data_type(), ix->meta.dimension());
  if (batch == 0) batch = 1;
  if (threads == 0) threads = 1;
  uint32_t nb = (count + batch - 1) / batch;
  std::atomic<uint32_t> next{0};
  std::atomic<int> err{0};
  const uint64_t fmod = g_fmod, frem = g_frem;
  const int has_thr = g_has_threshold;
  const float thr = g_threshold;
  auto work = [&]() {
    IndexRunner *run = ix->runner();
    auto ctx = run->create_context();
    if (!ctx) { err = -1; return; }
    ctx->set_topk(topk);
    if (fmod) ctx->set_filter([fmod, frem](uint64_t key) { return key % fmod == frem; });
    if (has_thr) ctx->set_threshold(thr);
    for (;;) {
      uint32_t b = next.fetch_add(1);
      if (b >= nb) break;
      uint32_t q0 = b * batch, qn = std::min(batch, count - q0);
      int rc = (qn == 1) ? run->search_impl(&tq[(size_t)q0 * oes], qm, ctx)
                         : run->search_impl(&tq[(size_t)q0 * oes], qm, qn, ctx);
      if (rc != 0) { err = rc; break; }
      for (uint32_t i = 0; i < qn; ++i) {
        const auto &res = ctx->result(i);
        size_t o = (size_t)(q0 + i) * topk;
        uint32_t c = (uint32_t)std::min<size_t>(res.size(), topk);
        if (out_counts) out_counts[q0 + i] = c;
        for (uint32_t j = 0; j < topk; ++j) {
          if (j < c) {
            float s = res[j].score();
            if (normalize && ix->metric->support_normalize()) ix->metric->normalize(&s);
            out_keys[o + j] = res[j].key(); out_scores[o + j] = s;
            if (out_index) out_index[o + j] = res[j].index();
          } else {
            out_keys[o + j] = ~0ull; out_scores[o + j] = __builtin_nanf("");
            if (out_index) out_index[o + j] = ~0u;
          }
        }
      }
    }
  };
  if (threads == 1) work();
  else {
    std::vector<std::thread> ts;
    for (uint32_t t = 0; t < threads; ++t) ts.emplace_back(work);
    for (auto &t : ts) t.join();
  }
  return err.load();
}

// search_bf_by_p_keys_impl through whichever searcher / streamer the index holds: the same key list for every query
// (FlatSearcher::search_bf_by_p_keys_impl, flat_searcher.cc:213-237).  With zref_set_group_by the call is a group-by
// search and the groups come back flattened.  Stored-format queries are made by the index's reformer.
int zref_index_search_keys(void *h, const void *queries, uint32_t count, const uint64_t *pks, uint32_t npks,
                           uint32_t topk, uint64_t *out_keys, float *out_scores, uint32_t *out_counts) {
  auto ix = (RefIndex *)h;
  size_t oes = ix->meta.element_size();
  std::string tq((size_t)count * oes, '\0');
  int r = zref_index_transform_queries(h, queries, count, &tq[0], tq.size());
  if (r < 0) return r;
  IndexQueryMeta qm(ix->meta.data_type(), ix->meta.dimension());
  IndexRunner *run = ix->runner();
  auto ctx = run->create_context();
  if (!ctx) return fail(-1, "context");
  ctx->set_topk(topk);
  const uint64_t fmod = g_fmod, frem = g_frem, gm = g_gmod;
  if (fmod) ctx->set_filter([fmod, frem](uint64_t key) { return key % fmod == frem; });
  if (gm) {
    ctx->set_group_params(g_gnum, g_gtopk);
    ctx->set_group_by([gm](uint64_t key) { return std::to_string(key % gm); });
  }
  int rc;
  if (pks) {
    std::vector<std::vector<uint64_t>> keys(count, std::vector<uint64_t>(pks, pks + npks));
    rc = run->search_bf_by_p_keys_impl(tq.data(), keys, qm, count, ctx);
  } else {
    rc = run->search_impl(tq.data(), qm, count, ctx);
  }
  if (rc != 0) return fail(rc, "search");
  for (uint32_t i = 0; i < count; ++i) {
    uint32_t c = 0;
    auto put = [&](const IndexDocument &doc) {
      if (c >= topk) return;
      out_keys[(size_t)i * topk + c] = doc.key();
      out_scores[(size_t)i * topk + c] = doc.score();
      ++c;
    };
    if (gm) { for (const auto &grp : ctx->group_result(i)) for (const auto &doc : grp.docs()) put(doc); }
    else { for (const auto &doc : ctx->result(i)) put(doc); }
    out_counts[i] = c;
    for (uint32_t j = c; j < topk; ++j) { out_keys[(size_t)i * topk + j] = ~0ull; out_scores[(size_t)i * topk + j] = __builtin_nanf(""); }
  }
  return 0;
}

// IndexSearcher::get_vector(key) and create_provider()->get_vector(key): the stored bytes of one document, twice
// (out = 2 * element_size bytes).  Returns element_size, or < 0 when either lookup fails.
int zref_index_get_vector(void *h, uint64_t key, void *out, uint64_t cap) {
  auto ix = (RefIndex *)h;
  const size_t es = ix->meta.element_size();
  if (cap < 2 * es) return fail(-31, "cap");
  IndexRunner *run = ix->runner();
  const void *v = run->get_vector(key);
  if (!v) return fail(-1, "get_vector");
  std::memcpy(out, v, es);
  auto prov = run->create_provider();
  if (!prov) return fail(-1, "create_provider");
  const void *pv = prov->get_vector(key);
  if (!pv) return fail(-1, "provider get_vector");
  std::memcpy((char *)out + es, pv, es);
  if (prov->count() != ix->n) return fail(-1, "provider count");
  return (int)es;
}

// ---- streamers (rows arrive between searches): drives any registered IndexStreamer class the way
// tests/core/algorithm/flat/flat_streamer_test.cc:180-212,1038-1060 does ------------------------------
struct RefStreamer {
  IndexStreamer::Pointer st;
  IndexStorage::Pointer storage;
  IndexContext::Pointer ctx;
  IndexMeta meta;
};

void *zref_streamer_open(const char *cls, int dtype, uint32_t dim, const char *metric, const char *path) {
  auto h = new RefStreamer();
  h->meta = IndexMeta((IndexMeta::DataType)dtype, dim);
  h->meta.set_metric(metric, 0, Params());
  h->st = IndexFactory::CreateStreamer(cls);
  if (!h->st) { delete h; fail(-1, std::string("no streamer class ") + cls); return nullptr; }
  h->storage = IndexFactory::CreateStorage("MMapFileStorage");
  Params sp;
  int r = h->storage ? h->storage->init(sp) : -1;
  if (r == 0) r = h->storage->open(path, true);
  if (r == 0) r = h->st->init(h->meta, Params());
  if (r == 0) r = h->st->open(h->storage);
  if (r != 0) { delete h; fail(r, "streamer init/open"); return nullptr; }
  h->ctx = h->st->create_context();
  if (!h->ctx) { delete h; fail(-1, "streamer context"); return nullptr; }
  return h;
}

int zref_streamer_add(void *hh, uint64_t key, const void *vec) {
  auto h = (RefStreamer *)hh;
  IndexQueryMeta qm(h->meta.data_type(), h->meta.dimension());
  return h->st->add_impl(key, vec, qm, h->ctx);
}

int zref_streamer_add_with_id(void *hh, uint32_t id, const void *vec) {
  auto h = (RefStreamer *)hh;
  IndexQueryMeta qm(h->meta.data_type(), h->meta.dimension());
  return h->st->add_with_id_impl(id, vec, qm, h->ctx);
}

int zref_streamer_search(void *hh, const void *queries, uint32_t count, uint32_t topk, uint64_t *out_keys,
                         float *out_scores, uint32_t *out_counts) {
  auto h = (RefStreamer *)hh;
  IndexQueryMeta qm(h->meta.data_type(), h->meta.dimension());
  auto ctx = h->st->create_context();
  if (!ctx) return fail(-1, "context");
  ctx->set_topk(topk);
  const uint64_t fmod = g_fmod, frem = g_frem;
  if (fmod) ctx->set_filter([fmod, frem](uint64_t key) { return key % fmod == frem; });
  if (g_has_threshold) ctx->set_threshold(g_threshold);
  if (g_gmod) {
    const uint64_t gm = g_gmod;
    ctx->set_group_params(g_gnum, g_gtopk);
    ctx->set_group_by([gm](uint64_t key) { return std::to_string(key % gm); });
  }
  int rc = count == 1 ? h->st->search_impl(queries, qm, ctx) : h->st->search_impl(queries, qm, count, ctx);
  if (rc != 0) return fail(rc, "streamer search");
  if (g_gmod) {  // groups flattened in result order: out_counts[i] = documents over all groups of query i
    for (uint32_t i = 0; i < count; ++i) {
      uint32_t c = 0;
      for (const auto &grp : ctx->group_result(i))
        for (const auto &doc : grp.docs()) {
          if (c >= topk) break;
          out_keys[(size_t)i * topk + c] = doc.key();
          out_scores[(size_t)i * topk + c] = doc.score();
          ++c;
        }
      out_counts[i] = c;
      for (uint32_t j = c; j < topk; ++j) { out_keys[(size_t)i * topk + j] = ~0ull; out_scores[(size_t)i * topk + j] = __builtin_nanf(""); }
    }
    return 0;
  }
  for (uint32_t i = 0; i < count; ++i) {
    const auto &res = ctx->result(i);
    const uint32_t c = (uint32_t)std::min<size_t>(res.size(), topk);
    out_counts[i] = c;
    for (uint32_t j = 0; j < topk; ++j) {
      out_keys[(size_t)i * topk + j] = j < c ? res[j].key() : ~0ull;
      out_scores[(size_t)i * topk + j] = j < c ? res[j].score() : __builtin_nanf("");
    }
  }
  return 0;
}

// search_bf_by_p_keys_impl: the same key list for every query (core_interface::Index::_dense_search with bf_pks,
// src/core/interface/index.cc:603-610)
int zref_streamer_search_keys(void *hh, const void *queries, uint32_t count, const uint64_t *pks, uint32_t npks,
                              uint32_t topk, uint64_t *out_keys, float *out_scores, uint32_t *out_counts) {
  auto h = (RefStreamer *)hh;
  IndexQueryMeta qm(h->meta.data_type(), h->meta.dimension());
  auto ctx = h->st->create_context();
  if (!ctx) return fail(-1, "context");
  ctx->set_topk(topk);
  const uint64_t fmod = g_fmod, frem = g_frem;
  if (fmod) ctx->set_filter([fmod, frem](uint64_t key) { return key % fmod == frem; });
  std::vector<std::vector<uint64_t>> keys(count, std::vector<uint64_t>(pks, pks + npks));
  int rc = h->st->search_bf_by_p_keys_impl(queries, keys, qm, count, ctx);
  if (rc != 0) return fail(rc, "streamer search_bf_by_p_keys");
  for (uint32_t i = 0; i < count; ++i) {
    const auto &res = ctx->result(i);
    const uint32_t c = (uint32_t)std::min<size_t>(res.size(), topk);
    out_counts[i] = c;
    for (uint32_t j = 0; j < topk; ++j) {
      out_keys[(size_t)i * topk + j] = j < c ? res[j].key() : ~0ull;
      out_scores[(size_t)i * topk + j] = j < c ? res[j].score() : __builtin_nanf("");
    }
  }
  return 0;
}

int zref_streamer_close(void *hh, int flush_first) {
  auto h = (RefStreamer *)hh;
  int r = 0;
  if (flush_first) r = h->st->flush(0);
  h->ctx.reset();
  int r2 = h->st->close();
  h->st->cleanup();
  if (h->storage) h->storage->close();
  delete h;
  return r ? r : r2;
}

// Export an IVF index the reference built (what a device store must ingest, via the
// reference's own IVFEntity reader, ivf_entity.h:89-170): centroids (row cid = centroid cid, in
// the centroid index's stored layout), posting lists as CSR (list_offsets[nlist+1]), rows in
// list order (stored layout, row-major) and their keys.  Call with NULL buffers for the sizes.
int zref_ivf_export(void *h, uint32_t *nlist, uint64_t *nrows, uint32_t *elem, uint32_t *c_elem,
                    int *c_dtype, uint32_t *c_dim, char *c_metric, uint32_t c_metric_cap,
                    void *centroids, uint64_t *list_offsets, void *rows, uint64_t *keys) {
  auto ix = (RefIndex *)h;
  auto storage = IndexFactory::CreateStorage("MMapFileReadStorage");
  if (!storage || storage->open(ix->path, false) != 0) return fail(-1, "storage open");
  IVFEntity ent;
  int r = ent.load(storage);
  if (r != 0) return fail(r, "entity load");
  const size_t nl = ent.inverted_list_count(), n = ent.vector_count(), es = ent.meta().element_size();
  *nlist = (uint32_t)nl; *nrows = n; *elem = (uint32_t)es;
  // centroid index = nested Flat index in segment "ivf.centroid" (ivf_searcher.cc:62-86)
  auto seg = storage->get(IVF_CENTROID_SEG_ID, 0);
  if (!seg) return fail(-1, "no centroid segment");
  IndexStorage::Pointer nested = std::make_shared<IndexSegmentStorage>(seg);
  if (nested->open(std::string(), false) != 0) return fail(-1, "nested open");
  IndexMeta cmeta;
  r = IndexHelper::DeserializeFromStorage(nested.get(), &cmeta);
  if (r != 0) return fail(r, "centroid meta");
  const size_t ces = cmeta.element_size();
  *c_elem = (uint32_t)ces; *c_dtype = (int)cmeta.data_type(); *c_dim = cmeta.dimension();
  if (c_metric && c_metric_cap) snprintf(c_metric, c_metric_cap, "%s", cmeta.metric_name().c_str());
  if (!centroids) return 0;
  auto fseg = nested->get("flat.features"), kseg = nested->get("flat.keys");
  if (!fseg || !kseg) return fail(-1, "centroid index is not a Flat index");
  const size_t nc = kseg->data_size() / sizeof(uint64_t);
  if (nc != nl || fseg->data_size() != nc * ces) return fail(-1, "centroid count mismatch");
  const void *fp = nullptr; const uint64_t *kp = nullptr;
  if (fseg->read(0, &fp, fseg->data_size()) != fseg->data_size()) return fail(-122, "read");
  if (kseg->read(0, (const void **)&kp, kseg->data_size()) != kseg->data_size()) return fail(-122, "read");
  const bool col = cmeta.major_order() == IndexMeta::MO_COLUMN;
  const size_t us = IndexMeta::AlignSizeof(cmeta.data_type()), units = ces / us;
  for (size_t i = 0; i < nc; ++i) {
    char *dst = (char *)centroids + (size_t)kp[i] * ces;
    if (kp[i] >= nc) return fail(-1, "centroid key out of range");
    const size_t blk = i / 32 * 32;
    if (col && blk + 32 <= nc) {
      const char *b = (const char *)fp + blk * ces;
      for (size_t u = 0; u < units; ++u) std::memcpy(dst + u * us, b + (u * 32 + (i - blk)) * us, us);
    } else {
      std::memcpy(dst, (const char *)fp + i * ces, ces);
    }
  }
  size_t pos = 0;
  for (size_t l = 0; l < nl; ++l) {
    auto lm = ent.inverted_list_meta(l);
    if (!lm) return fail(-122, "list meta");
    list_offsets[l] = pos;
    auto lk = ent.get_keys(lm->id_offset, lm->vector_count);
    if (!lk && lm->vector_count) return fail(-122, "keys");
    for (size_t j = 0; j < lm->vector_count; ++j, ++pos) {
      const void *v = ent.get_vector(lm->id_offset + j);
      if (!v) return fail(-122, "vector");
      std::memcpy((char *)rows + pos * es, v, es);
      keys[pos] = lk[j];
    }
  }
  list_offsets[nl] = pos;
  return pos == n ? 0 : fail(-1, "row count mismatch");
}

// Raw 1x1 distance of the index's metric (IndexMetric::distance()) between
// stored-format rows m[i] and ONE stored-format query: out[i], i<n.
int zref_index_distances(void *h, const void *m, uint64_t n, const void *q, float *out) {
  auto ix = (RefIndex *)h;
  auto f = ix->metric->distance();
  if (!f) return fail(-12, "no distance");
  size_t es = ix->meta.element_size();
  for (uint64_t i = 0; i < n; ++i)
    f((const char *)m + i * es, q, ix->meta.dimension(), out + i);
  return 0;
}

// Stand-alone metric distance: metric name + dtype + dimension (stored dim,
// tail included) [+ origin metric for QuantizedInteger].
int zref_metric_distances(const char *metric, const char *origin_metric, int dtype,
                          uint32_t dim, const void *m, uint64_t n, const void *q,
                          float *out) {
  IndexMeta meta;
  meta.set_meta((IndexMeta::DataType)dtype, dim);
  Params mp;
  if (origin_metric && *origin_metric)
    mp.set("proxima.quantized_integer.metric.origin_metric_name", std::string(origin_metric));
  meta.set_metric(metric, 0, mp);
  auto mt = IndexFactory::CreateMetric(metric);
  if (!mt) return fail(-12, "no metric");
  int r = mt->init(meta, mp);
  if (r != 0) return fail(r, "metric init");
  auto f = mt->distance();
  if (!f) return fail(-12, "no distance");
  size_t es = meta.element_size();
  for (uint64_t i = 0; i < n; ++i) f((const char *)m + i * es, q, dim, out + i);
  return 0;
}

// The reference's k-means (KmeansCluster::cluster / label, src/core/algorithm/cluster/kmeans_cluster.cc:469-535,
// 571-610) on caller rows from caller centroids: `centroids` holds k initial centroids going in and the
// updated ones coming out (an empty cluster's mean is 0/0 = NaN, vector_mean.h:217-226; nothing is purged);
// follows[k] = rows per cluster and *sse = the cost of the LAST clustering pass; labels[n] = label() against the
// returned centroids.  Returns 0 or a negative IndexError.
int zref_kmeans(const char *metric, const float *rows, uint64_t n, uint32_t dim, float *centroids, uint32_t k,
                uint32_t max_iterations, uint32_t threads, uint32_t *labels, uint64_t *follows, double *sse) {
  IndexMeta meta(IndexMeta::DT_FP32, dim);
  meta.set_metric(metric, 0, Params());
  auto cl = IndexFactory::CreateCluster("KmeansCluster");
  if (!cl) return fail(-1, "no KmeansCluster");
  Params p;
  p.set("proxima.kmeans.cluster.max_iterations", max_iterations);
  p.set("proxima.general.cluster.thread_count", threads ? threads : 1u);
  p.set("proxima.kmeans.cluster.purge_empty", false);
  p.set("proxima.kmeans.cluster.epsilon", 0.0);
  int r = cl->init(meta, p);
  if (r != 0) return fail(r, "cluster init");
  auto feats = std::make_shared<CoherentIndexFeatures>(meta, rows, (size_t)n * dim * sizeof(float));
  r = cl->mount(feats);
  if (r != 0) return fail(r, "cluster mount");
  IndexCluster::CentroidList cents;
  for (uint32_t c = 0; c < k; ++c) cents.emplace_back(centroids + (size_t)c * dim, (size_t)dim * sizeof(float));
  r = cl->cluster(cents);
  if (r != 0) return fail(r, "cluster");
  if (cents.size() != k) return fail(-1, "centroids were purged");
  double total = 0.0;
  for (uint32_t c = 0; c < k; ++c) {
    std::memcpy(centroids + (size_t)c * dim, cents[c].buffer().data(), (size_t)dim * sizeof(float));
    if (follows) follows[c] = cents[c].follows();
    total += cents[c].score();
  }
  if (sse) *sse = total;
  if (labels) {
    std::vector<uint32_t> out;
    r = cl->label(cents, &out);
    if (r != 0) return fail(r, "label");
    std::memcpy(labels, out.data(), (size_t)n * sizeof(uint32_t));
  }
  cl->cleanup();
  return 0;
}

// ---- core_interface::Index (src/core/interface): the layer the DB calls (vector_column_indexer.cc:195).  The index
// creates its streamer BY NAME (indexes/flat_index.cc:31, indexes/ivf_index.cc:38-39), so after a plugin's
// zvb200_install_overrides() the very same calls reach the GPU classes. ----------------------------------------
namespace ci = zvec::core_interface;

// zvb200_index_search_batch of the loaded plugin (the plugin is dlopen'ed RTLD_LOCAL by IndexPlugin::load, so the
// caller hands the address over)
typedef int (*zref_batch_fn_t)(void *, const void *, uint32_t, const void *, void *);
static zref_batch_fn_t g_batch_fn = nullptr;
void zref_set_batch_fn(void *fn) { g_batch_fn = (zref_batch_fn_t)fn; }

struct CiIndex {
  ci::Index::Pointer index;
  std::string path;
  bool ivf{false}, keep_file{false};
  uint32_t dim{0};
  size_t in_es{0};
};

// kind "Flat" | "IVF"; metric "L2" | "IP" | "COSINE"; quantizer "" | "fp16" | "int8"; dtype of the caller's vectors
// open_existing != 0: open the index file at `path` as it is (an IVF index another object trained), read-only
void *zref_ci_create(const char *kind, int dtype, uint32_t dim, const char *metric, const char *quantizer, int nlist,
                     const char *path, int open_existing) {
  auto h = std::make_unique<CiIndex>();
  h->ivf = std::string(kind) == "IVF";
  std::shared_ptr<ci::BaseIndexParam> p;
  if (h->ivf) {
    auto ip = std::make_shared<ci::IVFIndexParam>();
    ip->nlist = nlist;
    p = ip;
  } else {
    p = std::make_shared<ci::FlatIndexParam>();
  }
  const std::string m(metric), qz(quantizer ? quantizer : "");
  p->metric_type = m == "L2" ? ci::MetricType::kL2sq : m == "IP" ? ci::MetricType::kInnerProduct : ci::MetricType::kCosine;
  p->dimension = (int)dim;
  p->data_type = (IndexMeta::DataType)dtype;
  p->quantizer_param.type = qz == "fp16" ? ci::QuantizerType::kFP16 : qz == "int8" ? ci::QuantizerType::kInt8
                                                                                   : ci::QuantizerType::kNone;
  h->index = ci::IndexFactory::CreateAndInitIndex(*p);
  if (!h->index) { fail(-1, "CreateAndInitIndex"); return nullptr; }
  h->path = path;
  h->dim = dim;
  h->in_es = IndexQueryMeta((IndexMeta::DataType)dtype, dim).element_size();
  if (!open_existing) ::unlink(path);
  ci::StorageOptions so;
  so.type = ci::StorageOptions::StorageType::kMMAP;
  so.create_new = !open_existing;
  so.read_only = open_existing != 0;
  h->keep_file = open_existing != 0;
  if (h->index->Open(path, so) != 0) { fail(-1, "Index::Open"); return nullptr; }
  return h.release();
}

int zref_ci_add(void *hh, const void *rows, uint64_t n, uint32_t first_id) {
  auto h = (CiIndex *)hh;
  for (uint64_t i = 0; i < n; ++i) {
    ci::VectorData v;
    v.vector = ci::DenseVector{(const char *)rows + i * h->in_es};
    int r = h->index->Add(v, first_id + (uint32_t)i);
    if (r != 0) return fail(r, "Index::Add");
  }
  return 0;
}

int zref_ci_train(void *hh) { return ((CiIndex *)hh)->index->Train(); }
uint32_t zref_ci_count(void *hh) { return ((CiIndex *)hh)->index->GetDocCount(); }

int zref_ci_fetch(void *hh, uint32_t doc_id, void *out, uint64_t cap) {
  auto h = (CiIndex *)hh;
  ci::VectorDataBuffer b;
  int r = h->index->Fetch(doc_id, &b);
  if (r != 0) return fail(r, "Index::Fetch");
  auto &d = std::get<ci::DenseVectorBuffer>(b.vector_buffer).data;
  if (d.size() > cap) return fail(-31, "cap");
  std::memcpy(out, d.data(), d.size());
  return (int)d.size();
}

void zref_ci_destroy(void *hh) {
  auto h = (CiIndex *)hh;
  if (h->index) h->index->Close();
  h->index.reset();
  if (!h->keep_file) ::unlink(h->path.c_str());
  delete h;
}

// Index::Search per query (batch = 0: what the reference's callers do, index.cc:617 count = 1) or ONE
// zvb200_index_search_batch call (batch = 1: integration/index_batch.cc, bound by symbol from the loaded plugin).
// bf_pks / npks: FlatQueryParam::bf_pks; refiner: another zref_ci index (the reference Flat index) + scale_factor;
// the filter of zref_set_filter applies; out_vectors (optional): count * topk * in_es bytes, the fetched vectors
// reverted to the input layout (SearchResult::reverted_vector_list_, or the stored bytes when no reformer reverts).
int zref_ci_search(void *hh, const void *queries, uint32_t count, uint32_t topk, int batch, const uint64_t *bf_pks,
                   uint32_t npks, int is_linear, float radius, void *refiner, float scale_factor, int fetch_vector,
                   uint64_t *out_keys, float *out_scores, uint32_t *out_counts, void *out_vectors) {
  auto h = (CiIndex *)hh;
  const uint64_t fmod = g_fmod, frem = g_frem;
  auto make_param = [&]() -> ci::BaseIndexQueryParam::Pointer {
    ci::BaseIndexQueryParam::Pointer p;
    if (h->ivf) { auto ip = std::make_shared<ci::IVFQueryParam>(); ip->nprobe = 0; p = ip; }
    else p = std::make_shared<ci::FlatQueryParam>();
    p->topk = topk;
    p->is_linear = is_linear != 0;
    p->radius = radius;
    p->fetch_vector = fetch_vector != 0;
    if (fmod) p->filter = std::make_shared<ci::IndexFilter>();
    if (fmod) p->filter->set([fmod, frem](uint64_t key) { return key % fmod == frem; });
    if (bf_pks) p->bf_pks = std::make_shared<std::vector<uint64_t>>(bf_pks, bf_pks + npks);
    if (refiner) {
      p->refiner_param = std::make_shared<ci::RefinerParam>();
      p->refiner_param->scale_factor_ = scale_factor;
      p->refiner_param->reference_index = ((CiIndex *)refiner)->index;
    }
    return p;
  };
  std::vector<ci::SearchResult> results(count);
  if (batch) {
    auto fn = g_batch_fn;
    if (!fn) return fail(-1, "zref_set_batch_fn(zvb200_index_search_batch of the loaded plugin) first");
    auto p = make_param();
    int r = fn(h->index.get(), queries, count, &p, &results);
    if (r != 0) return fail(r, "SearchBatch");
  } else {
    for (uint32_t i = 0; i < count; ++i) {
      auto p = make_param();  // Index::Search moves the filter out of its param (flat_index.cc:60-62)
      ci::VectorData v;
      v.vector = ci::DenseVector{(const char *)queries + i * h->in_es};
      int r = h->index->Search(v, p, &results[i]);
      if (r != 0) return fail(r, "Index::Search");
    }
  }
  for (uint32_t i = 0; i < count; ++i) {
    const auto &docs = results[i].doc_list_;
    const uint32_t c = (uint32_t)std::min<size_t>(docs.size(), topk);
    out_counts[i] = c;
    for (uint32_t j = 0; j < topk; ++j) {
      const size_t o = (size_t)i * topk + j;
      out_keys[o] = j < c ? docs[j].key() : ~0ull;
      out_scores[o] = j < c ? docs[j].score() : __builtin_nanf("");
      if (out_vectors && j < c) {
        char *dst = (char *)out_vectors + o * h->in_es;
        if (j < results[i].reverted_vector_list_.size())
          std::memcpy(dst, results[i].reverted_vector_list_[j].data(),
                      std::min(h->in_es, results[i].reverted_vector_list_[j].size()));
        else if (docs[j].vector())
          std::memcpy(dst, docs[j].vector(), h->in_es);
      }
    }
  }
  return 0;
}

}  // extern "C"
