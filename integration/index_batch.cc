// SearchBatch: see index_batch.h.  Compiled against the reference's headers where they lie (never copied) into the
// plugin library, next to the GPU searcher / streamer classes it feeds.
#include "index_batch.h"

#include <zvec/core/framework/index_error.h>
#include <zvec/core/framework/index_factory.h>

#include <cmath>
#include <map>
#include <mutex>

#include "algorithm/ivf/ivf_params.h"

namespace zvec::core_interface {

namespace {

// What Index::Init builds and keeps private (index.cc:47-191): the reformer and the (query) metric of the index,
// re-created from the streamer's own IndexMeta, plus one search context per calling thread (index.cc:24-45).
struct Side {
  core::IndexStreamer::Pointer streamer;
  core::IndexReformer::Pointer reformer;
  core::IndexMetric::Pointer metric;
  core::IndexQueryMeta input_meta;
  core::IndexContext::Pointer context;
};

Side *side_of(Index &index) {
  thread_local std::map<const Index *, Side> sides;
  auto streamer = index.index_searcher();
  if (!streamer) return nullptr;
  Side &s = sides[&index];
  if (s.streamer != streamer) {  // first use on this thread, or the index was re-created at this address
    s = Side();
    s.streamer = streamer;
    const core::IndexMeta &meta = streamer->meta();
    const auto p = index.GetParam();
    s.input_meta.set_meta(p->data_type, p->dimension);
    if (!meta.reformer_name().empty()) {
      s.reformer = core::IndexFactory::CreateReformer(meta.reformer_name());
      if (!s.reformer || s.reformer->init(meta.reformer_params()) != 0) return nullptr;
    }
    s.metric = core::IndexFactory::CreateMetric(meta.metric_name());
    if (!s.metric || s.metric->init(meta, meta.metric_params()) != 0) return nullptr;
    if (s.metric->query_metric()) s.metric = s.metric->query_metric();  // index.cc:101-103
  }
  if (!s.context) s.context = streamer->create_context();
  return s.context ? &s : nullptr;
}

// FlatIndex / IVFIndex::_prepare_for_search (indexes/flat_index.cc:46-68, indexes/ivf_index.cc:239-264) through the
// context's public interface; the caller's filter is copied, not moved out of its param
void prepare(core::IndexContext &ctx, const BaseIndexQueryParam &p, uint32_t topk) {
  ctx.set_topk(topk);
  ctx.set_fetch_vector(p.fetch_vector);
  if (p.filter) ctx.set_filter(*p.filter); else ctx.reset_filter();
  if (p.radius > 0.0f) ctx.set_threshold(p.radius); else ctx.reset_threshold();
  if (auto *ivf = dynamic_cast<const IVFQueryParam *>(&p)) {
    if (ivf->nprobe > 0) {
      ailego::Params params;
      params.set(core::PARAM_IVF_BUILDER_CENTROID_COUNT, ivf->nprobe);  // the key the reference sets (ivf_index.cc:257-262)
      ctx.update(params);
    }
  }
}

int run(Index &index, const void *queries, uint32_t count, const BaseIndexQueryParam &p, uint32_t topk,
        const std::vector<std::vector<uint64_t>> *keys, std::vector<SearchResult> *results) {
  if (!queries || !count || !results) return core::IndexError_InvalidArgument;
  Side *s = side_of(index);
  if (!s) return core::IndexError_Runtime;
  prepare(*s->context, p, topk);
  const size_t in_es = s->input_meta.element_size();
  // the reformed queries, contiguous (flat_searcher_context.h:875-884)
  std::string buf, one;
  core::IndexQueryMeta new_meta = s->input_meta;
  const void *q = queries;
  if (s->reformer) {
    for (uint32_t i = 0; i < count; ++i) {  // IndexReformer::transform per query, as _dense_search does (index.cc:592-600)
      if (s->reformer->transform(static_cast<const char *>(queries) + i * in_es, s->input_meta, &one, &new_meta) != 0)
        return core::IndexError_Runtime;
      buf.append(one);
    }
    q = buf.data();
  }
  int rc;
  if (keys) rc = s->streamer->search_bf_by_p_keys_impl(q, *keys, new_meta, count, s->context);
  else if (p.is_linear) rc = s->streamer->search_bf_impl(q, new_meta, count, s->context);
  else rc = s->streamer->search_impl(q, new_meta, count, s->context);  // ONE call: the GPU sees the batch
  if (rc != 0) return core::IndexError_Runtime;
  results->assign(count, SearchResult());
  for (uint32_t i = 0; i < count; ++i) {
    SearchResult &r = (*results)[i];
    const char *orig = static_cast<const char *>(queries) + i * in_es;
    if (s->reformer && s->context->fetch_vector() && s->reformer->need_revert()) {  // index.cc:634-647
      const auto &docs = s->context->result(i);
      r.reverted_vector_list_.resize(docs.size());
      for (size_t j = 0; j < docs.size(); ++j) {
        r.reverted_vector_list_[j].resize(s->input_meta.dimension() * s->input_meta.unit_size());
        if (s->reformer->revert(docs[j].vector(), new_meta, &r.reverted_vector_list_[j]) != 0)
          return core::IndexError_Runtime;
      }
    }
    r.doc_list_ = std::move(*s->context->mutable_result(i));
    if (s->metric->support_normalize())  // index.cc:624-628
      for (auto &doc : r.doc_list_) s->metric->normalize(doc.mutable_score());
    if (s->reformer && s->reformer->normalize(orig, s->input_meta, r.doc_list_) != 0) return core::IndexError_Runtime;
  }
  return 0;
}

}  // namespace

int SearchBatchByKeys(Index &index, const void *queries, uint32_t count, const BaseIndexQueryParam::Pointer &param,
                      const std::vector<std::vector<uint64_t>> &keys, std::vector<SearchResult> *results) {
  if (!param || keys.size() != count) return core::IndexError_InvalidArgument;
  return run(index, queries, count, *param, param->topk, &keys, results);
}

int SearchBatch(Index &index, const void *queries, uint32_t count, const BaseIndexQueryParam::Pointer &param,
                std::vector<SearchResult> *results) {
  if (!param) return core::IndexError_InvalidArgument;
  if (param->bf_pks) {  // index.cc:603-610: the listed keys only, the same list for every query
    std::vector<std::vector<uint64_t>> keys(count, *param->bf_pks);
    return run(index, queries, count, *param, param->topk, &keys, results);
  }
  if (!param->refiner_param) return run(index, queries, count, *param, param->topk, nullptr, results);
  // the refiner flow of Index::Search (index.cc:408-445): coarse top floor(topk * scale_factor) from this index,
  // then the exact brute force over those keys on the reference (Flat) index
  auto &ref = param->refiner_param->reference_index;
  if (!ref || ref->GetParam()->index_type != IndexType::kFlat) return core::IndexError_Runtime;
  float scale = param->refiner_param->scale_factor_;
  if (scale == 0) scale = 1;  // Index::_get_coarse_search_topk (index.cc:794-801)
  const uint32_t coarse_topk = (uint32_t)std::floor(param->topk * scale);
  auto coarse_param = param->Clone();
  coarse_param->refiner_param = nullptr;
  coarse_param->fetch_vector = false;
  std::vector<SearchResult> coarse;
  int rc = run(index, queries, count, *coarse_param, coarse_topk, nullptr, &coarse);
  if (rc != 0) return rc;
  std::vector<std::vector<uint64_t>> keys(count);
  for (uint32_t i = 0; i < count; ++i)
    for (auto &doc : coarse[i].doc_list_) keys[i].push_back(doc.key());
  auto flat_param = std::make_shared<FlatQueryParam>();
  flat_param->topk = param->topk;
  flat_param->fetch_vector = param->fetch_vector;
  flat_param->filter = param->filter;
  return run(*ref, queries, count, *flat_param, flat_param->topk, &keys, results);
}

}  // namespace zvec::core_interface

extern "C" int zvb200_index_search_batch(void *index, const void *queries, uint32_t count, const void *param,
                                         void *results) {
  using namespace zvec::core_interface;
  if (!index || !param || !results) return zvec::core::IndexError_InvalidArgument;
  return SearchBatch(*static_cast<Index *>(index), queries, count,
                     *static_cast<const BaseIndexQueryParam::Pointer *>(param),
                     static_cast<std::vector<SearchResult> *>(results));
}
