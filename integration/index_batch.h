// Batch entry above `core`: core_interface::Index::Search hands the streamer ONE query per call
// (src/core/interface/index.cc:617 hard-codes count = 1; tools/core/bench.cc:262-297 loops over a "batch"), so only
// C++ callers of IndexRunner::search_impl(query, qmeta, count, ctx) ever reach the batched tensor-core scan.
// SearchBatch is that missing entry: the flow of Index::Search / _dense_search (index.cc:380-446, 581-652) with the
// per-query steps looped around ONE search_impl(count) call.  It is written against the reference's PUBLIC interface
// only (Index::index_searcher(), Index::GetParam(), the streamer's IndexMeta), so it works on every existing
// core_interface::Index object without touching the reference tree; INTEGRATION.md §6 shows the same code as the
// member function a maintainer would add next to _dense_search.
#pragma once
#include <zvec/core/interface/index.h>

#include <vector>

namespace zvec::core_interface {

//! `count` dense queries (row-major, the index's INPUT element layout: GetParam()->data_type x dimension) -> one
//! SearchResult per query, as `count` calls of index.Search(query_i, param, &results[i]) would fill them.
//! param: topk / filter / radius / is_linear / fetch_vector / bf_pks (the same key list for every query) /
//! refiner_param are honoured as Index::Search honours them.  Returns 0 or a core::IndexError code.
int SearchBatch(Index &index, const void *queries, uint32_t count, const BaseIndexQueryParam::Pointer &param,
                std::vector<SearchResult> *results);

//! the refiner's second stage: per-query key lists (what Index::Search builds from the coarse result, index.cc:430-442)
int SearchBatchByKeys(Index &index, const void *queries, uint32_t count, const BaseIndexQueryParam::Pointer &param,
                      const std::vector<std::vector<uint64_t>> &keys, std::vector<SearchResult> *results);

}  // namespace zvec::core_interface

//! C entry for hosts that bind by symbol (tests/bench shims): index = core_interface::Index*, param =
//! const BaseIndexQueryParam::Pointer*, results = std::vector<SearchResult>*
extern "C" int zvb200_index_search_batch(void *index, const void *queries, uint32_t count, const void *param,
                                         void *results);
