"""TEST INFRASTRUCTURE ONLY (see oracle/README): CPU restatement of the reference's multi-vector
fusion, python/zvec/extension/multi_vector_reranker.py.

  rrf(lists, topn, rank_constant)          <- RrfReRanker.rerank            (:62-89)
  weighted(lists, topn, metric, weights)   <- WeightedReRanker.rerank       (:135-165)
  normalize_score(score, metric)           <- WeightedReRanker._normalize_score (:167-174)

`lists` = per field, in field order, a list of (doc_id, score) in rank order — what the query executor
hands the reranker (python/zvec/executor/query_executor.py:178-206).  Pinned against the reference's own
classes: tests/golden/fusion.json is produced by tests/golden/make_fusion_golden.py, which loads the
reference module itself."""
import heapq
import math


def normalize_score(score, metric):
    if metric == "L2":
        return 1.0 - 2 * math.atan(score) / math.pi
    if metric == "IP":
        return 0.5 + math.atan(score) / math.pi
    if metric == "COSINE":
        return 1.0 - score / 2.0
    raise ValueError("Unsupported metric type")


def _top(totals, topn):
    # heapq.nlargest keeps first-seen order among equal scores (dict order = first-seen order)
    return heapq.nlargest(topn, totals.items(), key=lambda x: x[1])


def rrf(lists, topn, rank_constant=60):
    totals = {}
    for lst in lists:
        for rank, (doc_id, _score) in enumerate(lst):
            totals[doc_id] = totals.get(doc_id, 0.0) + 1.0 / (rank_constant + rank + 1)
    return _top(totals, topn)


def weighted(lists, topn, metric, weights):
    totals = {}
    for f, lst in enumerate(lists):
        for doc_id, score in lst:
            totals[doc_id] = totals.get(doc_id, 0.0) + normalize_score(score, metric) * weights[f]
    return _top(totals, topn)
