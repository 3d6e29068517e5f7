#!/usr/bin/env python3
"""Golden vectors for multi-vector fusion, produced by the REFERENCE's own classes
(/root/reference/python/zvec/extension/multi_vector_reranker.py, loaded as a file: the zvec package
itself needs its compiled _zvec extension, which cannot be built here — SURVEY 8c — so the three names
the module imports are provided as minimal stand-ins: Doc = a namedtuple with id/score/_replace,
MetricType = an enum with L2/IP/COSINE, RerankFunction = the real base class file).
Run in the build container only:  python tests/golden/make_fusion_golden.py > tests/golden/fusion.json"""
import collections, enum, importlib.util, json, random, sys, types

REF = "/root/reference/python/zvec"
Doc = collections.namedtuple("Doc", ["id", "score"])


class MetricType(enum.Enum):
    L2 = 1
    IP = 2
    COSINE = 3


def load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


for pkg in ("zvec", "zvec.model", "zvec.extension"):
    m = types.ModuleType(pkg); m.__path__ = []; sys.modules[pkg] = m
doc_mod = types.ModuleType("zvec.model.doc"); doc_mod.Doc = Doc; sys.modules["zvec.model.doc"] = doc_mod
typ_mod = types.ModuleType("zvec.typing"); typ_mod.MetricType = MetricType; sys.modules["zvec.typing"] = typ_mod
load("zvec.extension.rerank_function", REF + "/extension/rerank_function.py")
rr = load("zvec.extension.multi_vector_reranker", REF + "/extension/multi_vector_reranker.py")

rng = random.Random(20250)
cases = []
for ci in range(24):
    F = rng.choice([1, 2, 3, 4])
    k = rng.choice([1, 5, 10, 100])
    universe = rng.choice([k, 2 * k + 3, 50 * k])          # small universes force overlaps and score ties
    lists = []
    for f in range(F):
        n = k if rng.random() < 0.7 else rng.randrange(0, k + 1)   # short and empty lists
        ids = rng.sample(range(1, universe + 1), min(n, universe))
        sc = sorted(round(rng.uniform(-3, 3), rng.choice([1, 6])) for _ in ids)
        lists.append([[i, s] for i, s in zip(ids, sc)])
    topn = rng.choice([1, 10, 100])
    rc = rng.choice([0, 60])
    weights = [rng.choice([1.0, 0.25, 2.0, 0.7]) for _ in range(F)]
    qr = {"field%d" % f: [Doc(str(i), float(s)) for i, s in lst] for f, lst in enumerate(lists)}
    out = {"rrf": [[int(d.id), d.score] for d in rr.RrfReRanker(topn=topn, rank_constant=rc).rerank(qr)]}
    for mname in ("L2", "IP", "COSINE"):
        w = {"field%d" % f: weights[f] for f in range(F)}
        res = rr.WeightedReRanker(topn=topn, metric=MetricType[mname], weights=w).rerank(qr)
        out["weighted_" + mname] = [[int(d.id), d.score] for d in res]
    cases.append({"lists": lists, "k": k, "topn": topn, "rank_constant": rc, "weights": weights, "expect": out})
json.dump({"source": "reference RrfReRanker / WeightedReRanker, python/zvec/extension/multi_vector_reranker.py",
           "cases": cases}, sys.stdout)
