"""IVF: the GPU coarse + posting-list scan against the reference's own IVFSearcher
(oracle/_ref) on an index the reference's IVFBuilder trained and packed; the device store ingests
the posting lists exactly as the reference's IVFEntity reads them (RefIndex.ivf_export)."""
import math

import numpy as np
import pytest

import support as S
import zvec_b200 as z
from zvec_b200.binding import FLT_MAX

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(S.ref() is None, reason="oracle/_ref not available")]


def clustered(n, d, nc, seed):
    rng = np.random.default_rng(seed)
    centres = rng.standard_normal((nc, d)).astype(np.float32)
    x = centres[np.arange(n) % nc] + 0.3 * rng.standard_normal((n, d)).astype(np.float32)
    q = centres[rng.integers(0, nc, 64)] + 0.3 * rng.standard_normal((64, d)).astype(np.float32)
    return x.astype(np.float32), q.astype(np.float32)


def budget(n, nlist, ratio, bf_threshold=1000):
    """IVFSearcherContext::update (ivf_searcher_context.h:61-79)"""
    nprobe = max(int(round(nlist * ratio)), 1)
    return nprobe, max(bf_threshold, int(math.ceil(n * ratio)))


@pytest.mark.parametrize("metric,conv,origin", [
    ("SquaredEuclidean", "", None), ("Cosine", "CosineNormalizeConverter", None),
    ("SquaredEuclidean", "HalfFloatConverter", None), ("SquaredEuclidean", "Int8StreamingConverter", "SquaredEuclidean"),
])
@pytest.mark.parametrize("ratio", [0.05, 0.25])
@pytest.mark.parametrize("kernel", [0, 4])
def test_ivf_search_equals_reference(metric, conv, origin, ratio, kernel):
    n, d, nlist, k = 30000, 48, 100, 10
    x, qs = clustered(n, d, nlist, 3)
    ri = S.RefIndex(x, S.DT_FP32, d, metric, conv, kind="IVF",
                    build_params="proxima.ivf.builder.centroid_count=%d" % nlist,
                    search_params="proxima.ivf.searcher.scan_ratio=%g" % ratio)
    e = ri.ivf_export()
    assert e["c_metric"] == ri.metric_name and e["c_dtype"] == ri.dtype and e["c_dim"] == ri.dim
    mid = z.metric_from_names(ri.metric_name, origin)
    st = z.Store(ri.dtype, mid, ri.dim)
    st.append(e["rows"], e["keys"]); st.seal()
    st.ivf_attach(e["centroids"], e["list_offsets"])
    ctx = z.Context(st)
    ctx.set_kernel(kernel)
    tq = ri.transform_queries(qs)
    nprobe, max_scan = budget(n, e["nlist"], ratio)
    rk, rs, _, rc = ri.search(qs, k)
    stored = ri.stored()                      # key order
    for batch in (1, len(qs)):
        for b0 in range(0, len(qs), batch):
            gk, gs, gi, gc = ctx.ivf_search(tq[b0:b0 + batch], k, nprobe, max_scan)
            assert ctx.last_kernel() == ("scan_lists_stream" if kernel == 4 else "scan_lists_lm" if batch >= 8 else "scan_lists")
            for j in range(gk.shape[0]):
                i = b0 + j
                assert gc[j] == rc[i]
                alld = S.oracle_distances(mid, ri.dtype, stored, ri.dim, tq[i])
                nv = int(gc[j])
                gkey = np.full(k, 0xFFFFFFFF, np.uint32); gkey[:nv] = gk[j][:nv].astype(np.uint32)
                rkey = np.full(k, 0xFFFFFFFF, np.uint32); rkey[:nv] = rk[i][:nv].astype(np.uint32)
                S.check_topk_parity(gkey, gs[j], rkey, rs[i], all_dist=alld, exact=ri.dtype == S.DT_INT8,
                                    what="ivf %s/%s ratio %g q%d" % (conv, metric, ratio, i))
                # index() = position in list order: keys[index] == key
                assert np.array_equal(e["keys"][gi[j][:nv].astype(np.int64)], gk[j][:nv])
    ri.close()


def test_ivf_argument_errors():
    st = z.Store(z.DT_FP32, z.METRIC_L2SQ, 8)
    st.append(np.zeros((10, 8), np.float32)); st.seal()
    ctx = z.Context(st)
    with pytest.raises(z.ZvecB200Error) as ei:
        ctx.ivf_search(np.zeros((1, 8), np.float32), 5, 4, 100)
    assert ei.value.code == -21                       # NoReady: nothing attached
    with pytest.raises(z.ZvecB200Error) as ei:
        st.ivf_attach(np.zeros((2, 8), np.float32), [0, 4, 9])
    assert ei.value.code == -24                       # Mismatch: offsets do not cover the rows
    st.ivf_attach(np.zeros((2, 8), np.float32), [0, 4, 10])
    k_, s_, i_, c_ = ctx.ivf_search(np.zeros((1, 8), np.float32), 5, 1, 1)
    assert c_[0] == 4                                  # one list of 4 rows scanned


@pytest.mark.parametrize("kind,d", [("fp32_l2", 48), ("fp16_ip", 200), ("qi8_l2", 96), ("fp32_cos", 1536)])
@pytest.mark.parametrize("k", [10, 100])
def test_list_major_scan_equals_query_major_bit_for_bit(kind, d, k):
    """K4b (scan_lists_lm.cu) streams every probed list once for up to 8 of the batch's queries; the scores come from
    the same arithmetic as the query-major scan, so ids AND scores are identical — with deleted rows, a filter, a tight
    scan budget, empty lists and lists probed by more queries than one pass takes."""
    from test_flat_gpu import make_stored
    n, nlist, q, nprobe = 20000, 50, 200, 9
    metric, dt, sdim, stored, conv = make_stored(kind, n, d, seed=31)
    rng = np.random.default_rng(32)
    cuts = np.sort(rng.choice(np.arange(1, n), nlist - 1, replace=False))
    offs = np.concatenate([[0], cuts, [n]]).astype(np.uint64)
    offs[7] = offs[6]                                                      # an empty list
    cent = np.ascontiguousarray(conv(rng.standard_normal((nlist, d)).astype(np.float32)))
    queries = np.ascontiguousarray(conv(rng.standard_normal((q, d)).astype(np.float32)))
    st = z.Store(dt, metric, sdim)
    st.append(stored); st.seal()
    st.ivf_attach(cent, offs)
    for r in (5, 77, 4000):
        st.set_deleted(r, True)
    fb = np.zeros((n + 31) // 32, np.uint32); fb[10:40] = 0x0f0f0f0f
    ctx = z.Context(st)
    for max_scan, filt in ((n, None), (2500, fb)):
        ctx.set_kernel(1)
        a = ctx.ivf_search(queries, k, nprobe, max_scan, filter_bitmap=filt)
        assert ctx.last_kernel() == "scan_lists"
        ctx.set_kernel(0)
        b = ctx.ivf_search(queries, k, nprobe, max_scan, filter_bitmap=filt)
        assert ctx.last_kernel() == "scan_lists_lm"
        assert np.array_equal(a[3], b[3]) and np.array_equal(a[2], b[2]) and np.array_equal(a[0], b[0])
        assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))
        assert a[3].max() > 0
    ctx.close(); st.close()


@pytest.mark.parametrize("metric", [z.METRIC_L2SQ, z.METRIC_IP])
@pytest.mark.parametrize("per_list", [False, True])
def test_typed_centroids_split_queries_and_score_scales(metric, per_list):
    """zvb200_ivf_attach_typed + zvb200_ivf_search_split + zvb200_store_set_score_scale (the builder-quantised IVF
    index: int8 posting lists under fp32 centroids, IVFEntity::search's distances[k] * norm_val) against numpy: the
    int8 arithmetic is exact, the product with the scale is one fp32 multiplication, so the scores are bit for bit
    — in both posting-list scans, in the flat scan of the same store, with a threshold on the SCALED scores."""
    n, d, cd, nlist, q, k, nprobe, max_scan = 12000, 40, 24, 30, 24, 10, 4, 1500
    rng = np.random.default_rng(77)
    rows = rng.integers(-128, 128, (n, d), dtype=np.int8)
    cuts = np.sort(rng.choice(np.arange(1, n), nlist - 1, replace=False))
    offs = np.concatenate([[0], cuts, [n]]).astype(np.uint64)
    cent = rng.standard_normal((nlist, cd)).astype(np.float32)
    cq = rng.standard_normal((q, cd)).astype(np.float32)
    lq = rng.integers(-128, 128, (q, d), dtype=np.int8)
    scales = (0.5 + rng.random(nlist)).astype(np.float32) if per_list else None
    gscale = np.float32(0.37)
    st = z.Store(z.DT_INT8, metric, d)
    st.append(rows.view(np.uint8)); st.seal()
    st.ivf_attach_typed(cent.view(np.uint8), offs, z.DT_FP32, z.METRIC_L2SQ, cd)
    with pytest.raises(z.ZvecB200Error) as ei:             # one query layout cannot serve both stages
        z.Context(st).ivf_search(lq.view(np.uint8), k, nprobe, max_scan)
    assert ei.value.code == -24
    st.set_score_scale(1.0 if per_list else float(gscale), scales)
    list_of = np.searchsorted(offs[1:].astype(np.int64), np.arange(n), side="right")
    row_scale = scales[list_of] if per_list else np.full(n, gscale, np.float32)
    r32, q32 = rows.astype(np.int64), lq.astype(np.int64)
    if metric == z.METRIC_L2SQ:
        dist = ((r32 ** 2).sum(1)[None, :] - 2 * (q32 @ r32.T) + (q32 ** 2).sum(1)[:, None]).astype(np.float32)
    else:
        dist = (-(q32 @ r32.T)).astype(np.float32)
    scaled = dist * row_scale[None, :]                      # fp32 x fp32, rounded once
    cd2 = ((cq[:, None, :].astype(np.float64) - cent[None, :, :]) ** 2).sum(2)
    thr = float(np.median(scaled))

    def expect(i, threshold, whole):
        if whole:
            cand = np.arange(n)
        else:
            cand, seen = [], 0
            for l in np.argsort(cd2[i], kind="stable")[:nprobe]:
                if seen >= max_scan:
                    break
                cand.append(np.arange(int(offs[l]), int(offs[l + 1]))); seen += int(offs[l + 1] - offs[l])
            cand = np.concatenate(cand)
        cand = cand[scaled[i][cand] <= np.float32(threshold)]
        order = np.lexsort((cand, scaled[i][cand]))[:k]           # (score, row): the packed keys' order
        return cand[order], scaled[i][cand[order]]

    ctx = z.Context(st)
    for threshold in (FLT_MAX, thr):
        for kernel, batch, name in ((1, q, "scan_lists"), (0, q, "scan_lists_lm"), (0, 1, "scan_lists")):
            ctx.set_kernel(kernel)
            for b0 in range(0, q, batch):
                gk, gs, gi, gc = ctx.ivf_search_split(cq[b0:b0 + batch], lq[b0:b0 + batch].view(np.uint8), k, nprobe,
                                                      max_scan, threshold=threshold)
                assert ctx.last_kernel() == name
                for j in range(gk.shape[0]):
                    ei_, es_ = expect(b0 + j, threshold, False)
                    assert gc[j] == len(ei_)
                    assert np.array_equal(gi[j][:len(ei_)], ei_), (name, threshold, b0 + j)
                    assert np.array_equal(gs[j][:len(ei_)].view(np.uint32), es_.view(np.uint32))
        ctx.set_kernel(0)
        gk, gs, gi, gc = ctx.search(lq.view(np.uint8), k, threshold=threshold)   # search_bf_impl: every list
        assert ctx.last_kernel() == "scan_gemv"                                 # a scaled store stays on the GEMV scan
        for j in range(q):
            ei_, es_ = expect(j, threshold, True)
            assert gc[j] == len(ei_) and np.array_equal(gi[j][:len(ei_)], ei_)
            assert np.array_equal(gs[j][:len(ei_)].view(np.uint32), es_.view(np.uint32))
    if per_list:                                           # list scales cover the attached rows only: no append
        with pytest.raises(z.ZvecB200Error) as ei:
            st.append(rows[:1].view(np.uint8))
        assert ei.value.code == -12
    st.ivf_attach_typed(cent.view(np.uint8), offs, z.DT_FP32, z.METRIC_L2SQ, cd)  # re-attaching drops the list scales
    st.set_score_scale(1.0, None)
    gk, gs, gi, gc = ctx.search(lq[:2].view(np.uint8), k)
    assert np.array_equal(gs[0].view(np.uint32), np.sort(dist[0])[:k].view(np.uint32))
    ctx.close(); st.close()
