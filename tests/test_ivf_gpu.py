"""IVF: the GPU coarse + posting-list scan against the reference's own IVFSearcher
(oracle/_ref) on an index the reference's IVFBuilder trained and packed; the device store ingests
the posting lists exactly as the reference's IVFEntity reads them (RefIndex.ivf_export)."""
import math

import numpy as np
import pytest

import support as S
import zvec_b200 as z

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
