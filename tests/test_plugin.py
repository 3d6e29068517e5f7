"""The drop-in boundary, exercised the way a zvec maintainer would: the plugin .so
(integration/gpu_flat_plugin.cc) is loaded by the reference's OWN IndexPluginBroker, the index
is built by the reference's FlatBuilder, and IndexFactory::CreateSearcher("GpuFlatSearcher")
searches it.  Invariants follow tests/core/algorithm/flat/flat_searcher_test.cpp:93-1272 of the
reference: row-major index == column-major index == batch == non-batch == pass-all filter,
same key(), same index(), scores within tolerance of FlatSearcher's."""
import numpy as np
import pytest

import support as S

needs_plugin = pytest.mark.skipif(S.ref() is None or not S.load_plugin(),
                                  reason="oracle/_ref or integration/_build not built here")


@needs_plugin
def test_plugin_registers_with_the_reference_factory():
    assert S.ref().zref_has_searcher(b"GpuFlatSearcher") == 1
    assert S.ref().zref_has_searcher(b"FlatSearcher") == 1          # stock class still there


@needs_plugin
def test_plugin_fails_loudly_without_a_device():
    import zvec_b200 as z
    if z.lib().zvb200_device_count() > 0:
        pytest.skip("a device is present")
    x = np.random.default_rng(0).random((64, 8), dtype=np.float32)
    with pytest.raises(RuntimeError, match="searcher load"):
        S.RefIndex(x, S.DT_FP32, 8, "SquaredEuclidean", "", kind="Flat:GpuFlatSearcher")


CASES = [("", "SquaredEuclidean"), ("", "InnerProduct"), ("HalfFloatConverter", "InnerProduct"),
         ("HalfFloatConverter", "SquaredEuclidean"), ("CosineNormalizeConverter", "Cosine"),
         ("Int8StreamingConverter", "SquaredEuclidean"), ("Int8StreamingConverter", "InnerProduct")]


def _tie_free(n, d, seed):
    # the reference's own trick for tie-free data (flat_searcher_test.cpp:106-111) would make L2
    # trivially ordered; a continuous distribution is tie-free with probability 1
    rng = np.random.default_rng(seed)
    return rng.uniform(-1, 1, (n, d)).astype(np.float32)


@needs_plugin
@pytest.mark.gpu
@pytest.mark.parametrize("conv,metric", CASES)
@pytest.mark.parametrize("column", [False, True])
def test_gpu_searcher_equals_flat_searcher(conv, metric, column):
    if column and metric == "Cosine":
        # CosineMetric has no MxN kernels (cosine_metric.cc:194-199): the reference's own FlatSearcher
        # cannot serve a column-major cosine index (it dereferences a null distance function)
        pytest.skip("the reference cannot build/search column-major Cosine")
    n, d, q, k = 10007, 64, 40, 10           # 10007 = 312 column tiles + a 23-row row-major tail
    x, qs = _tie_free(n, d, 5), _tie_free(q, d, 6)
    ref = S.RefIndex(x, S.DT_FP32, d, metric, conv, kind="Flat", column_major=column)
    gpu = S.RefIndex(x, S.DT_FP32, d, metric, conv, kind="Flat:GpuFlatSearcher", column_major=column)
    rk, rs, ri, rc = ref.search(qs, k, batch=1)
    exact = ref.dtype == S.DT_INT8
    stored, tq = S.RefIndex(x, S.DT_FP32, d, metric, conv).stored(), ref.transform_queries(qs)
    import zvec_b200 as z
    origin = metric if conv.startswith("Int8") else None
    mid = z.metric_from_names(ref.metric_name, origin)
    for batch in (1, 7, 40):                 # non-batch and batch calls of search_impl
        gk, gs, gi, gc = gpu.search(qs, k, batch=batch)
        assert np.array_equal(gc, rc)
        for i in range(q):
            alld = S.oracle_distances(mid, ref.dtype, stored, ref.dim, tq[i])
            S.check_topk_parity(gi[i], gs[i], ri[i], rs[i], all_dist=alld, exact=exact,
                                what="plugin %s/%s col=%s batch=%d q%d" % (conv, metric, column, batch, i))
            same = gi[i] == ri[i]
            assert np.array_equal(gk[i][same], rk[i][same])       # key() follows index()
    ref.close(); gpu.close()


@needs_plugin
@pytest.mark.gpu
def test_gpu_searcher_filter_threshold_and_normalized_scores():
    n, d, q, k = 5000, 32, 9, 20
    x, qs = _tie_free(n, d, 7), _tie_free(q, d, 8)
    ref = S.RefIndex(x, S.DT_FP32, d, "InnerProduct", "HalfFloatConverter", kind="Flat")
    gpu = S.RefIndex(x, S.DT_FP32, d, "InnerProduct", "HalfFloatConverter", kind="Flat:GpuFlatSearcher")
    L = S.ref()
    try:
        L.zref_set_filter(3, 1)              # IndexFilter: drop keys with key % 3 == 1
        rk, rs, ri, rc = ref.search(qs, k, normalize=True)
        gk, gs, gi, gc = gpu.search(qs, k, normalize=True)
        assert np.array_equal(rc, gc) and (gk[gk != np.uint64(2**64 - 1)] % 3 != 1).all()
        assert np.array_equal(rk, gk) and np.allclose(rs, gs, rtol=1e-4, atol=1e-6)
        assert (np.diff(gs, axis=1) <= 1e-6).all()            # normalize() flips IP back: descending
        L.zref_set_filter(1, 0)              # everything filtered
        _, _, _, gc = gpu.search(qs, k)
        assert (gc == 0).all()
        L.zref_set_filter(0, 0)
        L.zref_set_threshold(1, -6.0)        # radius in internal units (-IP <= -6): FlatSearcherContext has no metric to denormalise with
        rk, rs, ri, rc = ref.search(qs, k)
        gk, gs, gi, gc = gpu.search(qs, k)
        assert np.array_equal(rc, gc) and np.array_equal(rk, gk) and 0 < rc.max() < k
    finally:
        L.zref_set_filter(0, 0); L.zref_set_threshold(0, 0.0)
    ref.close(); gpu.close()
