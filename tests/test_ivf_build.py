"""IVF build on the device, second half: nearest-centroid labels (zvb200_ivf_assign), the dump order
(zvb200_ivf_pack) and GpuIVFBuilder (integration/gpu_ivf_builder.cc), which writes the reference's on-disk IVF format
through the reference's own IVFDumper.  The bar: from the SAME centroids the GPU-built file and the file the stock
IVFBuilder writes (one thread) hold the same centroids, list offsets, rows and keys; the stock IVFSearcher loads the
GPU-built file; GpuIVFSearcher over it answers like the stock searcher."""
import ctypes as C
import math

import numpy as np
import pytest

import support as S
import zvec_b200 as z
from test_ivf_gpu import clustered, budget

needs_ref = pytest.mark.skipif(S.ref() is None, reason="oracle/_ref not available")
needs_plugin = pytest.mark.skipif(S.ref() is None or not S.load_plugin(), reason="oracle/_ref or the plugin not built")


def _fixed(cents):
    L = S.ref()
    L.zref_set_fixed_centroids.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32]
    if cents is None:
        L.zref_set_fixed_centroids(None, 0, 0)
    else:
        cents = np.ascontiguousarray(cents)
        L.zref_set_fixed_centroids(cents.ctypes.data, cents.shape[0], cents.nbytes // cents.shape[0])


def _nearest(x, cents, metric="l2"):
    xd, cd = x.astype(np.float64), cents.astype(np.float64)
    if metric == "l2":
        d = (xd * xd).sum(1)[:, None] - 2.0 * xd @ cd.T + (cd * cd).sum(1)[None, :]
    else:
        d = 1.0 - xd[:, :-1] @ cd[:, :-1].T
    part = np.partition(d, 1, axis=1)
    clear = (part[:, 1] - part[:, 0]) > 1e-5 * np.maximum(np.abs(part[:, 1]), 1e-6)
    return d.argmin(1).astype(np.uint32), clear


@needs_ref
def test_stock_builder_from_fixed_centroids_labels_by_the_nearest_centroid():
    """pins the harness: IVFBuilder::train(trainer) + build with one thread = nearest centroid, ids ascending per list"""
    n, d, nlist = 6000, 24, 37
    x, _ = clustered(n, d, nlist, 5)
    cents = x[np.arange(nlist) * (n // nlist)].copy()
    _fixed(cents)
    try:
        ri = S.RefIndex(x, S.DT_FP32, d, "SquaredEuclidean", "", kind="IVF",
                        build_params="proxima.ivf.builder.centroid_count=%d;proxima.ivf.builder.thread_count=1" % nlist)
    finally:
        _fixed(None)
    e = ri.ivf_export()
    assert e["nlist"] == nlist and np.array_equal(e["centroids"].view(np.float32).reshape(nlist, d), cents)
    lab, clear = _nearest(x, cents)
    offs = e["list_offsets"].astype(np.int64)
    for l in range(nlist):
        ks = e["keys"][offs[l]:offs[l + 1]].astype(np.int64)
        assert np.array_equal(ks, np.sort(ks))                                   # ascending ids inside a list
        assert (lab[ks][clear[ks]] == l).all()
    ri.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n,nlist", [(1, 1), (5000, 7), (300001, 16384), (70000, 3)])
def test_ivf_pack_is_the_stable_order_by_list(n, nlist):
    rng = np.random.default_rng(n + nlist)
    labels = rng.integers(0, nlist, n).astype(np.uint32)
    if nlist > 4:
        labels[labels == 3] = 4                                                   # an empty list
    order, offs = z.ivf_pack(labels, nlist)
    exp = np.argsort(labels, kind="stable").astype(np.uint32)
    assert np.array_equal(order, exp)
    cnt = np.bincount(labels, minlength=nlist)
    assert np.array_equal(offs, np.concatenate([[0], np.cumsum(cnt)]).astype(np.uint64))
    with pytest.raises(z.ZvecB200Error):
        z.ivf_pack(np.array([0, nlist], np.uint32), nlist)                        # a label out of range


@pytest.mark.gpu
@pytest.mark.parametrize("metric", ["l2", "cos"])
def test_ivf_assign_is_the_nearest_centroid(metric):
    n, d, nlist = 50000, 40, 300
    x, _ = clustered(n, d, nlist, 8)
    if metric == "cos":
        x = S.cosine_normalize_fp32(x)
    cents = x[np.arange(nlist) * (n // nlist)].copy()
    lab = z.ivf_assign(x, cents, z.DT_FP32, z.METRIC_L2SQ if metric == "l2" else z.METRIC_COSINE, x.shape[1])
    exp, clear = _nearest(x, cents, metric)
    assert clear.mean() > 0.99 and np.array_equal(lab[clear], exp[clear])
    assert lab.max() < nlist


@pytest.mark.gpu
@needs_plugin
@pytest.mark.parametrize("metric,conv,d", [("SquaredEuclidean", "", 48), ("Cosine", "CosineNormalizeConverter", 48),
                                           ("SquaredEuclidean", "", 600)])      # 600 dims: row-major posting blocks
def test_gpu_built_file_equals_the_stock_builders_file_from_the_same_centroids(metric, conv, d):
    n, nlist, k = 30000, 100, 10
    x, qs = clustered(n, d, nlist, 3)
    stored = S.cosine_normalize_fp32(x) if conv else x
    cents = stored[np.arange(nlist) * (n // nlist)].copy()
    bp = "proxima.ivf.builder.centroid_count=%d;proxima.ivf.builder.thread_count=1" % nlist
    sp = "proxima.ivf.searcher.scan_ratio=0.1"
    _fixed(cents)
    try:
        stock = S.RefIndex(x, S.DT_FP32, d, metric, conv, kind="IVF", build_params=bp, search_params=sp)
        l0 = z.lib().zvb200_kernel_launches()
        gpu = S.RefIndex(x, S.DT_FP32, d, metric, conv, kind="GpuIVF:IVFSearcher", build_params=bp, search_params=sp)
        assert z.lib().zvb200_kernel_launches() > l0, "GpuIVFBuilder did not reach the device"
        both = S.RefIndex(x, S.DT_FP32, d, metric, conv, kind="GpuIVF:GpuIVFSearcher", build_params=bp, search_params=sp)
    finally:
        _fixed(None)
    a, b = stock.ivf_export(), gpu.ivf_export()
    for key in ("nlist", "c_dtype", "c_dim", "c_metric"):
        assert a[key] == b[key], key
    assert np.array_equal(a["centroids"], b["centroids"])
    # the labels agree except where a row sits (numerically) between two centroids
    lab, clear = _nearest(stored, cents, "cos" if conv else "l2")
    if np.array_equal(a["list_offsets"], b["list_offsets"]):
        assert np.array_equal(a["keys"], b["keys"]) and np.array_equal(a["rows"], b["rows"])   # the same file content
    else:
        assert (~clear).any()
    offs = b["list_offsets"].astype(np.int64)
    for l in range(nlist):
        ks = b["keys"][offs[l]:offs[l + 1]].astype(np.int64)
        assert np.array_equal(ks, np.sort(ks)) and (lab[ks][clear[ks]] == l).all()
    # the stock searcher over the GPU-built file = over the stock-built file; GpuIVFSearcher over it too
    rk, rs, _, rc = stock.search(qs, k)
    gk, gs, _, gc = gpu.search(qs, k)
    bk, bs, _, bc = both.search(qs, k)
    if np.array_equal(a["list_offsets"], b["list_offsets"]):
        assert np.array_equal(rk, gk) and np.array_equal(rs.view(np.uint32), gs.view(np.uint32)) and np.array_equal(rc, gc)
    assert np.array_equal(gc, bc)
    same = gk == bk
    assert same.mean() > 0.995 and np.allclose(gs[same], bs[same], rtol=1e-4, atol=1e-6)
    for ix in (stock, gpu, both):
        ix.close()


@pytest.mark.gpu
@needs_plugin
def test_gpu_trained_index_is_loaded_and_searched_by_the_stock_searcher():
    """the whole build on the device: k-means, labels, packing — then the reference's IVFSearcher on the file"""
    n, d, nlist, k = 40000, 32, 64, 10
    x, qs = clustered(n, d, nlist, 9)
    bp = "proxima.ivf.builder.centroid_count=%d;proxima.ivf.builder.train_sample_count=20000" % nlist
    gpu = S.RefIndex(x, S.DT_FP32, d, "SquaredEuclidean", "", kind="GpuIVF:IVFSearcher", build_params=bp,
                     search_params="proxima.ivf.searcher.scan_ratio=0.2")
    e = gpu.ivf_export()
    assert 1 <= e["nlist"] <= nlist and int(e["list_offsets"][-1]) == n
    assert np.array_equal(np.sort(e["keys"]), np.arange(n, dtype=np.uint64))
    cents = e["centroids"].view(np.float32).reshape(e["nlist"], d)
    lab, clear = _nearest(x, cents)
    offs = e["list_offsets"].astype(np.int64)
    for l in range(e["nlist"]):
        ks = e["keys"][offs[l]:offs[l + 1]].astype(np.int64)
        assert (lab[ks][clear[ks]] == l).all()
    # the clusters are the data's: searching 20 % of the lists finds the true neighbours
    gk, gs, _, gc = gpu.search(qs, k)
    dd = ((qs[:, None, :].astype(np.float64) - x[None, :, :]) ** 2).sum(-1)
    true = np.argsort(dd, axis=1)[:, :k]
    recall = np.mean([len(set(gk[i].tolist()) & set(true[i].tolist())) / k for i in range(len(qs))])
    assert recall > 0.95, recall
    gpu.close()
