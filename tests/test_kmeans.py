"""k-means on the device against the reference's KmeansCluster (compiled from its own sources,
oracle/_ref) and the numpy restatement oracle/kmeans_oracle.py.  Sums are accumulated in double on all three
sides, so the centroids agree to fp32 rounding as long as the assignments agree; rows whose two nearest
centroids are closer than 1e-4 relative are near-ties and may be assigned either way."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import support as S

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import kmeans_oracle as KO  # noqa: E402


def _data(n, d, k, seed, spread=0.5):
    rng = np.random.default_rng(seed)
    centres = rng.standard_normal((k, d)).astype(np.float32)
    x = (centres[rng.integers(0, k, n)] + spread * rng.standard_normal((n, d))).astype(np.float32)
    init = x[rng.choice(n, k, replace=False)].copy()
    return x, init


def _ref_kmeans(metric_name, x, init, iters, threads=4):
    L = S.ref()
    L.zref_kmeans.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32,
                              C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
    n, d = x.shape
    k = len(init)
    cent = init.copy()
    lab = np.zeros(n, np.uint32); fol = np.zeros(k, np.uint64); sse = C.c_double()
    rc = L.zref_kmeans(metric_name.encode(), x.ctypes.data, n, d, cent.ctypes.data, k, iters, threads,
                       lab.ctypes.data, fol.ctypes.data, C.byref(sse))
    assert rc == 0, L.zref_last_error()
    return cent, lab, fol, sse.value


def _near_tie_rows(d, rtol=1e-4):
    part = np.partition(d, 1, axis=1)
    return (part[:, 1] - part[:, 0]) <= rtol * np.maximum(np.abs(part[:, 1]), 1e-6)


needs_ref = pytest.mark.skipif(S.ref() is None, reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("iters", [1, 4])
def test_oracle_kmeans_equals_the_reference(iters):
    x, init = _data(6000, 24, 16, 3)
    rc, rl, rf, rs = _ref_kmeans("SquaredEuclidean", x, init, iters)
    oc, ol, of, os_, d = KO.kmeans(x, init, iters)
    ties = _near_tie_rows(d)
    assert np.array_equal(rl[~ties], ol[~ties])
    assert np.allclose(rc, oc, rtol=1e-5, atol=1e-6, equal_nan=True)
    assert np.array_equal(rf, of) or ties.any()
    assert abs(rs - os_) <= 1e-5 * abs(rs)


def test_oracle_kmeans_empty_cluster_is_nan():
    x, init = _data(500, 8, 4, 4)
    init[3] = 1e6                                   # nobody's nearest centroid
    c, cnt, cost, lab, _ = KO.lloyd_pass(x, init)
    assert cnt[3] == 0 and np.isnan(c[3]).all() and not np.isnan(c[:3]).any()
    c2, cnt2, _, _, _ = KO.lloyd_pass(x, c)         # a NaN centroid is never chosen again
    assert cnt2[3] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("n,d,k,iters", [(20000, 32, 64, 1), (20000, 32, 64, 5), (5000, 100, 7, 3), (70000, 16, 300, 2)])
def test_gpu_kmeans_equals_oracle_and_reference(n, d, k, iters):
    import zvec_b200 as z
    x, init = _data(n, d, k, 10 + k)
    gc, gl, gf, gs, git = z.kmeans(x, init, max_iterations=iters, epsilon=0.0)
    assert git == iters
    oc, ol, of, os_, dmat = KO.kmeans(x, init, iters)
    ties = _near_tie_rows(dmat)
    assert ties.mean() < 0.01
    assert np.array_equal(gl[~ties], ol[~ties])
    # identical assignments in every pass <=> identical follows; then the centroids agree to fp32 rounding
    if np.array_equal(gf, of):
        assert np.allclose(gc, oc, rtol=1e-5, atol=1e-6, equal_nan=True)
    else:
        assert np.abs(gf.astype(np.int64) - of.astype(np.int64)).sum() <= 4 * max(1, int(ties.sum()))
        assert np.allclose(gc, oc, rtol=1e-2, atol=1e-3, equal_nan=True)
    assert abs(gs - os_) <= 1e-4 * abs(os_)
    if S.ref() is not None:
        rc, rl, rf, rs = _ref_kmeans("SquaredEuclidean", x, init, iters)
        assert np.array_equal(gl[~ties], rl[~ties])
        if np.array_equal(gf, rf):
            assert np.allclose(gc, rc, rtol=1e-5, atol=1e-6, equal_nan=True)
        assert abs(gs - rs) <= 1e-4 * abs(rs)


@pytest.mark.gpu
def test_gpu_kmeans_empty_cluster_nan_and_epsilon_stop():
    import zvec_b200 as z
    x, init = _data(4000, 8, 4, 21)
    init[3] = 1e6
    gc, gl, gf, gs, git = z.kmeans(x, init, max_iterations=2, epsilon=0.0)
    assert gf[3] == 0 and np.isnan(gc[3]).all() and not np.isnan(gc[:3]).any()
    assert (gl != 3).all()
    oc, ol, of, os_, _ = KO.kmeans(x, init, 2)
    assert np.array_equal(gf, of) and np.allclose(gc, oc, rtol=1e-5, atol=1e-6, equal_nan=True)
    # a huge epsilon stops after the second pass (the first compares with cost 0, kmeans_cluster.cc:510-529)
    _, _, _, _, it2 = z.kmeans(x, init, max_iterations=20, epsilon=1e30)
    assert it2 == 1


@pytest.mark.gpu
def test_gpu_kmeans_feeds_the_gpu_ivf_index():
    """train on the device, group the rows by label, attach: the IVF search then finds each row in its own list"""
    import zvec_b200 as z
    n, d, k = 30000, 32, 64
    x, init = _data(n, d, k, 33, spread=0.3)
    cent, lab, fol, _, _ = z.kmeans(x, init, max_iterations=8, epsilon=0.0)
    fol = np.bincount(lab, minlength=k)              # list sizes under label() (follows is the last pass's)
    keep = fol > 0                                   # IVFBuilder drops empty clusters (purge)
    remap = np.cumsum(keep) - 1
    order = np.argsort(lab, kind="stable")
    offs = np.zeros(int(keep.sum()) + 1, np.uint64)
    offs[1:] = np.cumsum(fol[keep])
    st = z.Store(z.DT_FP32, z.METRIC_L2SQ, d)
    st.append(x[order], order.astype(np.uint64)); st.seal()
    st.ivf_attach(cent[keep], offs)
    ctx = z.Context(st)
    qs = x[:200]
    keys, scores, index, counts = ctx.ivf_search(qs, 1, 1, n)        # 1 probe: the row's own list
    assert (counts == 1).all()
    assert (keys[:, 0] == np.arange(200)).mean() > 0.99 and (scores[:, 0] <= 1e-6).mean() > 0.99
    assert np.array_equal(remap[lab[:200]], np.searchsorted(offs, index[:, 0].astype(np.uint64), side="right") - 1)
    ctx.close(); st.close()


# ---- golden vectors written by the reference's KmeansCluster (tests/golden/make_kmeans_golden.py) ----
def _golden_cases():
    import json
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kmeans.json")))
    for c in g["cases"]:
        n, d, k = c["n"], c["dim"], c["k"]
        x = (np.array(c["rows_times_8"], np.float32) / 8).reshape(n, d)
        init = (np.array(c["init_times_8"], np.float32) / 8).reshape(k, d)
        ref = np.array(c["centroids_bits"], np.uint32).view(np.float32).reshape(k, d)
        yield c, x, init, ref


def _same_centroids(got, ref):
    assert np.array_equal(np.isnan(got), np.isnan(ref))
    live = ~np.isnan(ref)
    assert np.array_equal(got.view(np.uint32)[live], ref.view(np.uint32)[live])     # sums are exact in double: bit for bit


def test_oracle_kmeans_reproduces_the_golden_vectors():
    for c, x, init, ref in _golden_cases():
        oc, ol, of, os_, _ = KO.kmeans(x, init, c["iterations"])
        _same_centroids(oc, ref)
        assert np.array_equal(ol, np.array(c["labels"], np.uint32)) and of.tolist() == c["follows"]
        assert abs(os_ - c["sse"]) <= 1e-5 * abs(c["sse"])
@pytest.mark.gpu
def test_gpu_kmeans_reproduces_the_golden_vectors():
    import zvec_b200 as z
    for c, x, init, ref in _golden_cases():
        gc, gl, gf, gs, git = z.kmeans(x, init, max_iterations=c["iterations"], epsilon=0.0)
        _same_centroids(gc, ref)
        assert np.array_equal(gl, np.array(c["labels"], np.uint32)) and gf.astype(int).tolist() == c["follows"]
        assert abs(gs - c["sse"]) <= 1e-5 * abs(c["sse"])


# ---- edge cases of KmeansCluster the device path must share -------------------------------------
@needs_ref
@pytest.mark.parametrize("metric,name", [("ip", "InnerProduct"), ("l2", "Euclidean")])
def test_oracle_kmeans_other_metrics_equal_the_reference(metric, name):
    """InnerProduct and Euclidean assignment passes pinned to the compiled reference (only SquaredEuclidean was)"""
    x, init = _data(5000, 20, 12, 9)
    rc, rl, rf, rs = _ref_kmeans(name, x, init, 2)
    c = init.copy()
    for _ in range(2):
        dm = KO.distances(x, c, "ip" if metric == "ip" else "l2sq")
        if metric == "l2":
            dm = np.sqrt(dm)
        lab = np.where(np.isnan(dm), np.inf, dm).argmin(1)
        sums = np.zeros((len(c), x.shape[1])); np.add.at(sums, lab, x.astype(np.float64))
        cnt = np.bincount(lab, minlength=len(c))
        with np.errstate(invalid="ignore", divide="ignore"):
            c = (sums / cnt[:, None]).astype(np.float32)
    ties = _near_tie_rows(dm)
    if not ties.any():
        assert np.allclose(rc, c, rtol=1e-5, atol=1e-6, equal_nan=True)
        assert np.array_equal(rf, cnt.astype(np.uint64))


@needs_ref
def test_oracle_kmeans_moves_empty_clusters_to_the_tail_like_the_reference():
    """InnerProduct k-means starves clusters; cluster() ends with PurgeCentroids (kmeans_cluster.cc:363-383, 533), which
    swaps the empty ones to the tail even with purge_empty = false — centroids, follows and labels come back in THAT
    order.  Zero passes rotate the list (nobody follows anything)."""
    x, init = _data(20000, 32, 40, 10)
    rc, rl, rf, rs = _ref_kmeans("InnerProduct", x, init, 3, threads=8)
    assert (rf == 0).sum() >= 2 and not (rf[:-3] == 0).any()          # the case tests something
    oc, ol, of, os_, d = KO.kmeans(x, init, 3, "ip")
    assert np.array_equal(of, rf) and np.allclose(oc, rc, rtol=1e-5, atol=1e-6, equal_nan=True)
    ties = _near_tie_rows(np.where(np.isnan(d), np.inf, d))
    assert np.array_equal(ol[~ties], rl[~ties])
    y, init2 = _data(300, 8, 5, 22)
    rc0, _, rf0, _ = _ref_kmeans("SquaredEuclidean", y, init2, 0)
    oc0, _, of0, _, _ = KO.kmeans(y, init2, 0)
    assert np.array_equal(rc0, oc0) and not rf0.any() and not of0.any()


@needs_ref
def test_reference_kmeans_puts_an_all_nan_row_into_cluster_0():
    """LinearSeeker::seek starts at (column 0, FLT_MAX) and only moves on a strictly smaller score
    (linear_seeker.cc:25-41): a row whose distances are all NaN follows centroid 0 and adds FLT_MAX to the cost."""
    x, init = _data(400, 8, 4, 21)
    x[7] = np.nan
    rc, rl, rf, rs = _ref_kmeans("SquaredEuclidean", x, init, 1)
    assert rl[7] == 0 or np.isnan(rc[0]).any()       # cluster 0 took the NaN row: its mean is NaN
    assert np.isnan(rc[0]).all() and rs >= np.finfo(np.float32).max * 0.99


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("metric,name", [(1, "Euclidean"), (2, "InnerProduct")])
def test_gpu_kmeans_other_metrics_equal_the_reference(metric, name):
    import zvec_b200 as z
    x, init = _data(20000, 32, 40, 10)
    rc, rl, rf, rs = _ref_kmeans(name, x, init, 3, threads=8)
    gc, gl, gf, gs, it = z.kmeans(x, init, metric=metric, max_iterations=3, epsilon=0.0)
    assert it == 3
    if np.array_equal(gf, rf):
        assert np.allclose(gc, rc, rtol=1e-5, atol=1e-6, equal_nan=True)
        assert abs(gs - rs) <= 1e-4 * abs(rs)
    assert (gl == rl).mean() > 0.999


@pytest.mark.gpu
@needs_ref
def test_gpu_kmeans_nan_row_and_zero_iterations_follow_the_reference():
    import zvec_b200 as z
    x, init = _data(400, 8, 4, 21)
    x[7] = np.nan
    rc, rl, rf, rs = _ref_kmeans("SquaredEuclidean", x, init, 1)
    gc, gl, gf, gs, it = z.kmeans(x, init, max_iterations=1, epsilon=0.0)
    assert np.array_equal(gf, rf)                     # the NaN row was counted: cluster 0 has the reference's follows
    assert np.isnan(gc[0]).all() and np.allclose(gc[1:], rc[1:], rtol=1e-5, atol=1e-6, equal_nan=True)
    assert gs >= np.finfo(np.float32).max * 0.99
    assert gl.max() < 4                               # labels never carry 0xFFFFFFFF
    # max_iterations = 0: the loop of KmeansCluster::cluster does not run — nothing follows, and PurgeCentroids
    # (kmeans_cluster.cc:363-383) still swaps every 'empty' centroid to the tail: the list comes back rotated
    y, init2 = _data(300, 8, 5, 22)
    rc0, rl0, rf0, rs0 = _ref_kmeans("SquaredEuclidean", y, init2, 0)
    gc, gl, gf, gs, it = z.kmeans(y, init2, max_iterations=0)
    assert it == 0 and np.array_equal(gc, rc0) and gs == 0.0 and not gf.any()
    assert np.array_equal(gc, init2[KO.purge_order(np.zeros(5))])
    ol, _, _ = KO.assign(y, gc)
    assert (gl == ol).mean() > 0.99
