"""Scenarios that drive core_interface::Index (the layer collection.query() calls, vector_column_indexer.cc:195)
through oracle/ref_shim.cc.  Run twice by tests/test_core_interface.py: with the stock streamer classes, and — in a
process of its own — after zvb200_install_overrides(), when the SAME calls reach the GPU classes by name."""
import numpy as np

import support as S

N, D, Q, K = 3000, 32, 20, 10


def data(seed=11):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (N, D)).astype(np.float32)          # continuous: tie-free with probability 1
    qs = rng.uniform(-1, 1, (Q, D)).astype(np.float32)
    return x, qs


def in_thread(fn, *args, **kw):
    """core_interface keeps ONE search context per (thread, index type) (index.cc:24-45) and _prepare_for_search only
    ever SETS a filter / radius on it (flat_index.cc:60-65): what one search set stays for the next searches of that
    thread, whatever index they go to.  Every suite therefore runs on a thread of its own (a fresh context), and orders
    its searches so that nothing depends on a leftover."""
    import threading
    box = {}

    def run():
        try:
            box["out"] = fn(*args, **kw)
        except BaseException as e:  # noqa: BLE001 - re-raised on the caller's thread
            box["err"] = e
    t = threading.Thread(target=run)
    t.start(); t.join()
    if "err" in box:
        raise box["err"]
    return box["out"]


def flat_suite(tmp, metric, quantizer, batch=False):
    return in_thread(_flat_suite, tmp, metric, quantizer, batch)


def refiner_suite(tmp, batch=False):
    return in_thread(_refiner_suite, tmp, batch)


def _flat_suite(tmp, metric, quantizer, batch=False):
    """-> dict name -> (keys, scores, counts[, vectors]) for one Flat index"""
    x, qs = data()
    out = {}
    ix = S.CiIndex("Flat", D, metric, quantizer, path="%s/ci_%s_%s.index" % (tmp, metric, quantizer or "none"))
    ix.add(x[:2000], 0)
    out["first"] = ix.search(qs, K, batch=batch)               # rows arrive between searches
    ix.add(x[2000:], 2000)
    assert ix.count() == N
    out["plain"] = ix.search(qs, K, batch=batch)
    out["linear"] = ix.search(qs, K, batch=batch, linear=True)
    pks = np.arange(5, N, 13, dtype=np.uint64)
    pks = np.concatenate([pks, np.array([N + 7, 2 ** 40], np.uint64)])     # keys the index does not hold
    out["bf_pks"] = ix.search(qs, K, batch=batch, bf_pks=pks)
    out["fetch_vector"] = ix.search(qs[:4], 5, batch=batch, fetch_vector=True)
    out["fetch"] = (np.stack([ix.fetch(i) for i in (0, 17, N - 1)]),)
    # from here on every search names its own filter (a filter, once set, stays on the thread's context)
    try:
        S.ref().zref_set_filter(3, 1)
        out["filter"] = ix.search(qs, K, batch=batch)
        S.ref().zref_set_filter(2, 0)
        out["bf_pks_filter"] = ix.search(qs, K, batch=batch, bf_pks=pks)
        if metric == "L2":          # last: a radius, once set, stays too
            S.ref().zref_set_filter(5, 0)
            r = float(np.median(out["plain"][1][:, K // 2]))
            out["radius"] = ix.search(qs, K, batch=batch, radius=r)
    finally:
        S.ref().zref_set_filter(0, 0)
    ix.close()
    return out


def _refiner_suite(tmp, batch=False):
    """int8-quantised index refined by an exact Flat index (Index::Search with refiner_param, index.cc:408-445)"""
    x, qs = data(12)
    coarse = S.CiIndex("Flat", D, "L2", "int8", path=tmp + "/ci_ref_coarse.index")
    exact = S.CiIndex("Flat", D, "L2", "", path=tmp + "/ci_ref_exact.index")
    coarse.add(x, 0); exact.add(x, 0)
    out = {"refined": coarse.search(qs, K, batch=batch, refiner=exact, scale=4.0),
           "coarse_only": coarse.search(qs, K, batch=batch)}
    coarse.close(); exact.close()
    return out


def compare(a, b, what, exact=False, rtol=1e-4):
    """two suites' results: same counts and keys, scores bit-equal (exact) or within rtol"""
    assert a.keys() == b.keys(), what
    for name in a:
        ra, rb = a[name], b[name]
        if name == "fetch":
            assert np.array_equal(ra[0], rb[0]), "%s/%s" % (what, name)
            continue
        (ak, as_, ac), (bk, bs, bc) = ra[:3], rb[:3]
        assert np.array_equal(ac, bc), "%s/%s: counts %s vs %s" % (what, name, ac, bc)
        for i in range(len(ac)):
            n = int(ac[i])
            if exact:
                assert np.array_equal(as_[i][:n].view(np.uint32), bs[i][:n].view(np.uint32)), "%s/%s q%d scores" % (what, name, i)
            else:
                assert np.allclose(as_[i][:n], bs[i][:n], rtol=rtol, atol=1e-5), "%s/%s q%d scores" % (what, name, i)
            assert np.array_equal(ak[i][:n], bk[i][:n]), "%s/%s q%d keys\n%s\n%s" % (what, name, i, ak[i][:n], bk[i][:n])
        if len(ra) > 3:
            for i in range(len(ac)):
                n = int(ac[i])
                assert np.array_equal(ra[3][i][:n], rb[3][i][:n]), "%s/%s q%d fetched vectors" % (what, name, i)


def brute_force(x, qs, metric, k, keep=None):
    """numpy statement of what core_interface returns (user-facing scores): -> keys [q,k], scores [q,k]"""
    xd, qd = x.astype(np.float64), qs.astype(np.float64)
    if metric == "L2":
        s = ((qd[:, None, :] - xd[None, :, :]) ** 2).sum(-1)
        order = 1
    elif metric == "IP":
        s = qd @ xd.T
        order = -1
    else:
        s = 1.0 - (qd @ xd.T) / (np.linalg.norm(qd, axis=1)[:, None] * np.linalg.norm(xd, axis=1)[None, :])
        order = 1
    if keep is not None:
        s = np.where(keep[None, :], s, np.inf * order)
    idx = np.argsort(order * s, axis=1, kind="stable")[:, :k]
    return idx.astype(np.uint64), np.take_along_axis(s, idx, 1).astype(np.float32)
