"""The boundary one layer up: core_interface::Index (src/core/interface/index.cc), the object the DB layer calls
(vector_column_indexer.cc:195) — FlatIndex / IVFIndex created by IndexFactory::CreateAndInitIndex, Open, Add, Train,
Search (plain, linear, filter, radius, bf_pks, refiner_param, fetch_vector), Fetch.  The reference's six TUs of
src/core/interface are compiled into oracle/_ref by oracle/build_ref.py.

CPU tests pin the shim against numpy with the STOCK classes.  The GPU tests run every scenario twice in a process
of their own: stock classes first, then after zvb200_install_overrides() — the same calls, now answered by
GpuFlatStreamer / GpuIVFStreamer by name — and compare; plus the batch entry integration/index_batch.cc (N4)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import ci_scenarios as CS
import support as S

needs_ref = pytest.mark.skipif(S.ref() is None, reason="oracle/_ref not built here")
needs_plugin = pytest.mark.skipif(S.ref() is None or not S.load_plugin(),
                                  reason="oracle/_ref or integration/_build not built here")


@needs_ref
@pytest.mark.parametrize("metric", ["L2", "IP", "COSINE"])
def test_core_interface_flat_matches_numpy_with_the_stock_classes(metric, tmp_path):
    x, qs = CS.data()
    out = CS.flat_suite(str(tmp_path), metric, "")
    bk, bs = CS.brute_force(x, qs, metric, CS.K)
    k, s, c = out["plain"]
    assert np.all(c == CS.K) and np.array_equal(k, bk) and np.allclose(s, bs, rtol=1e-4, atol=1e-5)
    k, s, c = out["linear"]
    assert np.array_equal(k, bk)
    keep = np.arange(CS.N) % 3 != 1
    fk, fs = CS.brute_force(x, qs, metric, CS.K, keep)
    k, s, c = out["filter"]
    assert np.array_equal(k, fk) and np.allclose(s, fs, rtol=1e-4, atol=1e-5)
    pks = np.arange(5, CS.N, 13)
    keep = np.zeros(CS.N, bool); keep[pks] = True
    pk, ps = CS.brute_force(x, qs, metric, CS.K, keep)
    k, s, c = out["bf_pks"]
    assert np.array_equal(k, pk) and np.allclose(s, ps, rtol=1e-4, atol=1e-5)
    # fetch_vector hands the stored vectors back (cosine: the index reverts the normalisation)
    k, s, c, v = out["fetch_vector"]
    got = v.view(np.float32).reshape(4, 5, CS.D)
    assert np.allclose(got, x[k.astype(np.int64)], rtol=1e-5, atol=1e-6)
    assert np.allclose(out["fetch"][0].view(np.float32).reshape(3, CS.D), x[[0, 17, CS.N - 1]], rtol=1e-5, atol=1e-6)


@needs_ref
def test_core_interface_refiner_with_the_stock_classes(tmp_path):
    x, qs = CS.data(12)
    out = CS.refiner_suite(str(tmp_path))
    bk, bs = CS.brute_force(x, qs, "L2", CS.K)
    k, s, c = out["refined"]
    # the exact second stage re-scores 40 int8 candidates: its answers are exact fp32 distances of true neighbours
    assert np.all(c == CS.K)
    assert (k == bk).mean() > 0.95
    same = k == bk
    assert np.allclose(s[same], bs[same], rtol=1e-4, atol=1e-5)


@needs_plugin
def test_search_batch_of_one_equals_search_with_the_stock_classes(tmp_path):
    """integration/index_batch.cc on the stock streamer: count = 1 per call is Index::Search itself (the stock
    FlatStreamer's batch loop does not reset its heap between queries — INTEGRATION.md §5 — so count > 1 is compared
    on the GPU classes only)."""
    S.plugin_handle()
    x, qs = CS.data()
    def body():
        ix = S.CiIndex("Flat", CS.D, "L2", "fp16", path=str(tmp_path / "b1.index"))
        ix.add(x)
        for i in range(4):
            a = ix.search(qs[i:i + 1], CS.K)
            b = ix.search(qs[i:i + 1], CS.K, batch=True)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))
        pks = np.arange(0, 500, 3, dtype=np.uint64)
        a = ix.search(qs[:1], CS.K, bf_pks=pks)
        b = ix.search(qs[:1], CS.K, bf_pks=pks, batch=True)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))
        ix.close()
    CS.in_thread(body)


GPU_SCRIPT = r"""
import ctypes as C, os, sys, numpy as np
sys.path.insert(0, %(tests)r); sys.path.insert(0, %(root)r)
import support as S, ci_scenarios as CS
import zvec_b200 as z
tmp = %(tmp)r
plug = S.plugin_handle()
L = z.lib()

# ---- 1. every scenario with the stock classes ------------------------------------------------------------
cases = [("L2", ""), ("IP", ""), ("COSINE", ""), ("IP", "fp16"), ("L2", "fp16"), ("L2", "int8"), ("IP", "int8")]
stock = {c: CS.flat_suite(tmp + "/s", *c) for c in cases}
stock_ref = CS.refiner_suite(tmp + "/s")
# an IVF index trained by the stock builder; both streamer classes will open the same file
x, qs = CS.data(13)
ivf_path = tmp + "/ivf.index"
def ivf_suite(open_existing):
    ivf = S.CiIndex("IVF", CS.D, "L2", "", nlist=32, path=ivf_path, open_existing=open_existing)
    if not open_existing:
        ivf.add(x); assert ivf.train() == 0
    out = {"plain": ivf.search(qs, CS.K), "linear": ivf.search(qs, CS.K, linear=True),
           "fetch_vector": ivf.search(qs[:4], 5, fetch_vector=True)}
    S.ref().zref_set_filter(3, 1); out["filter"] = ivf.search(qs, CS.K); S.ref().zref_set_filter(0, 0)
    if open_existing: ivf.close()
    else: ivf.h = None   # keep the file: the destructor would unlink it
    return out
ivf_stock = CS.in_thread(ivf_suite, False)

# ---- 2. the same calls after the overrides: the stock NAMES now build the GPU classes ------------------------
assert plug.zvb200_install_overrides() == 0
before = L.zvb200_kernel_launches()
for c in cases:
    gpu = CS.flat_suite(tmp + "/g", *c)
    CS.compare(gpu, stock[c], "core_interface Flat %%s/%%s" %% c, exact=c[1] == "int8")
    gpu_b = CS.flat_suite(tmp + "/g", *c, batch=True)          # the batch entry, all queries in ONE search_impl call
    CS.compare(gpu_b, stock[c], "SearchBatch Flat %%s/%%s" %% c, exact=c[1] == "int8")
assert L.zvb200_kernel_launches() > before, "the stock names did not reach the GPU classes"
CS.compare(CS.refiner_suite(tmp + "/g"), stock_ref, "refiner")
CS.compare(CS.refiner_suite(tmp + "/g", batch=True), stock_ref, "refiner through SearchBatch")
before = L.zvb200_kernel_launches()
ivf_gpu = CS.in_thread(ivf_suite, True)
assert L.zvb200_kernel_launches() > before
CS.compare(ivf_gpu, ivf_stock, "core_interface IVF")
os.unlink(ivf_path)

# ---- 3. N4: 1024 fp32 queries through the batch entry = ONE tensor-core scan, results = 1024 single calls -----
rng = np.random.default_rng(21)
n, d, nq, k = 60000, 96, 1024, 10
xb = rng.standard_normal((n, d)).astype(np.float32)
qb = rng.standard_normal((nq, d)).astype(np.float32)
def n4():
    ix = S.CiIndex("Flat", d, "IP", "fp16", path=tmp + "/n4.index")
    ix.add(xb)
    singles = ix.search(qb, k)                                     # 1024 x Index::Search (count = 1 each)
    l0 = L.zvb200_kernel_launches()
    batched = ix.search(qb, k, batch=True)
    launches = L.zvb200_kernel_launches() - l0
    CS.compare({"n4": batched}, {"n4": singles}, "SearchBatch 1024 vs singles")
    # prep_queries + ONE scan_mma launch + merge_parts: the batch reached the GPU as one search_impl(count = 1024)
    assert launches <= 4, launches
    print("launches for 1024 queries through SearchBatch:", launches)
    ix.close()
CS.in_thread(n4)
print("CORE-INTERFACE-OK")
"""


@needs_plugin
@pytest.mark.gpu
def test_core_interface_reaches_the_gpu_classes_by_name(tmp_path):
    here = os.path.dirname(os.path.abspath(__file__))
    os.makedirs(tmp_path / "s"); os.makedirs(tmp_path / "g")
    code = GPU_SCRIPT % {"tests": here, "root": os.path.dirname(here), "tmp": str(tmp_path)}
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=1500)
    assert out.returncode == 0 and "CORE-INTERFACE-OK" in out.stdout, out.stdout[-3000:] + out.stderr[-6000:]
