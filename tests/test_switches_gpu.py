"""The round-2 fast paths (CUDA-graph replay, radix-select merge, direct coarse stage, whole-row preload, half-tile pools) against the
forms they replaced: the library reads its switches once per process, so each form runs in a process of its own on
the same seeded inputs and the results are compared bit for bit."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, %(root)r)
import zvec_b200 as z
out = {}
rng = np.random.default_rng(2024)
# flat fp32 L2, 1 and 3 queries, fourth call of each (a replayed graph when replay is on)
x = rng.standard_normal((50000, 64)).astype(np.float32)
q = rng.standard_normal((3, 64)).astype(np.float32)
st = z.Store(z.DT_FP32, z.METRIC_L2SQ, 64); st.append(x); st.seal(); ctx = z.Context(st)
for b in (1, 3):
    for _ in range(4):
        r = ctx.search(q[:b], 10)
    out["flat%%d_i" %% b], out["flat%%d_s" %% b] = r[2], r[1]
q40 = rng.standard_normal((40, 64)).astype(np.float32)
r = ctx.search(q40, 50)                      # tensor-core scan, shared-memory pools of a half query tile
assert ctx.last_kernel() == "scan_mma"
out["mma40_i"], out["mma40_s"] = r[2], r[1]
r = ctx.search(q, 10, threshold=float(np.median(out["flat3_s"])))
out["thr_i"], out["thr_s"], out["thr_c"] = r[2], r[1], r[3]
ctx.close(); st.close()
# short QuantizedInteger rows, enough of them for the whole-row preload form
n, d = 700000, 96
codes = rng.integers(-127, 128, (n, d), dtype=np.int8)
tail = np.zeros((n, 4), np.float32); tail[:, 0] = 1.0 / 127
cf = codes.astype(np.float32); tail[:, 2] = cf.sum(1); tail[:, 3] = (cf * cf).sum(1)
rows = np.concatenate([codes.view(np.uint8), tail.view(np.uint8).reshape(n, 16)], axis=1)
st = z.Store(z.DT_INT8, z.METRIC_QI8_L2SQ, d + 16); st.append(rows); st.seal(); ctx = z.Context(st)
for _ in range(4):
    r = ctx.search(rows[5:7], 100)
out["qi8_i"], out["qi8_s"] = r[2], r[1]
ctx.close(); st.close()
# IVF: coarse stage + plan + list scan at 1 and 3 queries
n, d, nlist = 40000, 64, 200
cent = rng.standard_normal((nlist, d)).astype(np.float32)
lab = np.sort(rng.integers(0, nlist, n))
x = (cent[lab] + 0.3 * rng.standard_normal((n, d))).astype(np.float32)
offs = np.searchsorted(lab, np.arange(nlist + 1)).astype(np.uint64)
qs = (cent[[3, 77, 150]] + 0.3 * rng.standard_normal((3, d))).astype(np.float32)
st = z.Store(z.DT_FP32, z.METRIC_L2SQ, d); st.append(x); st.seal(); st.ivf_attach(cent, offs); ctx = z.Context(st)
for b in (1, 3):
    for _ in range(4):
        r = ctx.ivf_search(qs[:b], 10, 16, 3000)
    out["ivf%%d_i" %% b], out["ivf%%d_s" %% b], out["ivf%%d_c" %% b] = r[2], r[1], r[3]
ctx.close(); st.close()
np.savez(sys.argv[1], **out)
"""


def _run(tmp_path, name, env):
    path = str(tmp_path / (name + ".npz"))
    e = dict(os.environ); e.update(env)
    p = subprocess.run([sys.executable, "-c", SCRIPT % {"root": ROOT}, path], env=e, capture_output=True, text=True,
                       timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    return np.load(path)


def test_fast_paths_equal_the_forms_they_replaced(tmp_path):
    base = _run(tmp_path, "default", {})
    assert base["ivf1_c"][0] == 10 and base["thr_c"].sum() > 0 and base["thr_c"].min() < 10
    for name, env in (("no_graph", {"ZVB200_GRAPH": "0"}), ("no_select", {"ZVB200_MERGE_SELECT": "0"}),
                      ("no_direct", {"ZVB200_COARSE_DIRECT": "0"}), ("no_pre", {"ZVB200_GEMV_PRE": "0"}),
                      ("full_pools", {"ZVB200_MMA_HALF_POOL": "0"})):
        other = _run(tmp_path, name, env)
        for key in base.files:
            a, b = base[key], other[key]
            same = np.array_equal(a.view(np.uint32), b.view(np.uint32)) if a.dtype == np.float32 else np.array_equal(a, b)
            assert same, (name, key)
