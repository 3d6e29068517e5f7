#!/usr/bin/env python3
"""Odd sizes through every round-2 kernel form, results asserted (compute-sanitizer is closed on the GPU pool, so this
is a plain run): whole-row preload + early tails (short QuantizedInteger rows), the all-keys coarse scan + select +
plan, the posting-list scan's preload form, the list-major scan, per-list score scales, graph replay."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zvec_b200 as z

rng = np.random.default_rng(5)
# short QuantizedInteger rows: odd row count, enough rows for the preload form
n, d = 650003, 96
codes = rng.integers(-127, 128, (n, d), dtype=np.int8)
tail = np.zeros((n, 4), np.float32); tail[:, 0] = 1.0 / 127
cf = codes.astype(np.float32); tail[:, 2] = cf.sum(1); tail[:, 3] = (cf * cf).sum(1)
rows = np.concatenate([codes.view(np.uint8), tail.view(np.uint8).reshape(n, 16)], axis=1)
st = z.Store(z.DT_INT8, z.METRIC_QI8_L2SQ, d + 16); st.append(rows); st.seal(); ctx = z.Context(st)
for b in (1, 2, 3):
    for _ in range(3):
        r = ctx.search(rows[11:11 + b], 100)
    assert r[2][0][0] == 11, r[2][0][:4]
print("qi8 short rows ok", ctx.last_kernel())
ctx.close(); st.close()
# IVF, fp32 long rows (preload form of the list scan), odd sizes, an empty list, 1 / 3 / 20 queries
n, d, nlist = 30011, 192, 157
cent = rng.standard_normal((nlist, d)).astype(np.float32)
lab = np.sort(rng.integers(0, nlist, n)); lab[lab == 5] = 6
x = (cent[lab] + 0.3 * rng.standard_normal((n, d))).astype(np.float32)
offs = np.searchsorted(lab, np.arange(nlist + 1)).astype(np.uint64)
qs = (cent[rng.integers(0, nlist, 20)] + 0.3 * rng.standard_normal((20, d))).astype(np.float32)
st = z.Store(z.DT_FP32, z.METRIC_L2SQ, d); st.append(x); st.seal(); st.ivf_attach(cent, offs); ctx = z.Context(st)
for b in (1, 3, 20):
    for _ in range(3):
        r = ctx.ivf_search(qs[:b], 10, 13, 2500)
    assert (r[3] == 10).all()
    print("ivf batch", b, ctx.last_kernel())
st.set_score_scale(1.0, (0.5 + rng.random(nlist)).astype(np.float32))
for b in (1, 20):
    r = ctx.ivf_search(qs[:b], 10, 13, 2500)
    assert (r[3] == 10).all()
r = ctx.search(qs[:2], 10, threshold=float(np.median(r[1])))
print("scaled ok", ctx.last_kernel(), r[3])
ctx.close(); st.close()
print("EDGE-PROBE-OK")
