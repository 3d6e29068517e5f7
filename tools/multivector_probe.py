#!/usr/bin/env python3
"""C5 timing: multi-vector query — 4 vector fields (4 stores of rows x 768 fp16 InnerProduct, seeds 2024..2027),
top-100 per field, fused with RRF (rank_constant 60) and with the weighted IP reranker; plus the
shared-matrix variant (4 vectors against ONE store, one pass at batch 4).
Usage: python tools/multivector_probe.py [rows] [dim] [batches]"""
import json, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zvec_b200 as z

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 768
batches = [int(b) for b in sys.argv[3].split(",")] if len(sys.argv) > 3 else [1, 64]
F, k, topn = 4, 100, 100
dev = torch.device("cuda", 0)
stores, ctxs = [], []
for f in range(F):
    g = torch.Generator(device=dev); g.manual_seed(2024 + f)
    st = z.Store(z.DT_FP16, z.METRIC_IP, dim); st.reserve(rows)
    for c0 in range(0, rows, 1_000_000):
        n = min(1_000_000, rows - c0)
        x = torch.nn.functional.normalize(torch.randn((n, dim), generator=g, device=dev), dim=1).half()
        torch.cuda.synchronize(); st.append_device(x.data_ptr(), n); del x
    st.seal(); stores.append(st); ctxs.append(z.Context(st))
g = torch.Generator(device=dev); g.manual_seed(4202)
B = max(batches)
hq = [torch.nn.functional.normalize(torch.randn((B, dim), generator=g, device=dev), dim=1).half().cpu().numpy() for _ in range(F)]
out = {"workload": "multi-vector: %d fields x %dx%d fp16 InnerProduct, top%d per field, fused top%d" % (F, rows, dim, k, topn),
       "bytes_per_multi_query_batch": F * rows * dim * 2, "results": []}


def multi(b, mode):
    keys = np.empty((F, b, k), np.uint64); scores = np.empty((F, b, k), np.float32); counts = np.empty((F, b), np.uint32)
    scan = 0.0
    for f in range(F):
        kk, ss, _, cc = ctxs[f].search(hq[f][:b], k)
        keys[f], scores[f], counts[f] = kk, ss, cc
        scan += ctxs[f].last_scan_ms()
    if mode == "rrf":
        r = ctxs[0].fuse_rrf(keys, counts, topn, 60)
    else:
        r = ctxs[0].fuse_weighted(keys, scores, counts, topn, z.METRIC_IP, np.ones(F), negate=True)
    return r, scan


for b in batches:
    for mode in ("rrf", "weighted_ip"):
        for _ in range(2):
            multi(b, mode)
        n_it = max(3, min(20, 200 // b))
        t0 = time.perf_counter(); scan = 0.0
        for _ in range(n_it):
            r, s = multi(b, mode); scan += s
        el = time.perf_counter() - t0
        out["results"].append({"batch": b, "fusion": mode, "multi_qps_e2e": b * n_it / el, "ms_per_call": 1e3 * el / n_it,
                               "scan_ms_per_call": scan / n_it, "scan_GBs": F * rows * dim * 2 / (scan / n_it / 1e3) / 1e9,
                               "kernel": ctxs[1].last_kernel(), "fused_docs": int(r[2][0])})
# shared-matrix variant: the 4 vectors of a query against one store, one pass (batch 4 per multi-query)
for b in batches:
    q4 = np.concatenate([hq[f][:b] for f in range(F)], axis=0)          # [F*b, dim], field-major
    for _ in range(2):
        ctxs[0].search(q4, k)
    n_it = max(3, min(20, 200 // b))
    t0 = time.perf_counter(); scan = 0.0
    for _ in range(n_it):
        kk, ss, _, cc = ctxs[0].search(q4, k)
        scan += ctxs[0].last_scan_ms()
        r = ctxs[0].fuse_rrf(kk.reshape(F, b, k), cc.reshape(F, b), topn, 60)
    el = time.perf_counter() - t0
    out["results"].append({"batch": b, "fusion": "rrf, shared matrix", "multi_qps_e2e": b * n_it / el,
                           "ms_per_call": 1e3 * el / n_it, "scan_ms_per_call": scan / n_it,
                           "scan_GBs": rows * dim * 2 / (scan / n_it / 1e3) / 1e9, "kernel": ctxs[0].last_kernel()})
print(json.dumps(out))
