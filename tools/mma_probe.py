#!/usr/bin/env python3
"""Perf triage of the tcgen05 scan: time zvb200_search_device on synthetic fp16 rows.
Usage: python tools/mma_probe.py [rows] [batch] [dim] [k]   (ZVB200_MMA_DEBUG selects probes)"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zvec_b200 as z

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
dim = int(sys.argv[3]) if len(sys.argv) > 3 else 768
k = int(sys.argv[4]) if len(sys.argv) > 4 else 10
kernel = int(os.environ.get("KERNEL", "3"))
dev = torch.device("cuda", 0)
st = z.Store(z.DT_FP16, z.METRIC_IP, dim)
st.reserve(rows)
g = torch.Generator(device=dev); g.manual_seed(1)
for c0 in range(0, rows, 500_000):
    n = min(500_000, rows - c0)
    x = torch.nn.functional.normalize(torch.randn((n, dim), generator=g, device=dev), dim=1).to(torch.float16)
    torch.cuda.synchronize()
    st.append_device(x.data_ptr(), n)
st.seal()
q = torch.nn.functional.normalize(torch.randn((batch, dim), generator=g, device=dev), dim=1).to(torch.float16)
os_ = torch.empty((batch, k), dtype=torch.float32, device=dev)
oi = torch.empty((batch, k), dtype=torch.int32, device=dev)
ctx = z.Context(st); ctx.set_kernel(kernel)
torch.cuda.synchronize()
ms = []
for i in range(8):
    ctx.search_device(q.data_ptr(), batch, k, os_.data_ptr(), oi.data_ptr())
    ctx.sync()
    ms.append(ctx.last_scan_ms())
best = min(ms[2:])
tf = 2.0 * rows * dim * batch / (best / 1e3) / 1e12
print("dbg=%s kernel=%s rows=%d batch=%d dim=%d k=%d: scan %.3f ms (best of %s)  %.1f TFLOP/s  %.1f GB/s algorithmic" % (
    os.environ.get("ZVB200_MMA_DEBUG", "0"), ctx.last_kernel(), rows, batch, dim, k, best,
    ["%.2f" % m for m in ms], tf, rows * dim * 2 / (best / 1e3) / 1e9))
