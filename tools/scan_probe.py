#!/usr/bin/env python3
"""Time zvb200_search_device for any store type on synthetic rows (device-generated).
Usage: python tools/scan_probe.py --rows N --dim D --batch B --k K --dtype fp32|fp16|int8q|int8 --metric l2|ip|cos --kernel 0..4"""
import argparse, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zvec_b200 as z

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_000_000)
ap.add_argument("--dim", type=int, default=128)
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--k", type=int, default=10)
ap.add_argument("--dtype", default="fp32")
ap.add_argument("--metric", default="l2")
ap.add_argument("--kernel", type=int, default=0)
ap.add_argument("--iters", type=int, default=10)
a = ap.parse_args()
dev = torch.device("cuda", 0)
if os.environ.get("ZVB200_PROBE_L2_GRAN"):   # cudaLimitMaxL2FetchGranularity (0x05): 32 / 64 / 128 bytes
    import ctypes
    torch.zeros(1, device=dev)
    rt = ctypes.CDLL("libcudart.so.12")
    print("cudaDeviceSetLimit ->", rt.cudaDeviceSetLimit(5, ctypes.c_size_t(int(os.environ["ZVB200_PROBE_L2_GRAN"]))))
    v = ctypes.c_size_t(0); rt.cudaDeviceGetLimit(ctypes.byref(v), 5); print("L2 fetch granularity now", v.value)
g = torch.Generator(device=dev); g.manual_seed(7)
d = a.dim
if a.dtype == "int8q":
    metric = {"l2": z.METRIC_QI8_L2SQ, "ip": z.METRIC_QI8_IP}[a.metric]; dt = z.DT_INT8; sdim = d + 16
elif a.dtype == "int8":
    metric = {"l2": z.METRIC_L2SQ, "ip": z.METRIC_IP}[a.metric]; dt = z.DT_INT8; sdim = d
else:
    metric = {"l2": z.METRIC_L2SQ, "ip": z.METRIC_IP, "cos": z.METRIC_COSINE, "cosine": z.METRIC_COSINE}[a.metric]
    dt = z.DT_FP32 if a.dtype == "fp32" else z.DT_FP16
    sdim = d + ((1 if dt == z.DT_FP32 else 2) if a.metric == "cos" else 0)


def make(n):
    if dt == z.DT_INT8:
        codes = torch.randint(-127, 128, (n, d), generator=g, device=dev, dtype=torch.int8)
        if a.dtype == "int8":
            return codes
        tail = torch.empty((n, 4), device=dev, dtype=torch.float32)
        tail[:, 0] = 1.0 / 127; tail[:, 1] = 0.0
        cf = codes.float()
        tail[:, 2] = cf.sum(1); tail[:, 3] = (cf * cf).sum(1)
        return torch.cat([codes.view(torch.uint8), tail.view(torch.uint8).view(n, 16)], dim=1).contiguous()
    x = torch.randn((n, d), generator=g, device=dev)
    if a.metric == "cos":
        nrm = x.norm(dim=1, keepdim=True)
        if dt == z.DT_FP32:
            return torch.cat([x / nrm, nrm], dim=1).contiguous()
        h = (x / nrm).half()
        return torch.cat([h.view(torch.uint8).view(n, 2 * d), nrm.view(torch.uint8).view(n, 4)], dim=1).contiguous()
    return x.contiguous() if dt == z.DT_FP32 else x.half().contiguous()


st = z.Store(dt, metric, sdim)
st.reserve(a.rows)
CH = 1_000_000
for c0 in range(0, a.rows, CH):
    x = make(min(CH, a.rows - c0)); torch.cuda.synchronize()
    st.append_device(x.data_ptr(), x.shape[0]); del x
st.seal()
q = make(a.batch)
os_ = torch.empty((a.batch, a.k), dtype=torch.float32, device=dev)
oi = torch.empty((a.batch, a.k), dtype=torch.int32, device=dev)
ctx = z.Context(st); ctx.set_kernel(a.kernel)
torch.cuda.synchronize()
ms = []
for i in range(a.iters):
    ctx.search_device(q.data_ptr(), a.batch, a.k, os_.data_ptr(), oi.data_ptr()); ctx.sync()
    ms.append(ctx.last_scan_ms())
best = min(ms[2:])
print("%s %s rows=%d dim=%d batch=%d k=%d kernel=%s: scan %.3f ms  %.1f GB/s algorithmic  %.1f TFLOP/s  (%.1f QPS)" % (
    a.dtype, a.metric, a.rows, d, a.batch, a.k, ctx.last_kernel(), best, st.scan_bytes() / (best / 1e3) / 1e9,
    2.0 * a.rows * d * a.batch / (best / 1e3) / 1e12, a.batch / (best / 1e3)))
