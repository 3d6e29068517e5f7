#!/usr/bin/env python3
"""Time Lloyd iterations on the device (zvb200_kmeans_device) on synthetic clustered rows.
Usage: python tools/kmeans_probe.py [rows] [dim] [k] [iterations]"""
import json, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zvec_b200 as z

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
k = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 5
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(5)
centres = torch.randn((k, d), generator=g, device=dev)
x = torch.empty((n, d), device=dev)
CH = 1 << 18
for o in range(0, n, CH):
    m = min(CH, n - o)
    sel = torch.randint(0, k, (m,), generator=g, device=dev)
    x[o:o + m] = centres[sel] + 0.3 * torch.randn((m, d), generator=g, device=dev)
init = x[torch.randperm(n, generator=g, device=dev)[:k]].cpu().numpy()
torch.cuda.synchronize()
z.kmeans(None, init, max_iterations=1, epsilon=0.0, rows_device_ptr=x.data_ptr(), n=n)     # warm-up
t0 = time.perf_counter()
cent, lab, fol, sse, it = z.kmeans(None, init, max_iterations=iters, epsilon=0.0, rows_device_ptr=x.data_ptr(), n=n)
el = time.perf_counter() - t0
passes = it + 1                                      # + the label() pass
print(json.dumps({"workload": "k-means %dx%d fp32, k=%d, %d iterations + label pass" % (n, d, k, it),
                  "seconds": el, "ms_per_pass": 1e3 * el / passes,
                  "assign_TFLOPs": 2.0 * n * d * k * passes / el / 1e12,
                  "empty_clusters": int((fol == 0).sum()), "sse": sse}))
