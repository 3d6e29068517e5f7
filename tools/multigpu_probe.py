#!/usr/bin/env python3
"""Multi-GPU timing of the two paths bench.py does not cover, one process per GPU under torchrun:
  ivf : C4-shaped IVF, every rank holds a slice of EVERY posting list (zvb200_ivf_set_budget_sizes keeps the list
        visits identical to the unsharded index), per-rank zvb200_ivf_search, all-gather of (score, global row),
        merge by (score, global row)
  mv  : C5 multi-vector, 4 fields x rows x 768 fp16, every field row-sharded, per-field gather + device merge
        (zvb200_merge_topk_device), fusion (zvb200_fuse_rrf) on rank 0
Usage: torchrun --nproc-per-node N tools/multigpu_probe.py ivf|mv [rows] [batch]"""
import json, math, os, sys, time
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zvec_b200 as z
from zvec_b200 import sharding

mode = sys.argv[1]
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, iters):
    for _ in range(3):
        fn()
    barrier(); t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    barrier(); el = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([el], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); el = float(t.item())
    return el / iters


if mode == "ivf":
    rows = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
    batch = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
    dim, nlist, nprobe, k = 1536, 16384, 64, 10
    sizes = np.full(nlist, rows // nlist, np.int64); sizes[: rows % nlist] += 1
    off = np.zeros(nlist + 1, np.uint64); off[1:] = np.cumsum(sizes)
    lo, hi, loff = sharding.ivf_shard_plan(off, rank, world)
    g = torch.Generator(device=dev); g.manual_seed(1536)          # same centres on every rank
    centres = torch.randn((nlist, dim), generator=g, device=dev)
    st = z.Store(z.DT_FP32, z.METRIC_COSINE, dim + 1, device=local); st.reserve(int(loff[-1]))
    cnt = torch.from_numpy((hi - lo).astype(np.int64)).to(dev)
    gl = torch.Generator(device=dev); gl.manual_seed(7000 + rank)
    for l0 in range(0, nlist, 512):
        l1 = min(nlist, l0 + 512)
        rep = torch.repeat_interleave(torch.arange(l0, l1, device=dev), cnt[l0:l1])
        x = centres[rep] + 0.3 * torch.randn((rep.numel(), dim), generator=gl, device=dev)
        nrm = x.norm(dim=1, keepdim=True)
        x = torch.cat([x / nrm, nrm], dim=1).contiguous(); torch.cuda.synchronize()
        st.append_device(x.data_ptr(), x.shape[0]); del x, rep, nrm
    st.seal()
    cn = centres.norm(dim=1, keepdim=True)
    st.ivf_attach(torch.cat([centres / cn, cn], dim=1).cpu().numpy(), loff.astype(np.uint64))
    st.ivf_set_budget_sizes(sizes.astype(np.uint64))
    ctx = z.Context(st)
    max_scan = max(1000, int(math.ceil(rows * nprobe / nlist)))
    gq = torch.Generator(device=dev); gq.manual_seed(99)
    qsel = torch.randint(0, nlist, (batch,), generator=gq, device=dev)
    q = centres[qsel] + 0.3 * torch.randn((batch, dim), generator=gq, device=dev)
    qn = q.norm(dim=1, keepdim=True)
    hq = torch.cat([q / qn, qn], dim=1).contiguous()               # pinned-host -> device inside the step (e2e)
    hq_host = hq.cpu().pin_memory()
    st.ivf_set_global_bases(lo.astype(np.uint64))
    scan = []
    dq = torch.empty_like(hq)
    l_s = torch.empty((batch, k), dtype=torch.float32, device=dev); l_i = torch.empty((batch, k), dtype=torch.int32, device=dev)
    l_r = torch.empty((batch, k), dtype=torch.int64, device=dev)
    a_s = torch.empty((world, batch, k), dtype=torch.float32, device=dev); a_r = torch.empty((world, batch, k), dtype=torch.int64, device=dev)
    f_s = torch.empty((batch, k), dtype=torch.float32, device=dev); f_r = torch.empty((batch, k), dtype=torch.int64, device=dev)

    def step():
        # everything stays on the device: search, local -> global rows, all-gather, merge by (score, global row)
        dq.copy_(hq_host, non_blocking=True); torch.cuda.synchronize()
        ctx.ivf_search_device(dq.data_ptr(), batch, k, nprobe, max_scan, l_s.data_ptr(), l_i.data_ptr())
        ctx.ivf_global_rows_device(l_i.data_ptr(), batch * k, l_r.data_ptr())
        ctx.sync()
        scan.append(ctx.last_scan_ms())
        if world > 1:
            dist.all_gather_into_tensor(a_s.view(world * batch, k), l_s); dist.all_gather_into_tensor(a_r.view(world * batch, k), l_r)
            torch.cuda.synchronize()
            ctx.merge_topk_rows_device(a_s.data_ptr(), a_r.data_ptr(), world, batch, k, f_s.data_ptr(), f_r.data_ptr()); ctx.sync()
            return f_s.cpu(), f_r.cpu()
        return l_s.cpu(), l_r.cpu()

    ms = timed(step, 5) * 1e3
    if rank == 0:
        print(json.dumps({"workload": "IVF nlist %d nprobe %d fp32 cosine %dx%d, every list sharded over %d GPU(s), batch %d" % (nlist, nprobe, rows, dim, world, batch),
                          "n_gpus": world, "qps_e2e": batch / (ms / 1e3), "ms_per_batch": ms, "list_scan_ms_rank0": float(np.mean(scan[3:])),
                          "merge": "device: zvb200_ivf_global_rows_device + NCCL all-gather + zvb200_merge_topk_rows_device"}))
else:
    rows = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
    batch = int(sys.argv[3]) if len(sys.argv) > 3 else 64
    dim, F, k, topn = 768, 4, 100, 100
    lo, hi = sharding.shard_bounds(rows, world)[rank]
    bases = sharding.shard_bases(rows, world)
    stores, ctxs = [], []
    for f in range(F):
        gf = torch.Generator(device=dev); gf.manual_seed(2024 + f)
        st = z.Store(z.DT_FP16, z.METRIC_IP, dim, device=local); st.reserve(hi - lo)
        for c0 in range(0, rows, 1_000_000):                    # same stream on every rank: global row = its position
            n = min(1_000_000, rows - c0)
            x = torch.nn.functional.normalize(torch.randn((n, dim), generator=gf, device=dev), dim=1).half()
            a, b = max(lo, c0), min(hi, c0 + n)
            if b > a:
                part = x[a - c0:b - c0].contiguous(); torch.cuda.synchronize(); st.append_device(part.data_ptr(), b - a)
            del x
        st.seal(); stores.append(st); ctxs.append(z.Context(st))
    gq = torch.Generator(device=dev); gq.manual_seed(4202)
    dq = [torch.nn.functional.normalize(torch.randn((batch, dim), generator=gq, device=dev), dim=1).half() for _ in range(F)]
    ls = torch.empty((batch, k), dtype=torch.float32, device=dev); li = torch.empty((batch, k), dtype=torch.int32, device=dev)
    a_s = torch.empty((world, batch, k), dtype=torch.float32, device=dev); a_i = torch.empty((world, batch, k), dtype=torch.int32, device=dev)
    fs = torch.empty((batch, k), dtype=torch.float32, device=dev); fr = torch.empty((batch, k), dtype=torch.int64, device=dev)
    stream = torch.cuda.ExternalStream(ctxs[0].stream())
    scan = []

    def step():
        keys = np.empty((F, batch, k), np.uint64); counts = np.full((F, batch), k, np.uint32)
        tot = 0.0
        for f in range(F):
            ctxs[f].search_device(dq[f].data_ptr(), batch, k, ls.data_ptr(), li.data_ptr())
            ctxs[f].sync(); tot += ctxs[f].last_scan_ms()
            if world > 1:
                dist.all_gather_into_tensor(a_s.view(world * batch, k), ls); dist.all_gather_into_tensor(a_i.view(world * batch, k), li)
            else:
                a_s[0], a_i[0] = ls, li
            torch.cuda.synchronize()
            ctxs[f].merge_topk_device(a_s.data_ptr(), a_i.data_ptr(), world, bases, batch, k, fs.data_ptr(), fr.data_ptr())
            ctxs[f].sync()
            if rank == 0:
                keys[f] = fr.cpu().numpy().view(np.uint64)
        scan.append(tot)
        if rank == 0:
            return ctxs[0].fuse_rrf(keys, counts, topn, 60)
        return None

    ms = timed(step, 5) * 1e3
    if rank == 0:
        print(json.dumps({"workload": "multi-vector %d fields x %dx%d fp16 IP top%d fused, rows sharded over %d GPU(s), batch %d" % (F, rows, dim, k, world, batch),
                          "n_gpus": world, "multi_qps_e2e": batch / (ms / 1e3), "ms_per_batch": ms, "scan_ms_rank0": float(np.mean(scan[3:]))}))
for c in (ctxs if mode != "ivf" else [ctx]):
    c.close()
if world > 1:
    dist.barrier(); dist.destroy_process_group()
