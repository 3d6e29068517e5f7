#!/usr/bin/env python3
"""bench.py — QPS@top10 of the exhaustive scan on BASELINE.json configs[1]:
Flat fp16 inner product over synthetic 10M x 768 (Cohere-shape), batch 1024 (headline) and
batch 1, rows sharded over N GPUs of one node (strong scaling), per-GPU top-k merged with an
NCCL all-gather + the k-way merge kernel.

  python bench.py --gpus N --steps K --warmup W            # our CUDA path (torchrun for N>1)
  python bench.py --impl reference --gpus N ...            # the reference's CPU path, host cores

A "step" is one search of one batch.  `value` is whole-job QPS with queries already in HBM,
timed with CUDA events on the stream the kernels run on, max over ranks.  `e2e` is the same
metric through the host-buffer C-ABI call (H2D of the queries and D2H of the results inside
the timed region).  PyTorch is only plumbing here (device memory for synthetic data,
torch.distributed); every kernel in the timed region is from libzvec_b200.so.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DIM = 768
SEED_BASE, SEED_QUERY = 2024, 4202
CHUNK = 500_000  # rows generated per device chunk


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--topk", type=int, default=10)
    ap.add_argument("--kernel", type=int, default=0, help="0 auto, 1 gemv, 2 tile, 3 tcgen05")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-rows", type=int, default=200_000)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sustained": d["bf16_tflops_sustained"],
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nme, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# -----------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own FlatSearcher on the host cores
# -----------------------------------------------------------------------------------------
def host_queries(n):
    rng = np.random.default_rng(SEED_QUERY)
    q = rng.standard_normal((n, DIM)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q


class CpuArm:
    """The reference's CPU implementation of the path on this box's host cores: its own
    FlatSearcher over a HalfFloatConverter index (oracle/_ref, compiled from the reference's
    sources), driven like tools/core/bench_original.cc: T threads, one context each, looping
    search_impl(q, qmeta, batch, ctx).  Bounded sample: the first `cpu_sample_rows` rows of
    the workload; QPS is scaled linearly by rows (a flat scan is linear in N)."""

    def __init__(self, args):
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import support as S
        self.S, self.args = S, args
        self.cores = os.cpu_count() or 1
        self.ns = min(args.cpu_sample_rows, args.rows)
        rng = np.random.default_rng(SEED_BASE)
        x = rng.standard_normal((self.ns, DIM)).astype(np.float32)
        x /= np.linalg.norm(x, axis=1, keepdims=True)
        if S.ref() is not None:
            self.kind, self.lib = "reference", S.ref()._name
            self.ri = S.RefIndex(x, S.DT_FP32, DIM, "InnerProduct", "HalfFloatConverter")
            self.nq = max(self.cores * 4, 32)
            self.qs = host_queries(self.nq)
            self.ri.search(self.qs[:self.cores], args.topk, batch=1, threads=self.cores)  # page in the mmap
        else:  # the scalar restatement (1 thread)
            self.kind, self.cores, self.lib, self.ri = "port", 1, "libzvec_oracle.so", None
            self.xs = x.astype(np.float16).view(np.uint8).reshape(self.ns, -1)
            self.nq = 8
            self.qs = host_queries(8).astype(np.float16).view(np.uint8).reshape(8, -1)

    def measure(self, min_seconds, batch):
        S, args = self.S, self.args
        done, t0 = 0, time.perf_counter()
        while True:
            if self.ri is not None:
                self.ri.search(self.qs, args.topk, batch=batch, threads=self.cores)
            else:
                S.oracle_flat_search(S.M_IP, S.DT_FP16, self.xs, DIM, self.qs, args.topk)
            done += self.nq
            el = time.perf_counter() - t0
            if el >= min_seconds:
                break
        qps_sample = done / el
        return {"value": qps_sample * self.ns / args.rows, "unit": "queries/s", "cores": self.cores, "kind": self.kind,
                "sample": "%s FlatSearcher (row-major, HalfFloatConverter, batch %d per search_impl call, %d threads) "
                          "on %d of %d rows for %.1f s (%d queries); QPS scaled linearly by rows" %
                          (self.lib, batch, self.cores, self.ns, args.rows, el, done),
                "qps_on_sample": qps_sample}

    def best(self, seconds):
        """the faster of the reference's two call shapes (batch 1 = what collection.query issues,
        batch 32 = its largest tile) -- the reference arm gets its best configuration"""
        a = self.measure(seconds / 2, 1)
        b = self.measure(seconds / 2, 32) if self.ri is not None else a
        return a if a["value"] >= b["value"] else b

    def close(self):
        if self.ri is not None:
            self.ri.close()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm(args)
    # each "step" = a bounded sample timed for a few seconds; K+W steps stay within minutes
    per_step = max(2.0, min(10.0, 90.0 / max(1, args.steps + args.warmup)))
    for _ in range(min(args.warmup, 1)):
        arm.best(2.0)
    vals, base = [], None
    for _ in range(max(1, min(args.steps, 8))):
        base = arm.best(per_step)
        vals.append(base["value"])
    arm.close()
    v = float(np.mean(vals))
    base["value"] = v
    out = {"impl": "reference", "metric": "QPS@top10 (batch 1024)", "value": v, "unit": "queries/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": 1000.0 * args.batch / v, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f16", "data": "synthetic",
           "config": workload_config(args, args.gpus), "cpu_baseline": base,
           "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    emit(out)


def workload_config(args, n):
    return {"workload": "Flat fp16 InnerProduct, synthetic %dx%d (Cohere-shape), top%d, batch %d (headline) and "
                        "batch 1; rows sharded over %d GPU(s)" % (args.rows, DIM, args.topk, args.batch, n),
            "rows": args.rows, "dim": DIM, "topk": args.topk, "batch": args.batch, "shards": n,
            "l2_policy": "inputs larger than L2: every step streams the whole %.1f GB shard (L2 is 126 MB)" %
                         (args.rows / n * DIM * 2 / 1e9)}


def run_ours(args):
    import torch
    import torch.distributed as dist
    import zvec_b200 as z

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch N>1 with torch.distributed.run)" % (args.gpus, world))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pk = peaks()

    # ---- the shard: rows [lo, hi) of the global matrix, generated chunk by chunk on the device
    from zvec_b200 import sharding
    lo, hi = sharding.shard_bounds(args.rows, world)[rank]
    st = z.Store(z.DT_FP16, z.METRIC_IP, DIM, device=local)
    st.reserve(hi - lo)
    g = torch.Generator(device=dev)
    c0 = (lo // CHUNK) * CHUNK
    while c0 < hi:
        g.manual_seed(SEED_BASE + c0 // CHUNK)           # data depends on the global row only
        n = min(CHUNK, args.rows - c0)
        x = torch.randn((n, DIM), generator=g, device=dev, dtype=torch.float32)
        x = torch.nn.functional.normalize(x, dim=1).to(torch.float16)
        a, b = max(lo, c0) - c0, min(hi, c0 + n) - c0
        xs = x[a:b].contiguous()
        torch.cuda.synchronize()
        st.append_device(xs.data_ptr(), b - a)
        del x, xs
        c0 += CHUNK
    st.seal()
    torch.cuda.empty_cache()
    nloc = st.count()
    shard_bytes = st.scan_bytes()

    hq = torch.from_numpy(host_queries(args.batch)).to(torch.float16).pin_memory()
    dq = hq.to(dev)
    k, B = args.topk, args.batch
    ctx = z.Context(st)
    ctx.set_kernel(args.kernel)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    loc_s = torch.empty((B, k), dtype=torch.float32, device=dev)
    loc_i = torch.empty((B, k), dtype=torch.int32, device=dev)
    all_s = torch.empty((world, B, k), dtype=torch.float32, device=dev)
    all_i = torch.empty((world, B, k), dtype=torch.int32, device=dev)
    fin_s = torch.empty((B, k), dtype=torch.float32, device=dev)
    fin_r = torch.empty((B, k), dtype=torch.int64, device=dev)
    h_s = torch.empty((B, k), dtype=torch.float32).pin_memory()
    h_r = torch.empty((B, k), dtype=torch.int64).pin_memory()
    bases = sharding.shard_bases(args.rows, world)

    def step(q):
        """one batch of q queries already in HBM -> final top-k in HBM (all on ctx's stream)"""
        ctx.search_device(dq.data_ptr(), q, k, loc_s.data_ptr(), loc_i.data_ptr())
        if world > 1:
            with torch.cuda.stream(stream):
                dist.all_gather_into_tensor(all_s[:, :q].contiguous() if q != B else all_s, loc_s[:q] if q != B else loc_s)
                dist.all_gather_into_tensor(all_i[:, :q].contiguous() if q != B else all_i, loc_i[:q] if q != B else loc_i)
            ctx.merge_topk_device(all_s.data_ptr(), all_i.data_ptr(), world, bases, q, k, fin_s.data_ptr(), fin_r.data_ptr())

    # gather buffers for q != B need their own tight tensors
    small = {}

    def step_q(q):
        if q == B or world == 1:
            return step(q)
        if q not in small:
            small[q] = (torch.empty((world, q, k), dtype=torch.float32, device=dev),
                        torch.empty((world, q, k), dtype=torch.int32, device=dev))
        ss, si = small[q]
        ctx.search_device(dq.data_ptr(), q, k, loc_s.data_ptr(), loc_i.data_ptr())
        with torch.cuda.stream(stream):
            dist.all_gather_into_tensor(ss, loc_s.view(-1)[: q * k].view(q, k))
            dist.all_gather_into_tensor(si, loc_i.view(-1)[: q * k].view(q, k))
        ctx.merge_topk_device(ss.data_ptr(), si.data_ptr(), world, bases, q, k, fin_s.data_ptr(), fin_r.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(q, steps, warmup):
        for _ in range(warmup):
            step_q(q)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = z.lib().zvb200_kernel_launches()
        e0.record(stream)
        for _ in range(steps):
            step_q(q)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        launches = z.lib().zvb200_kernel_launches() - l0
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_b, launches = timed(B, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    kernel_b = ctx.last_kernel()
    # dominant-kernel duration, measured live with CUDA events around the scan kernel only
    scan_ms = []
    for _ in range(args.steps):
        step_q(B)
        scan_ms.append(ctx.last_scan_ms())
    scan_b = float(np.mean(scan_ms))
    # batch 1
    s1 = max(args.steps, 30)
    ms_1, _ = timed(1, s1, max(args.warmup, 3))
    kernel_1 = ctx.last_kernel()
    scan_ms = []
    for _ in range(10):
        step_q(1)
        scan_ms.append(ctx.last_scan_ms())
    scan_1 = float(np.mean(scan_ms))

    # ---- e2e: the call a user makes — host buffers in, host results out
    def e2e_step(q):
        if world == 1:
            ctx.search(hq.numpy()[:q], k)        # zvb200_search: H2D + kernels + D2H + sync
            return
        with torch.cuda.stream(stream):
            dq[:q].copy_(hq[:q], non_blocking=True)
        step_q(q)
        with torch.cuda.stream(stream):
            h_s[:q].copy_(fin_s[:q], non_blocking=True)
            h_r[:q].copy_(fin_r[:q], non_blocking=True)
        stream.synchronize()

    def e2e_timed(q, steps):
        for _ in range(3):
            e2e_step(q)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_step(q)
        barrier()
        el = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([el], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            el = float(t.item())
        return el

    el_b = e2e_timed(B, args.steps)
    el_1 = e2e_timed(1, s1)

    if rank == 0:
        def traffic(kernel, batch):
            """measured DRAM bytes per launch (ncu capture committed under profiles/), when one exists for
            exactly this kernel and shard shape; else null"""
            try:
                with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                    for e in json.load(f)["entries"]:
                        if (e["kernel"], e["rows"], e["dim"], e["batch"]) == (kernel, nloc, DIM, batch):
                            return e["bytes"]
            except (OSError, ValueError, KeyError):
                pass
            return None
        qps_b = B * args.steps / (ms_b / 1e3)
        qps_1 = s1 / (ms_1 / 1e3)
        flops = 2.0 * nloc * DIM * B                       # per launch, this rank's shard
        tf = flops / (scan_b / 1e3) / 1e12
        gbs1 = shard_bytes / (scan_1 / 1e3) / 1e9
        out = {
            "metric": "QPS@top10 (batch 1024)", "value": qps_b, "unit": "queries/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_b / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16",
            "data": "synthetic", "config": workload_config(args, world),
            "kernel": kernel_b, "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": B * args.steps / el_b, "unit": "queries/s", "h2d_bytes_per_step": B * DIM * 2,
                    "d2h_bytes_per_step": B * k * (8 if world == 1 else 12),
                    "via": "zvb200_search (host buffers)" if world == 1 else
                           "pinned H2D + zvb200_search_device + NCCL all-gather + zvb200_merge_topk_device + D2H"},
            "roofline": {"bound": "tensor", "achieved": tf, "peak": pk["tf_sustained"], "unit": "TFLOP/s",
                         "frac": tf / pk["tf_sustained"], "traffic": traffic(kernel_b, B), "kernel": kernel_b,
                         "kernel_ms": scan_b, "peak_source": pk["source"] + ", sustained bf16",
                         "algorithmic": "2*rows*dim*batch flop per launch = %.4g" % flops},
            "batch1": {"value": qps_1, "unit": "queries/s", "steps": s1, "ms_per_step": ms_1 / s1, "kernel": kernel_1,
                       "e2e": {"value": s1 / el_1, "unit": "queries/s", "h2d_bytes_per_step": DIM * 2,
                               "d2h_bytes_per_step": k * (8 if world == 1 else 12)},
                       "roofline": {"bound": "hbm", "achieved": gbs1, "peak": pk["hbm_gbs"], "unit": "GB/s",
                                    "frac": gbs1 / pk["hbm_gbs"], "traffic": traffic(kernel_1, 1), "kernel": kernel_1,
                                    "kernel_ms": scan_1, "peak_source": pk["source"],
                                    "algorithmic": "rows*dim*2 B per launch = %.4g" % shard_bytes}},
        }
        if world == 1 and not args.no_cpu_baseline:
            arm = CpuArm(args)
            out["cpu_baseline"] = arm.best(20.0)
            arm.close()
        emit(out)
    # torch tensors that were used on the context's stream must go before that stream does
    del loc_s, loc_i, all_s, all_i, fin_s, fin_r, h_s, h_r, dq, hq
    small.clear()
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close(); st.close()


def emit(obj):
    """the ONE JSON line of the contract, on the process's real stdout"""
    os.write(_REAL_STDOUT, (json.dumps(obj) + "\n").encode())


# Libraries may chatter on fd 1 (NCCL prints its version banner there when NCCL_DEBUG is set): keep the real
# stdout for the JSON line only and send everything else to stderr.
_REAL_STDOUT = 1


if __name__ == "__main__":
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
