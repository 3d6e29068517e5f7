#!/usr/bin/env python3
"""bench.py — QPS@top10 of the exhaustive scan on BASELINE.json configs[1]:
Flat fp16 inner product over synthetic 10M x 768 (Cohere-shape), batch 1024 (headline) and batch 1, rows sharded
over N GPUs of one node (strong scaling: the 10M rows are divided), per-GPU top-k exchanged and merged INSIDE the
library (zvb200_comm_* / zvb200_search_sharded_device: peer mailboxes over NVLink, or one ncclAllGather).

  python bench.py --gpus N --steps K --warmup W            # our CUDA path (torchrun for N>1)
  python bench.py --impl reference --gpus N ...            # the reference's CPU path, host cores

A "step" is `batches_per_step` searches of one batch each (declared in config; 1 when a batch is long enough, more
when the shards are small, so that the timed region lasts >= ~1.5 s at every N and the GPU reaches its sustained
power state).  `value` is whole-job QPS with queries already in HBM, timed with CUDA events on the stream the
kernels run on, max over ranks.  `e2e` is the same metric through the host-buffer call (H2D of the queries and D2H
of the results inside the timed region).  `parity` compares the product's final ids / scores for ALL queries of the
timed batch, at the timed launch geometry and N, with an independent fp32 torch.matmul + topk over the same rows;
a mismatch exits non-zero.  At N = 1 the line also carries `extra`: driver-measured blocks for the other BASELINE
configs (C1 fp32 L2 1Mx128, C3 int8-quantised L2 100Mx96 top-100, C4 IVF 10Mx1536, C5 multi-vector), each with its
own roofline and parity sample.  PyTorch is only plumbing here (device memory for synthetic data, the independent
check, torch.distributed for the rendezvous); every kernel in a timed region is from libzvec_b200.so.
"""
import argparse
import json
import math
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
    ap.add_argument("--no-extras", action="store_true", help="skip the C1 / C3 / C4 / C5 blocks (N = 1 only anyway)")
    ap.add_argument("--extras-scale", type=float, default=1.0, help="row-count factor of the extra configs (testing)")
    ap.add_argument("--cpu-sample-rows", type=int, default=1_000_000, help="rows of the cpu_baseline sample (1/10)")
    ap.add_argument("--target-seconds", type=float, default=1.5, help="minimum length of a timed region")
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


def emit(obj):
    """the ONE JSON line of the contract, on the process's real stdout"""
    os.write(_REAL_STDOUT, (json.dumps(obj) + "\n").encode())


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# Libraries may chatter on fd 1 (NCCL prints its version banner there when NCCL_DEBUG is set): keep the real
# stdout for the JSON line only and send everything else to stderr.
_REAL_STDOUT = 1


def workload_config(args, n, batches_per_step=1):
    return {"workload": "Flat fp16 InnerProduct, synthetic %dx%d (Cohere-shape), top%d, batch %d (headline) and "
                        "batch 1; rows sharded over %d GPU(s)" % (args.rows, DIM, args.topk, args.batch, n),
            "rows": args.rows, "dim": DIM, "topk": args.topk, "batch": args.batch, "shards": n,
            "batches_per_step": batches_per_step,
            "l2_policy": "inputs larger than L2: every batch streams the whole %.1f GB shard (L2 is 126 MB)" %
                         (args.rows / n * DIM * 2 / 1e9)}


# -----------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own FlatSearcher on the host cores
# -----------------------------------------------------------------------------------------
def host_queries_fp16(n):
    rng = np.random.default_rng(SEED_QUERY)
    q = rng.standard_normal((n, DIM)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q.astype(np.float16)


def host_rows_fp16(rows):
    """the workload's rows on the host: N(0,1) row-normalised in fp32, then fp16, generated in 1M-row
    chunks with torch's multi-threaded CPU generator (numpy's would take a minute for 7.7e9 samples)"""
    import torch
    out = np.empty((rows, DIM), np.float16)
    g = torch.Generator(); g.manual_seed(SEED_BASE)
    step = 1_000_000
    for c0 in range(0, rows, step):
        n = min(step, rows - c0)
        x = torch.randn((n, DIM), generator=g, dtype=torch.float32)
        x = torch.nn.functional.normalize(x, dim=1).to(torch.float16)
        out[c0:c0 + n] = x.numpy()
    return out


class CpuArm:
    """The reference's CPU implementation of the path on this box's host cores: its own FlatSearcher over the
    stored fp16 rows (oracle/_ref, compiled from the reference's sources; the index lives in memory through the
    reference's MemoryDumper / MemoryReadStorage), driven like tools/core/bench_original.cc: T threads, one
    context each, one search_impl(q, qmeta, b, ctx) call per thread and step."""

    def __init__(self, args, rows):
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        os.environ["ZREF_MEMORY_INDEX"] = "1"
        os.environ["ZREF_NO_STORED_COPY"] = "1"
        import support as S
        self.S, self.args = S, args
        self.cores = os.cpu_count() or 1
        self.rows = rows
        t0 = time.perf_counter()
        x = host_rows_fp16(rows)
        if S.ref() is not None:
            self.kind, self.lib = "reference", S.ref()._name
            self.ri = S.RefIndex(x, S.DT_FP16, DIM, "InnerProduct", "")
            self.x = x          # the holder reads the caller's memory during the build only; keep it for safety
        else:  # the scalar restatement (1 thread)
            self.kind, self.cores, self.lib, self.ri = "port", 1, "libzvec_oracle.so", None
            self.xs = x.view(np.uint8).reshape(rows, -1)
        self.build_s = time.perf_counter() - t0
        self.qs = host_queries_fp16(self.cores * 8)

    def step(self, b):
        """T threads x one search_impl call of b queries -> (queries done, seconds)"""
        S, args = self.S, self.args
        nq = self.cores * b
        q = self.qs[:nq]
        t0 = time.perf_counter()
        if self.ri is not None:
            self.ri.search(q, args.topk, batch=b, threads=self.cores)
        else:
            S.oracle_flat_search(S.M_IP, S.DT_FP16, self.xs, DIM, q.view(np.uint8).reshape(nq, -1), args.topk)
        return nq, time.perf_counter() - t0

    def describe(self, b, steps, seconds, queries):
        scaled = self.rows != self.args.rows
        return ("%s FlatSearcher (row-major fp16 index in memory, %d threads x one search_impl call of %d quer%s per "
                "step) on %s%d of %d rows: %d steps, %d queries, %.1f s%s" %
                (self.lib, self.cores, b, "y" if b == 1 else "ies", "a slice of " if scaled else "ALL ", self.rows,
                 self.args.rows, steps, queries, seconds,
                 "; QPS scaled linearly by rows (a flat scan is linear in N)" if scaled else ""))

    def close(self):
        if self.ri is not None:
            self.ri.close()


def host_ram_ok(rows):
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        return True
    return avail > rows * DIM * 2 * 2.2 + (8 << 30)     # the rows + the index's copy + slack


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rows = args.rows
    if not host_ram_ok(rows):
        rows = max(args.rows // 10, 1)                   # a 1/10 slice, flagged "scaled"
    arm = CpuArm(args, rows)
    log("reference arm: index of %d rows built in %.1f s" % (rows, arm.build_s))
    # warm-up steps double as the shape calibration: 1, 2, 4 queries per search_impl call; the timed steps use the
    # shape with the best QPS whose step stays below ~6 s
    shapes, best_b, best_qps = [1, 2, 4], 1, 0.0
    for i in range(max(args.warmup, 1)):
        b = shapes[i] if i < len(shapes) else best_b
        nq, el = arm.step(b)
        qps = nq / el
        log("reference warm-up: %d queries per call -> %.2f QPS (%.2f s)" % (b, qps, el))
        if qps > best_qps and (el <= 6.0 or b == 1):
            best_b, best_qps = b, qps
    done, total = 0, 0.0
    for _ in range(args.steps):
        nq, el = arm.step(best_b)
        done += nq; total += el
    scale = rows / args.rows
    v = done / total * scale
    base = {"value": v, "unit": "queries/s", "cores": arm.cores, "kind": arm.kind,
            "sample": arm.describe(best_b, args.steps, total, done), "scaled": rows != args.rows,
            "scale_factor": args.rows / rows, "queries_per_step": arm.cores * best_b}
    arm.close()
    out = {"impl": "reference", "metric": "QPS@top10 (batch 1024)", "value": v, "unit": "queries/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": 1000.0 * total / args.steps, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f16", "data": "synthetic",
           "config": workload_config(args, args.gpus), "cpu_baseline": base,
           "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    emit(out)


# -----------------------------------------------------------------------------------------
# the independent check: fp32 torch.matmul + topk over the same rows, compared under the documented rule
# -----------------------------------------------------------------------------------------
def merge_candidates(scores, ids, k):
    """[q, m] candidate (score, global id) -> the k best per query by (score, id) ascending"""
    import torch
    # stable two-key sort: by id first, then (stable) by score
    o = torch.argsort(ids, dim=1, stable=True)
    s1, i1 = torch.gather(scores, 1, o), torch.gather(ids, 1, o)
    o = torch.argsort(s1, dim=1, stable=True)
    return torch.gather(s1, 1, o)[:, :k].contiguous(), torch.gather(i1, 1, o)[:, :k].contiguous()


def compare_topk(got_ids, got_scores, ref_ids, ref_scores, k, rtol=1e-4, exact=False):
    """The parity rule of DESIGN.md §3 on host arrays.  ref_* hold MORE than k entries per query (k + margin, sorted by
    (score, id)) so that a near-tie straddling position k can be recognised.
    fp: score j within rtol of the reference's score j; an id that differs from the reference's id at its position
    must be among the reference's candidates with a reference score within rtol of the score it was returned with.
    exact: scores bit-identical, ids equal as sets inside every run of equal scores (the last run may be cut).
    -> (mismatching queries, max relative score error)"""
    got_ids = np.asarray(got_ids).astype(np.int64); ref_ids = np.asarray(ref_ids).astype(np.int64)
    got_scores = np.asarray(got_scores, np.float32); ref_scores = np.asarray(ref_scores, np.float32)
    bad, max_rel = 0, 0.0
    for i in range(got_ids.shape[0]):
        gs, gi, rs, ri = got_scores[i][:k], got_ids[i][:k], ref_scores[i], ref_ids[i]
        denom = np.maximum(np.abs(rs[:k]), 1e-6)
        rel = np.abs(gs - rs[:k]) / denom
        ok = True
        if exact:
            ok = np.array_equal(gs.view(np.uint32), rs[:k].view(np.uint32))
            if ok and not np.array_equal(gi, ri[:k]):
                last = rs[k - 1]
                for s in np.unique(rs[:k]):
                    a, b = set(gi[gs == s].tolist()), set(ri[rs == s].tolist())     # b: the WHOLE run, beyond k too
                    ok = ok and (a == b if s != last else a <= b)
        else:
            max_rel = max(max_rel, float(rel.max()) if rel.size else 0.0)
            ok = bool(np.all(rel <= rtol + 1e-7))
            if ok and not np.array_equal(gi, ri[:k]):
                pos = {int(r): j for j, r in enumerate(ri)}
                for j in np.nonzero(gi != ri[:k])[0]:
                    p = pos.get(int(gi[j]))
                    if p is None or abs(float(rs[p]) - float(gs[j])) > rtol * max(abs(float(gs[j])), 1e-6) + 1e-7:
                        ok = False
                        break
        bad += 0 if ok else 1
    return bad, max_rel


class C2Data:
    """The C2 rows, chunk by chunk on the device: data depends on the global row only (seed per 500k-row chunk)."""

    def __init__(self, dev, rows, seed_base=SEED_BASE):
        import torch
        self.torch, self.dev, self.rows, self.seed_base = torch, dev, rows, seed_base
        self.g = torch.Generator(device=dev)

    def chunk(self, c0):
        torch = self.torch
        self.g.manual_seed(self.seed_base + c0 // CHUNK)
        n = min(CHUNK, self.rows - c0)
        x = torch.randn((n, DIM), generator=self.g, device=self.dev, dtype=torch.float32)
        return torch.nn.functional.normalize(x, dim=1).to(torch.float16)

    def fill(self, st, lo, hi):
        torch = self.torch
        c0 = (lo // CHUNK) * CHUNK
        while c0 < hi:
            x = self.chunk(c0)
            a, b = max(lo, c0) - c0, min(hi, c0 + x.shape[0]) - c0
            xs = x[a:b].contiguous()
            torch.cuda.synchronize()
            st.append_device(xs.data_ptr(), b - a)
            del x, xs
            c0 += CHUNK

    def reference_topk(self, q16, lo, hi, m):
        """-(q . x) in fp32 (no TF32) over rows [lo, hi): the m best per query as (scores, global ids)"""
        torch = self.torch
        q32 = q16.float()
        best_s = best_i = None
        c0 = (lo // CHUNK) * CHUNK
        while c0 < hi:
            x = self.chunk(c0)
            a, b = max(lo, c0) - c0, min(hi, c0 + x.shape[0]) - c0
            for s0 in range(a, b, 250_000):
                s1 = min(b, s0 + 250_000)
                sc = -(q32 @ x[s0:s1].float().T)
                mm = min(m, s1 - s0)
                ts, ti = torch.topk(sc, mm, dim=1, largest=False)
                ti = ti + (c0 + s0)
                best_s = ts if best_s is None else torch.cat([best_s, ts], 1)
                best_i = ti if best_i is None else torch.cat([best_i, ti], 1)
                if best_s.shape[1] > 8 * m:
                    best_s, best_i = merge_candidates(best_s, best_i, m)
                del sc
            del x
            c0 += CHUNK
        if best_s is None:
            return (torch.full((q16.shape[0], m), float("inf"), device=self.dev),
                    torch.full((q16.shape[0], m), -1, dtype=torch.int64, device=self.dev))
        s, i = merge_candidates(best_s, best_i, m)
        if s.shape[1] < m:
            pad = m - s.shape[1]
            s = torch.cat([s, torch.full((s.shape[0], pad), float("inf"), device=self.dev)], 1)
            i = torch.cat([i, torch.full((i.shape[0], pad), -1, dtype=torch.int64, device=self.dev)], 1)
        return s, i


def run_ours(args):
    import torch
    import torch.distributed as dist
    import zvec_b200 as z

    torch.backends.cuda.matmul.allow_tf32 = False        # the independent check multiplies in true fp32
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
    L = z.lib()

    # ---- the shard: rows [lo, hi) of the global matrix, generated chunk by chunk on the device
    from zvec_b200 import sharding
    lo, hi = sharding.shard_bounds(args.rows, world)[rank]
    data = C2Data(dev, args.rows)
    st = z.Store(z.DT_FP16, z.METRIC_IP, DIM, device=local)
    st.reserve(hi - lo)
    data.fill(st, lo, hi)
    st.seal()
    torch.cuda.empty_cache()
    nloc = st.count()
    shard_bytes = st.scan_bytes()

    hq = torch.from_numpy(host_queries_fp16(args.batch)).pin_memory()
    dq = hq.to(dev)
    k, B = args.topk, args.batch
    ctx = z.Context(st)
    ctx.set_kernel(args.kernel)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    loc_s = torch.empty((B, k), dtype=torch.float32, device=dev)
    loc_i = torch.empty((B, k), dtype=torch.int32, device=dev)
    fin_s = torch.empty((B, k), dtype=torch.float32, device=dev)
    fin_r = torch.empty((B, k), dtype=torch.int64, device=dev)
    h_s = torch.empty((B, k), dtype=torch.float32).pin_memory()
    h_r = torch.empty((B, k), dtype=torch.int64).pin_memory()
    bases = sharding.shard_bases(args.rows, world)
    comm = None
    if world > 1:
        # the multi-GPU layer lives in the library: torch.distributed only carries the 128-byte id (and the barriers)
        uid = torch.from_numpy(z.comm_unique_id() if rank == 0 else np.zeros(z.binding.COMM_ID_BYTES, np.uint8)).to(dev)
        dist.broadcast(uid, 0)
        comm = z.Comm(ctx, rank, world, uid.cpu().numpy(), bases)
        log("rank %d: exchange transport = %s" % (rank, comm.transport()))

    def step_q(q):
        """one batch of q queries already in HBM -> final top-k in HBM (all on ctx's stream)"""
        if world == 1:
            ctx.search_device(dq.data_ptr(), q, k, loc_s.data_ptr(), loc_i.data_ptr())
        else:
            ctx.search_sharded_device(comm, dq.data_ptr(), q, k, fin_s.data_ptr(), fin_r.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(v):
        if world == 1:
            return v
        t = torch.tensor([v], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(q, steps, warmup, reps):
        for _ in range(warmup):
            for _ in range(reps):
                step_q(q)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = L.zvb200_kernel_launches()
        e0.record(stream)
        for _ in range(steps * reps):
            step_q(q)
        e1.record(stream)
        barrier()
        return allmax(e0.elapsed_time(e1)), L.zvb200_kernel_launches() - l0

    def calibrate(q, steps):
        """batches per step so that the timed region lasts >= target seconds (same value on every rank)"""
        for _ in range(2):
            step_q(q)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(3):
            step_q(q)
        e1.record(stream)
        barrier()
        per = allmax(e0.elapsed_time(e1)) / 3.0
        return max(1, int(math.ceil(args.target_seconds * 1e3 / (per * steps))))

    reps_b = calibrate(B, args.steps)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_b, launches = timed(B, args.steps, args.warmup, reps_b)
    clocks = sampler.stop() if rank == 0 else None
    kernel_b = ctx.last_kernel()
    # dominant-kernel duration, measured live with CUDA events around the scan kernel only, IN the power state of the
    # timed region: the last batch of the timed region itself, then the last batch of a few more back-to-back bursts
    # (a sync after every single batch lets the clocks recover and flatters the kernel by ~8 %)
    scan_ms = [ctx.last_scan_ms()]
    for _ in range(4):
        for _ in range(max(3, reps_b)):
            step_q(B)
        ctx.sync()
        scan_ms.append(ctx.last_scan_ms())
    scan_b = allmax(float(np.mean(scan_ms)))
    # batch 1
    s1 = max(args.steps, 30)
    reps_1 = calibrate(1, s1)
    ms_1, _ = timed(1, s1, max(args.warmup, 3), reps_1)
    kernel_1 = ctx.last_kernel()
    scan_ms = [ctx.last_scan_ms()]
    for _ in range(4):
        for _ in range(max(3, reps_1)):
            step_q(1)
        ctx.sync()
        scan_ms.append(ctx.last_scan_ms())
    scan_1 = allmax(float(np.mean(scan_ms)))

    # ---- e2e: the call a user makes — host buffers in, host results out
    def e2e_step(q):
        if world == 1:
            return ctx.search(hq.numpy()[:q], k)        # zvb200_search: H2D + kernels + D2H + sync
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
        return allmax(time.perf_counter() - t0)

    n_e2e_b, n_e2e_1 = args.steps * reps_b, s1 * min(reps_1, 20)
    el_b = e2e_timed(B, n_e2e_b)
    el_1 = e2e_timed(1, n_e2e_1)

    # ---- parity at the timed geometry: every query of the batch, this N, against fp32 torch.matmul + topk
    margin = 8
    step_q(B)
    ctx.sync()
    if world == 1:
        got_s, got_r = loc_s.cpu().numpy(), loc_i.cpu().numpy().astype(np.int64) & 0xFFFFFFFF
    else:
        got_s, got_r = fin_s.cpu().numpy(), fin_r.cpu().numpy()
        comm.status()
    rs, ri = data.reference_topk(dq, lo, hi, k + margin)
    if world > 1:
        all_s = [torch.empty_like(rs) for _ in range(world)]
        all_i = [torch.empty_like(ri) for _ in range(world)]
        dist.all_gather(all_s, rs); dist.all_gather(all_i, ri)
        rs, ri = merge_candidates(torch.cat(all_s, 1), torch.cat(all_i, 1), k + margin)
    parity = None
    if rank == 0:
        bad, max_rel = compare_topk(got_r, got_s, ri.cpu().numpy(), rs.cpu().numpy(), k)
        parity = {"queries": B, "mismatches": bad, "max_rel_err": max_rel, "rule": "ids equal except near-ties "
                  "(< 1e-4 relative), scores within 1e-4 relative of an fp32 torch.matmul + topk over the same rows, "
                  "merged over all %d shard(s) by (score, global row)" % world, "topk": k}
        # batch 1 too (the streaming kernel): the first query alone
    step_q(1)
    ctx.sync()
    if rank == 0:
        g1s = (loc_s if world == 1 else fin_s)[:1].cpu().numpy()
        g1r = ((loc_i[:1].cpu().numpy().astype(np.int64) & 0xFFFFFFFF) if world == 1 else fin_r[:1].cpu().numpy())
        bad1, rel1 = compare_topk(g1r, g1s, ri[:1].cpu().numpy(), rs[:1].cpu().numpy(), k)
        parity["batch1_mismatches"] = bad1
        parity["max_rel_err"] = max(parity["max_rel_err"], rel1)
    del rs, ri

    out = None
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
        qps_b = B * args.steps * reps_b / (ms_b / 1e3)
        qps_1 = s1 * reps_1 / (ms_1 / 1e3)
        flops = 2.0 * nloc * DIM * B                       # per launch, this rank's shard
        tf = flops / (scan_b / 1e3) / 1e12
        gbs1 = shard_bytes / (scan_1 / 1e3) / 1e9
        step_ms = ms_b / (args.steps * reps_b)
        out = {
            "metric": "QPS@top10 (batch 1024)", "value": qps_b, "unit": "queries/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_b / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16",
            "data": "synthetic", "config": workload_config(args, world, reps_b),
            "kernel": kernel_b, "clocks": clocks, "gpu_launches": int(launches),
            "ms_per_batch": step_ms, "ms_outside_scan_per_batch": step_ms - scan_b,
            "exchange": comm.transport() if comm else "none (1 GPU)",
            "parity": parity,
            "e2e": {"value": B * n_e2e_b / el_b, "unit": "queries/s", "h2d_bytes_per_step": B * DIM * 2 * reps_b,
                    "d2h_bytes_per_step": B * k * (8 if world == 1 else 12) * reps_b,
                    "via": "zvb200_search (host buffers)" if world == 1 else
                           "pinned H2D + zvb200_search_sharded_device (scan, exchange, merge inside the library) + D2H"},
            "roofline": {"bound": "tensor", "achieved": tf, "peak": pk["tf_sustained"], "unit": "TFLOP/s",
                         "frac": tf / pk["tf_sustained"], "traffic": traffic(kernel_b, B), "kernel": kernel_b,
                         "kernel_ms": scan_b, "peak_source": pk["source"] + ", sustained bf16 (the kernel is timed "
                         "inside a >= 1.5 s region); burst peak %.1f" % pk["tf_burst"],
                         "frac_of_burst": tf / pk["tf_burst"],
                         "algorithmic": "2*rows*dim*batch flop per launch = %.4g" % flops},
            "batch1": {"value": qps_1, "unit": "queries/s", "steps": s1, "batches_per_step": reps_1,
                       "ms_per_step": ms_1 / s1, "ms_per_batch": ms_1 / (s1 * reps_1), "kernel": kernel_1,
                       "e2e": {"value": n_e2e_1 / el_1, "unit": "queries/s", "h2d_bytes_per_step": DIM * 2,
                               "d2h_bytes_per_step": k * (8 if world == 1 else 12)},
                       "roofline": {"bound": "hbm", "achieved": gbs1, "peak": pk["hbm_gbs"], "unit": "GB/s",
                                    "frac": gbs1 / pk["hbm_gbs"], "traffic": traffic(kernel_1, 1), "kernel": kernel_1,
                                    "kernel_ms": scan_1, "peak_source": pk["source"] + (
                                        "; above 1: that peak is a COPY measurement (read + write), a read-only stream goes faster"
                                        if gbs1 > pk["hbm_gbs"] else ""),
                                    "algorithmic": "rows*dim*2 B per launch = %.4g" % shard_bytes}},
        }
    # ---- the other BASELINE configs, driver-measured (1 GPU): C5 reuses the C2 store as its first field
    if world == 1 and not args.no_extras:
        import bench_extras
        out["extra"] = bench_extras.run_all(args, z, torch, dev, pk, st, ctx, data, compare_topk, merge_candidates, log)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # free the GPU side first: the CPU leg needs the host's memory bandwidth, not ours
        arm_rows = min(args.cpu_sample_rows, args.rows)
        arm = CpuArm(args, arm_rows)
        arm.step(1)
        done, total, t_end = 0, 0.0, time.perf_counter() + 12.0
        steps = 0
        while time.perf_counter() < t_end or steps < 2:
            nq, el = arm.step(1)
            done += nq; total += el; steps += 1
        out["cpu_baseline"] = {"value": done / total * arm_rows / args.rows, "unit": "queries/s", "cores": arm.cores,
                               "kind": arm.kind, "sample": arm.describe(1, steps, total, done),
                               "scaled": arm_rows != args.rows, "scale_factor": args.rows / arm_rows,
                               "qps_on_sample": done / total}
        arm.close()
    if rank == 0:
        emit(out)
    # torch tensors that were used on the context's stream must go before that stream does
    del loc_s, loc_i, fin_s, fin_r, h_s, h_r, dq, hq
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        comm.close()
        dist.barrier()
        dist.destroy_process_group()
    ctx.close(); st.close()
    if rank == 0:
        bad = [("C2", parity)] if parity and (parity["mismatches"] or parity.get("batch1_mismatches")) else []
        for name, blk in (out.get("extra") or {}).items():
            p = blk.get("parity") or {}
            if p.get("mismatches") or p.get("batch1_mismatches") or p.get("fusion_mismatches"):
                bad.append((name, p))
        if bad:
            log("PARITY MISMATCH: %s" % json.dumps(bad))
            sys.exit(3)


if __name__ == "__main__":
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
