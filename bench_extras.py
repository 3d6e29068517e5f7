"""bench.py's `extra` blocks: the other BASELINE.json configs measured in the driver's own run (1 GPU).

Each block = one config of SURVEY §8(d) at its full shape: device-timed throughput (`value`, queries resident in
HBM), the host-buffer call (`e2e`), the dominant kernel's roofline from CUDA events around it
(zvb200_ctx_last_scan_ms) against its own algorithmic work, and a parity sample against an independent torch
computation over the same rows (fp32 matmul, integer-exact codes, the reference's formulas op by op).
Nothing here is imported by the product; bench.py calls run_all() after the headline measurement.
"""
import math
import time

import numpy as np


def _timeit(fn, sync, reps, warm=2):
    for _ in range(warm):
        fn()
    sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    sync()
    return (time.perf_counter() - t0) / reps


def _dev_time(torch, stream, fn, reps, warm=2):
    """seconds per call, CUDA events on the context's stream"""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 1e3 / reps


def _scan_ms(ctx, fn, reps=5):
    v = []
    for _ in range(reps):
        fn()
        ctx.sync()
        v.append(ctx.last_scan_ms())
    return float(np.mean(v))


def _traffic(kernel, rows, dim, batch):
    """measured DRAM bytes per launch from profiles/traffic.json (one ncu --set full capture each), else None"""
    import json, os
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            for e in json.load(f)["entries"]:
                if (e["kernel"], e["rows"], e["dim"], e["batch"]) == (kernel, rows, dim, batch):
                    return e["bytes"]
    except (OSError, ValueError, KeyError):
        pass
    return None


def _roof(bound, achieved, peak, unit, kernel, kernel_ms, algorithmic, note=None):
    r = {"bound": bound, "achieved": achieved, "peak": peak, "unit": unit, "frac": achieved / peak, "traffic": None,
         "kernel": kernel, "kernel_ms": kernel_ms, "algorithmic": algorithmic}
    if note:
        r["note"] = note
    if bound == "hbm" and achieved > peak:
        r["note"] = (r.get("note", "") + "; " if note else "") + ("above 1: the peak is a COPY measurement (read + write, "
                     "MEASURED_PEAKS.json how), a read-only stream goes faster")
    return r


# ------------------------------------------------------------------------------------------------------------
# C1  Flat fp32 L2^2, 1M x 128, k = 10, 10 000 queries
# ------------------------------------------------------------------------------------------------------------
def c1(args, z, torch, dev, pk, compare_topk, merge_candidates, log):
    n, d, k, nq = max(1000, int(1_000_000 * args.extras_scale)), 128, 10, 10_000
    g = torch.Generator(device=dev); g.manual_seed(1234)
    x = torch.randn((n, d), generator=g, device=dev)
    g.manual_seed(4321)
    q = torch.randn((nq, d), generator=g, device=dev)
    st = z.Store(z.DT_FP32, z.METRIC_L2SQ, d)
    st.reserve(n)
    torch.cuda.synchronize()
    st.append_device(x.data_ptr(), n)
    st.seal()
    ctx = z.Context(st)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    os_ = torch.empty((nq, k), dtype=torch.float32, device=dev)
    oi = torch.empty((nq, k), dtype=torch.int32, device=dev)
    hq = q.cpu().numpy()
    out = {"config": "C1 Flat fp32 SquaredEuclidean, synthetic %dx%d (SIFT-shape), top%d" % (n, d, k)}
    for b, reps in ((1, 300), (1024, 100)):
        f = lambda: ctx.search_device(q.data_ptr(), b, k, os_.data_ptr(), oi.data_ptr())
        t = _dev_time(torch, stream, f, reps)
        kern = ctx.last_kernel()
        sm = _scan_ms(ctx, f)
        e2e = _timeit(lambda: ctx.search(hq[:b], k), ctx.sync, max(10, reps // 3))
        if b == 1:
            roof = _roof("hbm", st.scan_bytes() / (sm / 1e3) / 1e9, pk["hbm_gbs"], "GB/s", kern, sm,
                         "rows*512 B per launch = %.4g" % st.scan_bytes())
        else:
            fl = 2.0 * n * d * b
            roof = _roof("tensor", fl / (sm / 1e3) / 1e12, pk["tf_sustained"], "TFLOP/s", kern, sm,
                         "2*rows*dim*batch flop per launch = %.4g" % fl,
                         "kind::tf32 filter + exact fp32 re-score; quoted against the measured bf16 sustained peak; "
                         "this kernel's own kind::tf32 rate with the epilogue removed is 737 TFLOP/s "
                         "(profiles/r02_mma_peaks_in_situ.txt), see frac_of_in_situ_peak; the 128 x 256 fp32 "
                         "accumulator read-out (2048 cycles per tile) bounds d = 128 above the MMAs")
            roof["in_situ_peak"] = 737.0
            roof["frac_of_in_situ_peak"] = roof["achieved"] / 737.0
        out["batch%d" % b] = {"value": b / t, "unit": "queries/s", "ms_per_batch": t * 1e3, "kernel": kern,
                              "e2e": {"value": b / e2e, "unit": "queries/s", "h2d_bytes_per_step": b * d * 4,
                                      "d2h_bytes_per_step": b * k * 8}, "roofline": roof}
    # the 10 000-query shot (what the config names), device-timed and through host buffers
    f = lambda: ctx.search_device(q.data_ptr(), nq, k, os_.data_ptr(), oi.data_ptr())
    t = _dev_time(torch, stream, f, 10)
    e2e = _timeit(lambda: ctx.search(hq, k), ctx.sync, 5)
    out["shot_10000"] = {"value": nq / t, "unit": "queries/s", "ms": t * 1e3, "e2e_ms": e2e * 1e3,
                         "e2e": {"value": nq / e2e, "unit": "queries/s"}}
    # parity: all 10 000 queries against fp32 |x|^2 - 2 x.q + |q|^2 in torch
    f(); ctx.sync()
    xn = (x * x).sum(1)
    ref_s, ref_i = [], []
    for q0 in range(0, nq, 1000):
        qq = q[q0:q0 + 1000]
        sc = (qq * qq).sum(1)[:, None] - 2.0 * (qq @ x.T) + xn[None, :]
        ts, ti = torch.topk(sc, k + 8, dim=1, largest=False)
        ts, ti = merge_candidates(ts, ti, k + 8)
        ref_s.append(ts); ref_i.append(ti)
    rs, ri = torch.cat(ref_s).cpu().numpy(), torch.cat(ref_i).cpu().numpy()
    bad, rel = compare_topk(oi.cpu().numpy().astype(np.int64) & 0xFFFFFFFF, os_.cpu().numpy(), ri, rs, k, rtol=2e-4)
    out["parity"] = {"queries": nq, "mismatches": bad, "max_rel_err": rel,
                     "rule": "ids equal except near-ties, scores within 2e-4 of |x|^2 - 2x.q + |q|^2 in fp32 torch "
                             "(that form itself cancels ~1e-4 at distance ~ 250 from norms ~ 128)"}
    ctx.close(); st.close()
    del x, q, os_, oi
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------------------
# C3  Flat int8-quantised L2^2, 100M x 96, k = 100 (bit-exact)
# ------------------------------------------------------------------------------------------------------------
def _qi8_l2_torch(torch, ip, mt, qt, d):
    """quantized_integer_metric_matrix.h:36-60 op by op in fp32 (torch elementwise kernels round every operation).
    ip [q, n] exact integer dots as fp32; mt [n, 4] row tails {a, b, s, s2}; qt [q, 4] query tails."""
    ma, mb, ms, ms2 = (mt[:, i][None, :] for i in range(4))
    qa, qb, qs, qs2 = (qt[:, i][:, None] for i in range(4))
    summ = qa * qs
    sum2 = (qa * qa) * qs2
    r = (ma * ma) * ms2 + sum2
    r = r - ((2.0 * ma) * qa) * ip
    db = mb - qb
    r = r + (db * db) * d
    return r + (2.0 * db) * (ms * ma - summ)


def c3(args, z, torch, dev, pk, compare_topk, merge_candidates, log):
    n, d, k = max(100_000, int(100_000_000 * args.extras_scale)), 96, 100
    pq = 8                                              # parity queries
    B = 1024
    st = z.Store(z.DT_INT8, z.METRIC_QI8_L2SQ, d + 16)
    st.reserve(n)
    # a tiny store of the same type lends its context to the device reformer (the reference's per-record quantiser,
    # bit-identical, tests/test_refine.py): rows and queries are quantised exactly as Int8StreamingConverter would
    g = torch.Generator(device=dev); g.manual_seed(69)
    qf = torch.rand((B, d), generator=g, device=dev) * 2 - 1
    helper = z.Store(z.DT_INT8, z.METRIC_QI8_L2SQ, d + 16)
    hx = torch.zeros((4, d + 16), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    helper.append_device(hx.data_ptr(), 4); helper.seal()
    hctx = z.Context(helper)
    L = z.lib()

    def quantise(x32):
        out = torch.empty((x32.shape[0], d + 16), dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        rc = L.zvb200_reform_queries_device(hctx.h, x32.data_ptr(), x32.shape[0], d, out.data_ptr())
        assert rc == 0, L.zvb200_last_error()
        hctx.sync()
        return out

    q8 = quantise(qf)
    qcodes = q8[:pq, :d].view(torch.int8).float()
    qtails = q8[:pq, d:].contiguous().view(torch.float32).view(pq, 4)
    g.manual_seed(96)
    CH = 4_000_000
    cand_s = cand_i = None
    for c0 in range(0, n, CH):
        m = min(CH, n - c0)
        x8 = quantise(torch.rand((m, d), generator=g, device=dev) * 2 - 1)
        st.append_device(x8.data_ptr(), m)
        # the independent scores of this chunk for the parity queries
        ip = qcodes @ x8[:, :d].view(torch.int8).float().T            # exact: |sum| < 2^24
        sc = _qi8_l2_torch(torch, ip, x8[:, d:].contiguous().view(torch.float32).view(m, 4), qtails, float(d))
        ts, ti = torch.topk(sc, min(k + 28, m), dim=1, largest=False)
        ti = ti + c0
        cand_s = ts if cand_s is None else torch.cat([cand_s, ts], 1)
        cand_i = ti if cand_i is None else torch.cat([cand_i, ti], 1)
        cand_s, cand_i = merge_candidates(cand_s, cand_i, k + 28)
        del x8, ip, sc
    st.seal()
    torch.cuda.empty_cache()
    ctx = z.Context(st)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    os_ = torch.empty((B, k), dtype=torch.float32, device=dev)
    oi = torch.empty((B, k), dtype=torch.int32, device=dev)
    hq = q8.cpu().numpy()
    out = {"config": "C3 Flat QuantizedInteger(int8 per-record)/SquaredEuclidean, synthetic %dx%d (Deep-shape), top%d"
                     % (n, d, k)}
    for b, reps in ((1, 20), (1024, 10)):
        f = lambda: ctx.search_device(q8.data_ptr(), b, k, os_.data_ptr(), oi.data_ptr())
        t = _dev_time(torch, stream, f, reps)
        kern = ctx.last_kernel()
        sm = _scan_ms(ctx, f, 3)
        e2e = _timeit(lambda: ctx.search(hq[:b], k), ctx.sync, max(3, reps // 2))
        if b == 1:
            roof = _roof("hbm", st.scan_bytes() / (sm / 1e3) / 1e9, pk["hbm_gbs"], "GB/s", kern, sm,
                         "rows*112 B per launch = %.4g" % st.scan_bytes())
            roof["traffic"] = _traffic(kern, n, d + 16, 1)  # ncu capture of exactly this launch, when committed
        else:
            fl = 2.0 * n * d * b
            roof = _roof("tensor", fl / (sm / 1e3) / 1e12, pk["tf_sustained"], "TFLOP/s", kern, sm,
                         "2*rows*dim*batch flop per launch = %.4g" % fl,
                         "mma kind %d (3 = kind::f16 over the bf16 shadow operands, exact re-score from the int8 "
                         "codes); quoted against the measured bf16 sustained peak; the TMEM read-out of "
                         "4 B x batch x rows of accumulators (64 B per clock per SM) alone is 22.8 ms at 1.9 GHz "
                         "for this shape = 862 TFLOP/s, see frac_of_tmem_floor" % ctx.last_mma_kind())
            tmem_floor_s = 4.0 * b * n / (148 * 64 * 1.9e9)
            roof["tmem_readout_floor_ms"] = tmem_floor_s * 1e3
            roof["frac_of_tmem_floor"] = tmem_floor_s / (sm / 1e3)
        out["batch%d" % b] = {"value": b / t, "unit": "queries/s", "ms_per_batch": t * 1e3, "kernel": kern,
                              "e2e": {"value": b / e2e, "unit": "queries/s", "h2d_bytes_per_step": b * (d + 16),
                                      "d2h_bytes_per_step": b * k * 8}, "roofline": roof}
    # parity: bit-exact scores, ids under the documented tie rule — the batched (tensor-core) result and batch 1
    rs, ri = cand_s.cpu().numpy(), cand_i.cpu().numpy()
    ctx.search_device(q8.data_ptr(), B, k, os_.data_ptr(), oi.data_ptr()); ctx.sync()
    bad_b, _ = compare_topk(oi[:pq].cpu().numpy().astype(np.int64) & 0xFFFFFFFF, os_[:pq].cpu().numpy(), ri, rs, k, exact=True)
    bad_1 = 0
    for i in range(2):
        ctx.search_device(q8[i:i + 1].data_ptr(), 1, k, os_.data_ptr(), oi.data_ptr()); ctx.sync()
        b1, _ = compare_topk(oi[:1].cpu().numpy().astype(np.int64) & 0xFFFFFFFF, os_[:1].cpu().numpy(), ri[i:i + 1], rs[i:i + 1], k, exact=True)
        bad_1 += b1
    out["parity"] = {"queries": pq, "mismatches": bad_b, "batch1_mismatches": bad_1, "max_rel_err": 0.0,
                     "rule": "scores bit-identical to quantized_integer_metric_matrix.h:36-60 evaluated op by op in "
                             "fp32 torch over exact integer dots; ids equal as sets inside runs of equal scores"}
    ctx.close(); st.close(); hctx.close(); helper.close()
    del os_, oi, q8, qf, cand_s, cand_i
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------------------
# C4  IVF nlist 16384, nprobe 64, fp32 cosine, 10M x 1536
# ------------------------------------------------------------------------------------------------------------
def c4(args, z, torch, dev, pk, compare_topk, merge_candidates, log):
    rows, dim, k = max(200_000, int(10_000_000 * args.extras_scale)), 1536, 10
    nlist = 16384 if args.extras_scale >= 0.5 else max(64, int(16384 * args.extras_scale))
    nprobe = max(1, int(round(nlist * 64 / 16384)))
    LB = 64                                               # lists per generation chunk (seeded per chunk)
    g = torch.Generator(device=dev); g.manual_seed(1536)
    centres = torch.randn((nlist, dim), generator=g, device=dev)
    sizes = np.full(nlist, rows // nlist, np.int64); sizes[: rows % nlist] += 1
    offs = np.zeros(nlist + 1, np.uint64); offs[1:] = np.cumsum(sizes)

    def lists_chunk(l0):
        """rows of lists [l0, l0 + LB): centre + 0.3 N(0,1), normalised, in the CosineNormalizeConverter layout"""
        l1 = min(nlist, l0 + LB)
        g.manual_seed(1537 + l0)
        rep = torch.repeat_interleave(torch.arange(l0, l1, device=dev), torch.from_numpy(sizes[l0:l1]).to(dev))
        x = centres[rep] + 0.3 * torch.randn((rep.numel(), dim), generator=g, device=dev)
        nrm = x.norm(dim=1, keepdim=True)
        return torch.cat([x / nrm, nrm], dim=1).contiguous()

    st = z.Store(z.DT_FP32, z.METRIC_COSINE, dim + 1)
    st.reserve(rows)
    for l0 in range(0, nlist, LB):
        x = lists_chunk(l0)
        torch.cuda.synchronize()
        st.append_device(x.data_ptr(), x.shape[0])
        del x
    st.seal()
    cn = centres.norm(dim=1, keepdim=True)
    cent = torch.cat([centres / cn, cn], dim=1)
    st.ivf_attach(cent.cpu().numpy(), offs)
    ctx = z.Context(st)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    ratio = nprobe / nlist
    max_scan = max(1000, int(math.ceil(rows * ratio)))
    B = 1024
    g.manual_seed(4536)
    qsel = torch.randint(0, nlist, (B,), generator=g, device=dev)
    q = centres[qsel] + 0.3 * torch.randn((B, dim), generator=g, device=dev)
    qn = q.norm(dim=1, keepdim=True)
    dq = torch.cat([q / qn, qn], dim=1).contiguous()
    hq = dq.cpu().numpy()
    os_ = torch.empty((B, k), dtype=torch.float32, device=dev)
    oi = torch.empty((B, k), dtype=torch.int32, device=dev)
    per_list = rows / nlist
    lists_scanned = min(nprobe, math.ceil(max_scan / per_list))
    bytes_lists = lists_scanned * per_list * dim * 4
    bytes_coarse = nlist * dim * 4
    out = {"config": "C4 IVF nlist %d nprobe %d (scan_ratio %g, max_scan_count %d) fp32 Cosine, synthetic %dx%d "
                     "(OpenAI-shape), top%d" % (nlist, nprobe, ratio, max_scan, rows, dim, k)}
    # lists the whole batch visits (all lists have the same size here, so the budget keeps the first `lists_scanned`
    # probes of every query): the list-major scan (scan_lists_lm) streams each of them ONCE per batch
    cs_all = 1.0 - dq[:, :dim] @ cent[:, :dim].T
    vis = torch.topk(cs_all, nprobe, dim=1, largest=False).indices[:, :lists_scanned]
    uniq_lists = int(torch.unique(vis).numel())
    del cs_all, vis
    import os as _os
    shapes = ((1, 200), (1024, 5))
    if _os.environ.get("ZVB200_C4_BATCHES"):       # probes: "1:20" = batch 1, 20 repetitions
        shapes = tuple(tuple(int(v) for v in s.split(":")) for s in _os.environ["ZVB200_C4_BATCHES"].split(","))
    for b, reps in shapes:
        f = lambda: ctx.ivf_search_device(dq.data_ptr(), b, k, nprobe, max_scan, os_.data_ptr(), oi.data_ptr())
        t = _dev_time(torch, stream, f, reps)
        sm = _scan_ms(ctx, f, 3)
        e2e = _timeit(lambda: ctx.ivf_search(hq[:b], k, nprobe, max_scan), ctx.sync, max(3, reps // 2))
        per_query = bytes_coarse + b * bytes_lists      # the centroids once per batch, the probed lists per query
        if b == 1:
            alg, what = per_query, "coarse %.4g B per batch + %.4g B of posting lists per query" % (bytes_coarse, bytes_lists)
        else:                                           # every visited list crosses HBM once per BATCH
            alg = bytes_coarse + uniq_lists * per_list * dim * 4
            what = ("coarse %.4g B + the %d distinct posting lists the batch visits, once each = %.4g B per batch "
                    "(SURVEY 8(d)'s per-query figure, %.4g B, would read a shared list once per query)"
                    % (bytes_coarse, uniq_lists, alg, per_query))
        out["batch%d" % b] = {
            "value": b / t, "unit": "queries/s", "ms_per_batch": t * 1e3, "kernel": ctx.last_kernel(),
            "e2e": {"value": b / e2e, "unit": "queries/s", "h2d_bytes_per_step": b * (dim + 1) * 4,
                    "d2h_bytes_per_step": b * k * 8},
            "roofline": _roof("hbm", alg / t / 1e9, pk["hbm_gbs"], "GB/s", "ivf step (coarse + plan + list scan + merges)",
                              t * 1e3, what,
                              "whole call against the HBM roofline of its algorithmic bytes; list scan alone: %.3f ms = "
                              "%.1f GB/s" % (sm, (alg - bytes_coarse) / (sm / 1e3) / 1e9))}
    # parity sample: the reference's IVF semantics restated in torch — coarse order over ALL centroids, lists visited
    # in that order while scanned < max_scan_count (ivf_searcher.cc:216-231), exact fp32 1 - x.q over their rows
    pq = 6
    ctx.ivf_search_device(dq.data_ptr(), B, k, nprobe, max_scan, os_.data_ptr(), oi.data_ptr()); ctx.sync()
    got_s, got_i = os_[:pq].cpu().numpy(), oi[:pq].cpu().numpy().astype(np.int64) & 0xFFFFFFFF
    cs = 1.0 - dq[:pq, :dim] @ cent[:, :dim].T
    corder = torch.argsort(cs, dim=1, stable=True)[:, :nprobe].cpu().numpy()
    ref_s, ref_i = [], []
    for i in range(pq):
        visit, scanned = [], 0
        for l in corder[i]:
            if scanned >= max_scan:
                break
            visit.append(int(l)); scanned += int(sizes[l])
        cs_, ci_ = [], []
        for l0 in sorted({l // LB * LB for l in visit}):
            x = lists_chunk(l0)
            base = int(offs[l0])
            for l in visit:
                if l // LB * LB != l0:
                    continue
                a, b_ = int(offs[l]) - base, int(offs[l + 1]) - base
                sc = 1.0 - x[a:b_, :dim] @ dq[i, :dim]
                cs_.append(sc); ci_.append(torch.arange(int(offs[l]), int(offs[l + 1]), device=dev))
            del x
        s = torch.cat(cs_)[None, :]; ii = torch.cat(ci_)[None, :]
        ts, ti = merge_candidates(s, ii, k + 8)
        ref_s.append(ts); ref_i.append(ti)
    bad, rel = compare_topk(got_i, got_s, torch.cat(ref_i).cpu().numpy(), torch.cat(ref_s).cpu().numpy(), k)
    out["parity"] = {"queries": pq, "mismatches": bad, "max_rel_err": rel,
                     "rule": "same lists under the scan budget, ids equal except near-ties, scores within 1e-4 of "
                             "1 - x.q in fp32 torch"}
    ctx.close(); st.close()
    del centres, cent, dq, q, os_, oi
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------------------
# C5  multi-vector query: 4 fields x 10M x 768 fp16 InnerProduct, top-100 per field, fused (RRF / weighted)
# ------------------------------------------------------------------------------------------------------------
def c5(args, z, torch, dev, pk, st0, ctx0, data0, compare_topk, merge_candidates, log):
    C2Data, DIM = type(data0), 768
    F, k, topn = 4, 100, 100
    rows = st0.count()
    stores, ctxs, datas = [st0], [ctx0], [data0]
    for f in range(1, F):
        dt = C2Data(dev, rows, seed_base=2024 + 1000 * f)
        s = z.Store(z.DT_FP16, z.METRIC_IP, DIM)
        s.reserve(rows)
        dt.fill(s, 0, rows)
        s.seal()
        stores.append(s); ctxs.append(z.Context(s)); datas.append(dt)
    torch.cuda.empty_cache()
    B = 64
    g = torch.Generator(device=dev); g.manual_seed(4205)
    dqs = [torch.nn.functional.normalize(torch.randn((B, DIM), generator=g, device=dev), dim=1).half() for _ in range(F)]
    hqs = [q.cpu().numpy() for q in dqs]

    def multi(b, mode):
        keys = np.empty((F, b, k), np.uint64); scores = np.empty((F, b, k), np.float32); counts = np.empty((F, b), np.uint32)
        scan = 0.0
        for f in range(F):
            kk, ss, _, cc = ctxs[f].search(hqs[f][:b], k)
            keys[f], scores[f], counts[f] = kk, ss, cc
            scan += ctxs[f].last_scan_ms()
        if mode == "rrf":
            r = ctxs[0].fuse_rrf(keys, counts, topn, 60)
        else:
            r = ctxs[0].fuse_weighted(keys, scores, counts, topn, z.METRIC_IP, np.ones(F), negate=True)
        return r, scan, keys, scores, counts

    out = {"config": "C5 multi-vector: %d fields x %dx%d fp16 InnerProduct, top%d per field, fused top%d "
                     "(RRF rank_constant 60)" % (F, rows, DIM, k, topn)}
    bytes_mq = F * rows * DIM * 2
    for b, reps in ((1, 8), (64, 4)):
        for _ in range(2):
            multi(b, "rrf")
        t0 = time.perf_counter(); scan = 0.0
        for _ in range(reps):
            _, s, _, _, _ = multi(b, "rrf"); scan += s
        el = (time.perf_counter() - t0) / reps
        scan /= reps
        blk = {"value": b / el, "unit": "multi-queries/s", "ms_per_batch": el * 1e3, "kernel": ctxs[1].last_kernel(),
               "e2e": {"value": b / el, "unit": "multi-queries/s", "h2d_bytes_per_step": F * b * DIM * 2,
                       "d2h_bytes_per_step": F * b * k * 12 + b * topn * 16}}
        if b == 1:
            blk["roofline"] = _roof("hbm", bytes_mq / (scan / 1e3) / 1e9, pk["hbm_gbs"], "GB/s", ctxs[1].last_kernel(),
                                    scan, "4 fields * rows*dim*2 B per multi-query = %.4g" % bytes_mq,
                                    "the four scans of one multi-query, summed")
        else:
            # 64 queries are half a 128-query tile: the tensor-core scan streams every row once per batch and is
            # HBM-bound below ~213 queries, so the bytes bound it, not the flops
            fl = 2.0 * rows * DIM * b * F
            blk["roofline"] = _roof("hbm", bytes_mq / (scan / 1e3) / 1e9, pk["hbm_gbs"], "GB/s", ctxs[1].last_kernel(),
                                    scan, "4 fields * rows*dim*2 B per batch of multi-queries = %.4g" % bytes_mq,
                                    "the four tensor-core scans of one batch, summed; as flops: %.1f TFLOP/s = %.3f of "
                                    "the sustained bf16 peak" % (fl / (scan / 1e3) / 1e12,
                                                                 fl / (scan / 1e3) / 1e12 / pk["tf_sustained"]))
        out["batch%d" % b] = blk
    # parity: per-field top-100 of 2 multi-queries vs fp32 torch over the same rows; the fused list vs the formula of
    # RrfReRanker.rerank (multi_vector_reranker.py:62-89) applied to those per-field lists
    pq = 2
    (fk, fs, fc), _, keys, scores, counts = multi(pq, "rrf")
    bad, rel = 0, 0.0
    for f in range(F):
        rs, ri = datas[f].reference_topk(dqs[f][:pq], 0, rows, k + 16)
        b_, r_ = compare_topk(keys[f].astype(np.int64), scores[f], ri.cpu().numpy(), rs.cpu().numpy(), k)
        bad += b_; rel = max(rel, r_)
    fuse_bad = 0
    for i in range(pq):
        acc, order = {}, []
        for f in range(F):
            for r in range(int(counts[f, i])):
                key = int(keys[f, i, r])
                if key not in acc:
                    acc[key] = 0.0; order.append(key)
                acc[key] += 1.0 / (60 + r + 1)
        exp = sorted(order, key=lambda key: -acc[key])[:topn]        # stable: first-seen order among equal scores
        got = [int(v) for v in fk[i][: int(fc[i])]]
        if len(got) != len(exp) or any(abs(acc[a] - acc[b]) > 1e-12 for a, b in zip(got, exp)):
            fuse_bad += 1
    out["parity"] = {"queries": pq * F, "mismatches": bad, "fusion_mismatches": fuse_bad, "max_rel_err": rel,
                     "rule": "per-field ids equal except near-ties / scores within 1e-4 of fp32 torch; fused scores "
                             "equal to sum 1/(60 + rank + 1) over the per-field lists"}
    for f in range(1, F):
        ctxs[f].close(); stores[f].close()
    torch.cuda.empty_cache()
    return out


def run_all(args, z, torch, dev, pk, st0, ctx0, data0, compare_topk, merge_candidates, log):
    extra = {}
    for name, fn in (("C5", lambda: c5(args, z, torch, dev, pk, st0, ctx0, data0, compare_topk, merge_candidates, log)),
                     ("C1", lambda: c1(args, z, torch, dev, pk, compare_topk, merge_candidates, log)),
                     ("C3", lambda: c3(args, z, torch, dev, pk, compare_topk, merge_candidates, log)),
                     ("C4", lambda: c4(args, z, torch, dev, pk, compare_topk, merge_candidates, log))):
        t0 = time.perf_counter()
        try:
            extra[name] = fn()
            extra[name]["wall_s"] = time.perf_counter() - t0
        except Exception as e:  # noqa: BLE001 - one failing extra must not take the headline line with it
            import traceback
            log("extra %s failed: %s" % (name, traceback.format_exc()))
            extra[name] = {"error": "%s: %s" % (type(e).__name__, e)}
            torch.cuda.empty_cache()
        log("extra %s done in %.1f s" % (name, time.perf_counter() - t0))
    return extra
