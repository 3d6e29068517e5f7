"""Multi-GPU layer, host logic: world_size-2 gloo processes on CPU (the oracle stands in for the
per-shard GPU scan; the exchange and merge code is the product's zvec_b200/sharding.py), plus a
GPU check of merge_shards_kernel against the numpy statement of the same merge."""
import os
import sys

import numpy as np
import pytest

import support as S
from zvec_b200 import sharding


def test_shard_bounds_cover_rows_once():
    for rows in (0, 1, 7, 8, 9, 1000, 10_000_000):
        for world in (1, 2, 3, 4, 8):
            b = sharding.shard_bounds(rows, world)
            assert len(b) == world and b[0][0] == 0 and b[-1][1] == rows
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            assert sharding.shard_bases(rows, world) == [lo for lo, _ in b]


def test_merge_topk_host_small():
    s = np.array([[[1.0, 3.0, np.nan]], [[1.0, 2.0, 5.0]]], np.float32)          # g=2, q=1, k=3
    ix = np.array([[[4, 0, 0xFFFFFFFF]], [[0, 1, 2]]], np.uint32)
    os_, or_ = sharding.merge_topk_host(s, ix, [0, 10], 4)
    assert list(os_[0]) == [1.0, 1.0, 2.0, 3.0] and list(or_[0]) == [4, 10, 11, 0]   # tie -> lower global row


def _worker(rank, world, port, n, d, q, k, kind, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(99)
        x = rng.standard_normal((n, d)).astype(np.float32)
        qs = rng.standard_normal((q, d)).astype(np.float32)
        if kind == "i8":   # heavy ties: the global row tie-break must hold across shards
            xb = np.clip(np.round(x * 3), -127, 127).astype(np.int8).view(np.uint8)
            qb = np.clip(np.round(qs * 3), -127, 127).astype(np.int8).view(np.uint8)
            metric, dt = S.M_L2SQ, S.DT_INT8
        else:
            xb, qb = x.astype(np.float16).view(np.uint8).reshape(n, -1), qs.astype(np.float16).view(np.uint8).reshape(q, -1)
            metric, dt = S.M_IP, S.DT_FP16
        lo, hi = sharding.shard_bounds(n, world)[rank]
        _, ls, li, _ = S.oracle_flat_search(metric, dt, xb[lo:hi], d, qb, k) if hi > lo else (
            None, np.full((q, k), np.nan, np.float32), np.full((q, k), 0xFFFFFFFF, np.uint32), None)
        all_s, all_i = sharding.gather_topk(torch.from_numpy(ls), torch.from_numpy(li.astype(np.int32)), world)
        ms, mr = sharding.merge_topk_host(all_s.numpy(), all_i.numpy(), sharding.shard_bases(n, world), k)
        if rank == 0:
            _, es, ei, _ = S.oracle_flat_search(metric, dt, xb, d, qb, k)
            ok = True
            for i in range(q):
                alld = S.oracle_distances(metric, dt, xb, d, qb[i])
                try:
                    S.check_topk_parity(mr[i].astype(np.uint32), ms[i], ei[i], es[i], all_dist=alld,
                                        exact=dt == S.DT_INT8, what="sharded q%d" % i)
                except AssertionError as e:
                    ok = False
                    out.put(str(e))
            out.put("ok" if ok else "fail")
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind,n", [("fp16", 3001), ("i8", 2000), ("fp16", 1)])
def test_two_rank_gloo_sharded_search_equals_unsharded(kind, n):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29600 + (os.getpid() % 300)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, 24, 5, 10, kind, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    msgs = []
    while not out.empty():
        msgs.append(out.get())
    assert msgs and msgs[-1] == "ok", msgs


@pytest.mark.gpu
@pytest.mark.parametrize("g,q,k", [(2, 5, 10), (8, 33, 100), (3, 1, 1), (4, 1024, 10)])
def test_merge_shards_kernel_equals_host_merge(g, q, k):
    import torch
    import zvec_b200 as z
    rng = np.random.default_rng(g * 1000 + k)
    s = np.sort(np.round(rng.standard_normal((g, q, k)) * 4).astype(np.float32), axis=2)   # many ties
    ix = rng.integers(0, 1000, (g, q, k)).astype(np.uint32)
    # rows inside a run of equal scores ascend, like every scan kernel emits them
    for a in range(g):
        for b in range(q):
            o = np.lexsort((ix[a, b], s[a, b])); s[a, b], ix[a, b] = s[a, b][o], ix[a, b][o]
    cut = rng.integers(0, k + 1, (g, q))
    for a in range(g):
        for b in range(q):
            s[a, b, cut[a, b]:] = np.nan; ix[a, b, cut[a, b]:] = 0xFFFFFFFF
    bases = [i * 1000 for i in range(g)]
    st = z.Store(z.DT_FP32, z.METRIC_L2SQ, 4); st.append(np.zeros((1, 4), np.float32)); st.seal()
    ctx = z.Context(st)
    ds, di = torch.from_numpy(s).cuda(), torch.from_numpy(ix.astype(np.int32)).cuda()
    os_ = torch.empty((q, k), dtype=torch.float32, device="cuda")
    or_ = torch.empty((q, k), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    ctx.merge_topk_device(ds.data_ptr(), di.data_ptr(), g, bases, q, k, os_.data_ptr(), or_.data_ptr())
    ctx.sync()
    es, er = sharding.merge_topk_host(s, ix, bases, k)
    assert np.array_equal(os_.cpu().numpy(), es, equal_nan=True)
    assert np.array_equal(or_.cpu().numpy().view(np.uint64), er)


# ---- sharded IVF: every shard holds a slice of every posting list ----------------------------------------
def test_ivf_shard_plan_partitions_every_list():
    rng = np.random.default_rng(3)
    sizes = rng.integers(0, 40, 50)
    off = np.zeros(51, np.uint64); off[1:] = np.cumsum(sizes)
    for world in (1, 2, 3, 8):
        seen = []
        for r in range(world):
            lo, hi, local = sharding.ivf_shard_plan(off, r, world)
            assert (lo <= hi).all() and local[-1] == (hi - lo).sum()
            rows = sharding.ivf_shard_rows(lo, hi)
            assert len(rows) == local[-1]
            # local -> global is the inverse of the local order
            back = sharding.ivf_local_to_global(np.arange(len(rows), dtype=np.uint32), lo, local)
            assert np.array_equal(back.astype(np.int64), rows)
            seen.append(rows)
        allrows = np.sort(np.concatenate(seen))
        assert np.array_equal(allrows, np.arange(int(off[-1])))            # every row exactly once
    keep = sharding.ivf_local_to_global(np.array([0xFFFFFFFF], np.uint32), *sharding.ivf_shard_plan(off, 0, 2)[::2])
    assert keep[0] == np.uint64(0xFFFFFFFFFFFFFFFF)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3])
def test_sharded_ivf_search_equals_unsharded(world):
    """G stores on one device, each holding its slice of every list + the whole index's list sizes as budget:
    merged by (score, global row) the result IS the unsharded zvb200_ivf_search result (same lists visited)."""
    import zvec_b200 as z
    rng = np.random.default_rng(11)
    n, d, nlist, k, q = 30000, 32, 80, 10, 24
    centres = rng.standard_normal((nlist, d)).astype(np.float32)
    assign = np.sort(rng.integers(0, nlist, n))
    x = (centres[assign] + 0.3 * rng.standard_normal((n, d))).astype(np.float32)
    off = np.zeros(nlist + 1, np.uint64); off[1:] = np.cumsum(np.bincount(assign, minlength=nlist))
    qs = (centres[rng.integers(0, nlist, q)] + 0.3 * rng.standard_normal((q, d))).astype(np.float32)
    nprobe, max_scan = 8, 2500

    def build(rows, offsets):
        st = z.Store(z.DT_FP32, z.METRIC_L2SQ, d)
        st.append(np.ascontiguousarray(rows).view(np.uint8).reshape(len(rows), -1)); st.seal()
        st.ivf_attach(centres.view(np.uint8).reshape(nlist, -1), np.asarray(offsets, np.uint64))
        return st

    whole = build(x, off); cw = z.Context(whole)
    _, es, ei, ec = cw.ivf_search(qs, k, nprobe, max_scan)
    sizes = (off[1:] - off[:-1]).astype(np.uint64)
    all_s, all_r, keep = [], [], []
    for r in range(world):
        lo, hi, local = sharding.ivf_shard_plan(off, r, world)
        st = build(x[sharding.ivf_shard_rows(lo, hi)], local)
        st.ivf_set_budget_sizes(sizes)
        c = z.Context(st); keep.append((st, c))
        _, s, i, _ = c.ivf_search(qs, k, nprobe, max_scan)
        all_s.append(s); all_r.append(sharding.ivf_local_to_global(i, lo, local))
    ms, mr = sharding.merge_topk_global(np.stack(all_s), np.stack(all_r), k)
    for j in range(q):
        nv = int(ec[j])
        assert np.array_equal(ms[j][:nv].view(np.uint32), es[j][:nv].view(np.uint32))
        assert np.array_equal(mr[j][:nv], ei[j][:nv].astype(np.uint64))
        assert np.isnan(ms[j][nv:]).all()
    # the same exchange without leaving the device: per-shard zvb200_ivf_search_device, local -> global rows through
    # the per-list bases, one merge kernel by (score, global row)
    import torch
    dev = torch.device("cuda", 0)
    dq = torch.from_numpy(qs).to(dev)
    g_s = torch.empty((world, q, k), dtype=torch.float32, device=dev)
    g_i = torch.empty((world, q, k), dtype=torch.int32, device=dev)
    g_r = torch.empty((world, q, k), dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    for r, (st, c) in enumerate(keep):
        lo, hi, local = sharding.ivf_shard_plan(off, r, world)
        st.ivf_set_global_bases(lo.astype(np.uint64))
        c.ivf_search_device(dq.data_ptr(), q, k, nprobe, max_scan, g_s[r].data_ptr(), g_i[r].data_ptr())
        c.ivf_global_rows_device(g_i[r].data_ptr(), q * k, g_r[r].data_ptr())
        c.sync()
    f_s = torch.empty((q, k), dtype=torch.float32, device=dev)
    f_r = torch.empty((q, k), dtype=torch.int64, device=dev)
    c0 = keep[0][1]
    c0.merge_topk_rows_device(g_s.data_ptr(), g_r.data_ptr(), world, q, k, f_s.data_ptr(), f_r.data_ptr()); c0.sync()
    ds, dr = f_s.cpu().numpy(), f_r.cpu().numpy().view(np.uint64)
    assert np.array_equal(ds.view(np.uint32), ms.view(np.uint32)) and np.array_equal(dr, mr)
    for st, c in keep:
        c.close(); st.close()
    cw.close(); whole.close()
