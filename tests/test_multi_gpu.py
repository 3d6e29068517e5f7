"""The multi-GPU layer inside the C ABI (J1): zvb200_store_create_sharded (one process, several devices) and
zvb200_comm_* / zvb200_search_sharded_device (one process per GPU; peer mailboxes or one ncclAllGather).
Both must return exactly what ONE store holding all the rows returns — bit for bit, ties included, because the merge
orders by (score, global row).  Tests that need two devices skip on a one-GPU box (gpurun --gpus 2 runs them)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import support as S
import zvec_b200 as z
from test_flat_gpu import make_stored

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ndev():
    return z.lib().zvb200_device_count()


two_gpus = pytest.mark.skipif(_ndev() < 2, reason="needs two CUDA devices")


def _single(kind, stored, sdim, metric, dt, queries, k, **kw):
    st = z.Store(dt, metric, sdim)
    st.append(stored); st.seal()
    ctx = z.Context(st)
    out = ctx.search(queries, k, **kw)
    ctx.close(); st.close()
    return out


def _equal(a, b, what):
    assert np.array_equal(a[3], b[3]), what
    assert np.array_equal(a[2], b[2]), what                                   # row indices, ties included
    assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), what   # scores bit for bit (NaN padding too)
    assert np.array_equal(a[0], b[0]), what


def test_one_device_list_is_a_plain_store():
    metric, dt, sdim, stored, conv = make_stored("fp16_ip", 5000, 64, 1)
    qs = conv(np.random.default_rng(2).standard_normal((9, 64)).astype(np.float32))
    st = z.Store(dt, metric, sdim, devices=[0])
    st.append(stored); st.seal()
    ctx = z.Context(st)
    _equal(ctx.search(qs, 10), _single("fp16_ip", stored, sdim, metric, dt, qs, 10), "devices=[0]")
    ctx.close(); st.close()


@two_gpus
@pytest.mark.parametrize("kind", ["fp32_l2", "fp16_ip", "fp32_cos", "qi8_l2", "i8_ip"])
def test_sharded_store_equals_one_store(kind):
    n, d, k = 70001, 72, 10
    G = min(_ndev(), 4)
    metric, dt, sdim, stored, conv = make_stored(kind, n, d, 11)
    rng = np.random.default_rng(12)
    qs_all = np.ascontiguousarray(conv(rng.standard_normal((300, d)).astype(np.float32)))
    stored[40000] = stored[123]; stored[69999] = stored[123]          # exact duplicates on different shards: ties
    keys = (np.arange(n, dtype=np.uint64) * 3 + 1)
    keys[500] = np.uint64(0xFFFFFFFFFFFFFFFF)                          # a hole
    st = z.Store(dt, metric, sdim, devices=list(range(G)))
    with pytest.raises(z.ZvecB200Error):
        st.append(stored[:10])                                         # reserve() first: it fixes the shard size
    st.reserve(n)
    for lo, hi in ((0, 10), (10, 33000), (33000, n)):
        st.append(stored[lo:hi], keys[lo:hi])
    st.seal()
    one = z.Store(dt, metric, sdim)
    one.append(stored, keys); one.seal()
    for row in (7, 25000, 69000):
        st.set_deleted(row, True); one.set_deleted(row, True)
    ctx, c1 = z.Context(st), z.Context(one)
    fb = np.zeros((n + 31) // 32, np.uint32)
    for r in rng.integers(0, n, 5000):
        fb[r >> 5] |= np.uint32(1 << (r & 31))
    for q in (1, 4, 40, 300):
        _equal(ctx.search(qs_all[:q], k), c1.search(qs_all[:q], k), "%s %d queries over %d devices" % (kind, q, G))
    _equal(ctx.search(qs_all[:40], k, filter_bitmap=fb), c1.search(qs_all[:40], k, filter_bitmap=fb), "filter")
    th = float(np.median(c1.search(qs_all[:40], k)[1][:, k // 2]))
    _equal(ctx.search(qs_all[:40], k, threshold=th), c1.search(qs_all[:40], k, threshold=th), "threshold")
    _equal(ctx.search(qs_all[:40], 100), c1.search(qs_all[:40], 100), "k = 100")
    cand = rng.integers(0, n, (6, 700)).astype(np.uint32)
    _equal(ctx.search_rows(qs_all[:6], cand, k), c1.search_rows(qs_all[:6], cand, k), "search_rows")
    ctx.set_row_limit(50000); c1.set_row_limit(50000)
    _equal(ctx.search(qs_all[:8], k), c1.search(qs_all[:8], k), "row limit")
    ctx.set_row_limit(0); c1.set_row_limit(0)
    assert np.array_equal(st.get_vector(65000), one.get_vector(65000)) and st.count() == n
    newrow = stored[1]
    st.write(66000, newrow, 99); one.write(66000, newrow, 99)
    _equal(ctx.search(qs_all[:8], k), c1.search(qs_all[:8], k), "after an in-place write")
    for c in (ctx, c1):
        c.close()
    st.close(); one.close()


COMM_SCRIPT = r"""
import os, sys, time, json, numpy as np
sys.path.insert(0, %(root)r); sys.path.insert(0, %(tests)r)
import torch
import zvec_b200 as z
from zvec_b200 import sharding
from test_flat_gpu import make_stored
rank, world, tmp = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
dev = torch.device("cuda", rank %% max(1, torch.cuda.device_count()))
torch.cuda.set_device(dev)
n, d, k = 50021, 64, 10
metric, dt, sdim, stored, conv = make_stored("fp16_ip", n, d, 5)
stored[30000] = stored[77]                                         # a tie across ranks
qs = np.ascontiguousarray(conv(np.random.default_rng(6).standard_normal((130, d)).astype(np.float32)))
lo, hi = sharding.shard_bounds(n, world)[rank]
st = z.Store(dt, metric, sdim, device=dev.index)
st.append(stored[lo:hi]); st.seal()
ctx = z.Context(st)
idf = os.path.join(tmp, "uid.npy")
if rank == 0:
    uid = z.comm_unique_id(); np.save(idf + ".tmp.npy", uid); os.replace(idf + ".tmp.npy", idf)
else:
    while not os.path.exists(idf): time.sleep(0.05)
    uid = np.load(idf)
comm = z.Comm(ctx, rank, world, uid, sharding.shard_bases(n, world))
dq = torch.from_numpy(qs).to(dev)
res = {}
for q in (1, 130):
    fs = torch.empty((q, k), dtype=torch.float32, device=dev); fr = torch.empty((q, k), dtype=torch.int64, device=dev)
    for _ in range(3):                                             # repeated: the two mailbox slots alternate
        torch.cuda.synchronize()
        ctx.search_sharded_device(comm, dq.data_ptr(), q, k, fs.data_ptr(), fr.data_ptr())
        ctx.sync()
    comm.status()
    res[q] = (fs.cpu().numpy(), fr.cpu().numpy())
np.savez(os.path.join(tmp, "out%%d.npz" %% rank), s1=res[1][0], r1=res[1][1], s130=res[130][0], r130=res[130][1],
         transport=np.array([z.lib().zvb200_comm_transport(comm.h)]))
# everybody is done searching before anybody unmaps a mailbox
open(os.path.join(tmp, "done%%d" %% rank), "w").close()
while not all(os.path.exists(os.path.join(tmp, "done%%d" %% r)) for r in range(world)): time.sleep(0.05)
comm.close(); ctx.close(); st.close()
print("COMM-OK", rank)
"""


def _run_comm(world, tmp_path, env_extra):
    here = os.path.dirname(os.path.abspath(__file__))
    code = COMM_SCRIPT % {"root": ROOT, "tests": here}
    env = dict(os.environ, **env_extra)
    procs = [subprocess.Popen([sys.executable, "-c", code, str(r), str(world), str(tmp_path)], env=env,
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(world)]
    outs = [p.communicate(timeout=600) for p in procs]
    for p, (o, e) in zip(procs, outs):
        assert p.returncode == 0 and "COMM-OK" in o, o[-2000:] + e[-4000:]
    metric, dt, sdim, stored, conv = make_stored("fp16_ip", 50021, 64, 5)
    stored[30000] = stored[77]
    qs = np.ascontiguousarray(conv(np.random.default_rng(6).standard_normal((130, 64)).astype(np.float32)))
    exp = _single("fp16_ip", stored, sdim, metric, dt, qs, 10)
    transports = set()
    for r in range(world):
        d = np.load(os.path.join(str(tmp_path), "out%d.npz" % r))
        transports.add(int(d["transport"][0]))
        for q, ks, kr in ((1, "s1", "r1"), (130, "s130", "r130")):
            assert np.array_equal(d[ks].view(np.uint32), exp[1][:q].view(np.uint32)), "rank %d q=%d scores" % (r, q)
            assert np.array_equal(d[kr].astype(np.uint64) & 0xFFFFFFFF, exp[2][:q].astype(np.uint64)), "rank %d q=%d rows" % (r, q)
    return transports


@pytest.mark.parametrize("mailbox", ["1", "0"])
def test_communicator_of_one_rank(mailbox, tmp_path):
    """world = 1: the whole path (scan -> push / all-gather -> merge) on one GPU"""
    t = _run_comm(1, tmp_path, {"ZVB200_COMM_MAILBOX": mailbox})
    assert t == ({1} if mailbox == "1" else {0})


@two_gpus
@pytest.mark.parametrize("mailbox", ["1", "0"])
def test_communicator_of_two_ranks_equals_one_store(mailbox, tmp_path):
    t = _run_comm(2, tmp_path, {"ZVB200_COMM_MAILBOX": mailbox})
    assert len(t) == 1                                             # both ranks agreed on one transport
    if mailbox == "0":
        assert t == {0}
