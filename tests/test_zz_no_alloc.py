"""zvb200_search_device_noalloc (no counterpart in the reference): the
flat scan with every scratch buffer in the library's static device array.  Collected last on purpose.
Checked through the C ABI against the existing zvb200_search_device on the same device
buffers, bit for bit, on two calls in a row — the second call finds the first call's contents in the shared scratch
and must not see them."""
import numpy as np
import pytest

import support as S
import zvec_b200 as z

pytestmark = pytest.mark.gpu


def _device_search(ctx, fn, dq, q, k, filt=None):
    import torch
    s = torch.full((q, k), -7.0, dtype=torch.float32, device="cuda")
    i = torch.full((q, k), 12345, dtype=torch.int32, device="cuda")
    fn(dq.data_ptr(), q, k, s.data_ptr(), i.data_ptr(), filter_ptr=filt.data_ptr() if filt is not None else None)
    ctx.sync()
    return s.cpu().numpy(), i.cpu().numpy().view(np.uint32)


@pytest.mark.parametrize("kind", ["fp32_l2", "fp16_ip", "qi8_l2"])
def test_no_alloc_entry_point_equals_search_device_twice_in_a_row(kind):
    import torch
    from test_flat_gpu import make_stored
    n, d = 30011, 96
    metric, dt, sdim, stored, conv = make_stored(kind, n, d, seed=5)
    rng = np.random.default_rng(6)
    st = z.Store(dt, metric, sdim)
    st.append(stored); st.seal()
    st.set_deleted(17, True)
    ctx = z.Context(st)
    ctx.set_kernel(1)                       # the existing entry point on the same kernel: bit-for-bit comparable
    fb = np.zeros((n + 31) // 32, np.uint32); fb[3] = 0x00f0f0f0
    dfilt = torch.from_numpy(fb.view(np.int32)).cuda()
    # two calls in a row with different shapes: the second is smaller in q and k, so stale lists of the first call
    # sit right behind the region it uses
    for q, k, filt in ((8, 128, None), (3, 10, dfilt), (8, 100, None)):
        queries = np.ascontiguousarray(conv(rng.standard_normal((q, d)).astype(np.float32)))
        dq = torch.from_numpy(queries.view(np.uint8).reshape(q, -1)).cuda()
        es, ei = _device_search(ctx, ctx.search_device, dq, q, k, filt)
        for rep in range(2):
            gs, gi = _device_search(ctx, ctx.search_device_noalloc, dq, q, k, filt)
            assert np.array_equal(gi, ei), (kind, q, k, rep)
            assert np.array_equal(gs.view(np.uint32), es.view(np.uint32)), (kind, q, k, rep)
        assert ctx.last_kernel() == "scan_gemv"
        assert not (ei == 17).any() and (np.isfinite(es).all())
    ctx.close(); st.close()


def test_no_alloc_entry_point_rejects_sizes_beyond_its_static_scratch():
    import torch
    rng = np.random.default_rng(7)
    x = rng.standard_normal((2000, 32)).astype(np.float32)
    st = z.Store(z.DT_FP32, z.METRIC_L2SQ, 32)
    st.append(x); st.seal()
    ctx = z.Context(st)
    dq = torch.from_numpy(x[:16].copy()).cuda()
    s = torch.zeros((16, 256), dtype=torch.float32, device="cuda")
    i = torch.zeros((16, 256), dtype=torch.int32, device="cuda")
    L = z.lib()
    import ctypes as C
    args = lambda q, k: (ctx.h, C.c_void_p(dq.data_ptr()), q, k, C.c_float(3e38), None, C.c_void_p(s.data_ptr()),
                         C.c_void_p(i.data_ptr()))
    assert L.zvb200_search_device_noalloc(*args(9, 10)) == -31          # more than 8 queries
    assert b"noalloc" in L.zvb200_last_error()
    assert L.zvb200_search_device_noalloc(*args(2, 129)) == -31         # topk beyond 128
    assert L.zvb200_search_device_noalloc(*args(8, 128)) == 0
    ctx.sync()
    ctx.close(); st.close()
    # rows so wide that 8 prepared queries + the lists pass 4 MiB
    wide = z.Store(z.DT_FP16, z.METRIC_IP, 40000)
    wide.append(rng.standard_normal((64, 40000)).astype(np.float16)); wide.seal()
    c2 = z.Context(wide)
    dq2 = torch.zeros((8, 40000), dtype=torch.float16, device="cuda")
    r = L.zvb200_search_device_noalloc(c2.h, C.c_void_p(dq2.data_ptr()), 8, 128, C.c_float(3e38), None,
                                       C.c_void_p(s.data_ptr()), C.c_void_p(i.data_ptr()))
    assert r == -31 and b"4 MiB" in L.zvb200_last_error()
    c2.close(); wide.close()
