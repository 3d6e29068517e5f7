"""GPU parity tests of SURVEY 8f N2: the query reformers on the device (IndexReformer::transform of
HalfFloatReformer / CosineReformer / IntegerStreamingReformer), the brute force over given rows
(FlatStreamer::search_bf_by_p_keys_impl) and the refiner flow (core_interface::Index::Search with a
refiner_param), against the oracle restatements and, where oracle/_ref is present, the reference's own
reformer objects."""
import ctypes as C

import numpy as np
import pytest

import support as S
import zvec_b200 as z

pytestmark = pytest.mark.gpu

FMAX = float(np.finfo(np.float32).max)


def _ctx(dt, metric, sdim, stored=None, keys=None):
    st = z.Store(dt, metric, sdim)
    if stored is not None:
        st.append(stored, keys)
    st.seal()
    return st, z.Context(st)


def _ulps(a, b):
    a = np.asarray(a, np.float32).view(np.int32).astype(np.int64)
    b = np.asarray(b, np.float32).view(np.int32).astype(np.int64)
    return np.abs(a - b)


@pytest.mark.parametrize("d", [1, 3, 33, 96, 128, 768, 1537])
@pytest.mark.parametrize("euclid", [True, False])
def test_reform_int8_record_is_bit_exact(d, euclid):
    rng = np.random.default_rng(d * 2 + euclid)
    x = (rng.standard_normal((37, d)) * rng.uniform(0.01, 30.0, (37, 1))).astype(np.float32)
    x[3] = 0.0                       # constant record: max - min clamps to epsilon
    x[4] = np.abs(x[4])              # one-sided range
    metric = S.M_QI8_L2SQ if euclid else S.M_QI8_IP
    st, ctx = _ctx(S.DT_INT8, metric, d + 16)
    got = ctx.reform_queries(x)
    exp = S.quantize_int8(x, euclid)
    assert got.shape == exp.shape
    bad = np.nonzero((got != exp).any(axis=1))[0]
    assert bad.size == 0, "records %s differ: %s vs %s" % (bad[:4], got[bad[0], -16:], exp[bad[0], -16:])
    assert ctx.last_kernel() == "reform"


@pytest.mark.parametrize("conv,metric_name,metric,dt,extra", [
    ("Int8StreamingConverter", "SquaredEuclidean", S.M_QI8_L2SQ, S.DT_INT8, 16),
    ("Int8StreamingConverter", "InnerProduct", S.M_QI8_IP, S.DT_INT8, 16),
    ("HalfFloatConverter", "InnerProduct", S.M_IP, S.DT_FP16, 0),
])
def test_reform_matches_the_reference_reformer_bit_for_bit(conv, metric_name, metric, dt, extra):
    if S.ref() is None:
        pytest.skip("oracle/_ref not available on this host")
    d = 96
    rng = np.random.default_rng(11)
    x = rng.standard_normal((64, d)).astype(np.float32)
    ri = S.RefIndex(x, S.DT_FP32, d, metric_name, conv)
    qs = (rng.standard_normal((50, d)) * 3).astype(np.float32)
    exp = ri.transform_queries(qs)
    st, ctx = _ctx(dt, metric, d + extra)
    got = ctx.reform_queries(qs)
    assert got.shape == exp.shape
    assert (got == exp).all()


def test_reform_fp16_rounds_like_the_reference():
    d = 64
    rng = np.random.default_rng(5)
    x = (rng.standard_normal((20, d)) * 10.0 ** rng.integers(-9, 6, (20, d))).astype(np.float32)
    x[0, :4] = [65504.0, 65520.0, 1e-8, -0.0]      # max half, rounds to inf, subnormal range, signed zero
    x[0, 4:8] = [np.inf, -np.inf, 2049.0, 2051.0]  # ties to even
    st, ctx = _ctx(S.DT_FP16, S.M_IP, d)
    got = ctx.reform_queries(x).view(np.uint16)
    O = S.oracle()
    exp = np.zeros((20, d), np.uint16)
    xc = np.ascontiguousarray(x)
    O.zo_fp32_to_fp16(C.c_void_p(xc.ctypes.data), xc.size, C.c_void_p(exp.ctypes.data))
    assert (got == exp).all()


@pytest.mark.parametrize("dt", [S.DT_FP32, S.DT_FP16])
@pytest.mark.parametrize("d", [5, 128, 1536])
def test_reform_cosine(dt, d):
    rng = np.random.default_rng(d)
    x = (rng.standard_normal((24, d)) * rng.uniform(0.1, 50, (24, 1))).astype(np.float32)
    x[2] = 0.0  # zero norm: left as it is (normalizer.h:45-51)
    exp = S.cosine_normalize_fp32(x)
    if dt == S.DT_FP32:
        st, ctx = _ctx(dt, S.M_COSINE, d + 1)
        got = ctx.reform_queries(x).view(np.float32)
        assert got.shape == (24, d + 1)
        assert _ulps(got[:, -1], exp[:, -1]).max() <= 1          # the norm
        assert _ulps(got[:, :-1], exp[:, :-1]).max() <= 2
        same = got[:, -1] == exp[:, -1]
        assert (got[same] == exp[same]).all()                    # same norm: same IEEE quotients
        assert (got[2] == 0).all()
    else:
        st, ctx = _ctx(dt, S.M_COSINE, d + 2)
        raw = ctx.reform_queries(x)
        assert raw.shape == (24, 2 * d + 4)
        norm = raw[:, 2 * d:].copy().view(np.float32)[:, 0]
        assert _ulps(norm, exp[:, -1]).max() <= 1
        h = raw[:, :2 * d].copy().view(np.float16).astype(np.float32)
        eh = exp[:, :-1].astype(np.float16).astype(np.float32)
        assert np.abs(h - eh).max() <= 2 ** -10 * np.abs(eh).max()
        same = norm == exp[:, -1]
        assert (h[same] == eh[same]).all()


def test_reform_rejects_what_the_reference_rejects():
    st, ctx = _ctx(S.DT_INT8, S.M_L2SQ, 32)   # plain int8: queries are int8 already
    with pytest.raises(z.ZvecB200Error):
        ctx.reform_queries(np.zeros((2, 32), np.float32))
    st, ctx = _ctx(S.DT_FP16, S.M_IP, 32)
    with pytest.raises(z.ZvecB200Error):
        ctx.reform_queries(np.zeros((2, 31), np.float32))   # dimension mismatch


KINDS = [("fp32_l2", S.M_L2SQ, S.DT_FP32, 0), ("fp16_ip", S.M_IP, S.DT_FP16, 0), ("fp32_cos", S.M_COSINE, S.DT_FP32, 1),
         ("fp16_cos", S.M_COSINE, S.DT_FP16, 2), ("qi8_l2", S.M_QI8_L2SQ, S.DT_INT8, 16), ("qi8_ip", S.M_QI8_IP, S.DT_INT8, 16)]


@pytest.mark.parametrize("kind,metric,dt,extra", KINDS)
@pytest.mark.parametrize("q", [3, 70])
def test_search_fp32_equals_reform_then_search(kind, metric, dt, extra, q):
    n, d, k = 6000, 80, 12
    rng = np.random.default_rng(17)
    x = rng.standard_normal((n, d)).astype(np.float32)
    st0, c0 = _ctx(dt, metric, d + extra)
    stored = c0.reform_queries(x)               # rows go through the same transform (converter == reformer here)
    st, ctx = _ctx(dt, metric, d + extra, stored)
    qs = rng.standard_normal((q, d)).astype(np.float32)
    k1, s1, i1, c1 = ctx.search_fp32(qs, k)
    k2, s2, i2, c2 = ctx.search(ctx.reform_queries(qs), k)
    assert (i1 == i2).all() and (s1.view(np.uint32) == s2.view(np.uint32)).all() and (c1 == c2).all()
    # and the scan itself agrees with the oracle on the reformed bytes
    qb = ctx.reform_queries(qs)
    _, es, ei, _ = S.oracle_flat_search(metric, dt, stored, d + extra, qb, k)
    for i in range(q):
        alld = S.oracle_distances(metric, dt, stored, d + extra, qb[i])
        S.check_topk_parity(i1[i], s1[i], ei[i], es[i], all_dist=alld, exact=dt == S.DT_INT8, what=kind)


@pytest.mark.parametrize("kind,metric,dt,extra", KINDS)
def test_search_rows_matches_oracle(kind, metric, dt, extra):
    n, d, k, q, m = 5000, 72, 10, 9, 64
    rng = np.random.default_rng(23)
    x = rng.standard_normal((n, d)).astype(np.float32)
    st0, c0 = _ctx(dt, metric, d + extra)
    stored = c0.reform_queries(x)
    keys = (np.arange(n, dtype=np.uint64) * 7 + 3)
    st, ctx = _ctx(dt, metric, d + extra, stored, keys)
    for r in (10, 11, 12):
        st.set_deleted(r, True)
    qb = c0.reform_queries(rng.standard_normal((q, d)).astype(np.float32))
    rows = rng.integers(0, n, (q, m)).astype(np.uint32)
    rows[:, 0] = 10                       # deleted
    rows[:, 1] = 0xFFFFFFFF               # a key the entity does not hold
    rows[:, 2] = n + 5                    # out of range
    rows[0, 3:] = 0xFFFFFFFF              # a query with no usable candidate but row 10 (deleted): empty result
    rows[1, 3:3 + 4] = rows[1, 7]         # fewer distinct candidates than k is fine too
    filt = np.zeros(n, bool); filt[rows[2, 5]] = True
    fb = np.zeros((n + 31) // 32, np.uint32)
    for r in np.nonzero(filt)[0]:
        fb[r >> 5] |= np.uint32(1 << (r & 31))
    gk, gs, gi, gc = ctx.search_rows(qb, rows, k, filter_bitmap=fb)
    assert ctx.last_kernel() == "scan_rows"
    for i in range(q):
        alld = S.oracle_distances(metric, dt, stored, d + extra, qb[i])
        cand = [int(r) for r in rows[i] if r < n and r not in (10, 11, 12) and not filt[r]]
        order = sorted(cand, key=lambda r: (alld[r], r))[:k]     # duplicates are scored twice, like the reference
        assert gc[i] == len(order), (i, gc[i], len(order))
        exp_s = np.array([alld[r] for r in order], np.float32)
        got_s = gs[i, :gc[i]]
        if dt == S.DT_INT8:
            assert (got_s.view(np.uint32) == exp_s.view(np.uint32)).all()
        else:
            assert np.allclose(got_s, exp_s, rtol=1e-4, atol=1e-6)
        assert sorted(gi[i, :gc[i]].tolist()) == sorted(order) or np.allclose(
            [alld[r] for r in gi[i, :gc[i]]], exp_s, rtol=1e-4, atol=1e-6)
        assert (gk[i, :gc[i]] == keys[gi[i, :gc[i]]]).all()
        assert (gi[i, gc[i]:] == 0xFFFFFFFF).all()
    assert gc[0] == 0


@pytest.mark.parametrize("coarse_metric,exact_metric", [(S.M_QI8_L2SQ, S.M_L2SQ), (S.M_QI8_IP, S.M_IP)])
@pytest.mark.parametrize("q,scale", [(4, 4.0), (70, 2.5), (5, 0.0)])
def test_refined_search_follows_the_reference_flow(coarse_metric, exact_metric, q, scale):
    """Index::Search with a refiner: coarse top floor(k*scale) on the int8 index, exact fp32 scores of those, best k."""
    n, d, k = 20000, 96, 10
    rng = np.random.default_rng(31)
    x = rng.standard_normal((n, d)).astype(np.float32)
    qs = rng.standard_normal((q, d)).astype(np.float32)
    euclid = coarse_metric == S.M_QI8_L2SQ
    stored_c = S.quantize_int8(x, euclid)
    keys = np.arange(n, dtype=np.uint64) + 1000
    stc, cc = _ctx(S.DT_INT8, coarse_metric, d + 16, stored_c, keys)
    ste, ce = _ctx(S.DT_FP32, exact_metric, d, x, keys)
    gk, gs, gi, gc = cc.search_refined(ce, qs, k, scale)
    m = int(np.floor(k * (scale if scale else 1.0)))
    qc = S.quantize_int8(qs, euclid)
    _, cs, ci, _ = S.oracle_flat_search(coarse_metric, S.DT_INT8, stored_c, d + 16, qc, m)
    xb = x.view(np.uint8).reshape(n, -1)
    for i in range(q):
        calld = S.oracle_distances(coarse_metric, S.DT_INT8, stored_c, d + 16, qc[i])
        alld = S.oracle_distances(exact_metric, S.DT_FP32, xb, d, qs[i].view(np.uint8))
        # candidates: the coarse top-m; rows tied with its last score are interchangeable
        kth = cs[i, m - 1]
        sure = set(int(r) for r in np.nonzero(calld < kth)[0])
        maybe = set(int(r) for r in np.nonzero(calld == kth)[0])
        got = [int(r) for r in gi[i, :gc[i]]]
        assert gc[i] == min(k, m)
        assert set(got) <= (sure | maybe)
        assert np.allclose(gs[i, :gc[i]], [alld[r] for r in got], rtol=1e-4, atol=1e-6)   # exact-store scores
        assert (np.diff(gs[i, :gc[i]]) >= 0).all()
        if len(maybe) == 1:   # no tie at the coarse boundary: the result is fully determined
            cand = sorted(sure | maybe, key=lambda r: (alld[r], r))[:k]
            S.check_topk_parity(gi[i], gs[i], np.array(cand + [0xFFFFFFFF] * (k - len(cand)), np.uint32),
                                np.array([alld[r] for r in cand] + [np.nan] * (k - len(cand)), np.float32),
                                all_dist=None, exact=False, what="refined")
        assert (gk[i, :gc[i]] == keys[gi[i, :gc[i]]]).all()


def test_refined_beats_coarse_recall():
    """what the refiner is for: int8 speed, fp32 ranking — recall against the exact top-k rises with scale_factor"""
    n, d, k, q = 30000, 64, 10, 40
    rng = np.random.default_rng(41)
    x = rng.standard_normal((n, d)).astype(np.float32)
    qs = rng.standard_normal((q, d)).astype(np.float32)
    stc, cc = _ctx(S.DT_INT8, S.M_QI8_L2SQ, d + 16, S.quantize_int8(x, True))
    ste, ce = _ctx(S.DT_FP32, S.M_L2SQ, d, x)
    _, _, exact_i, _ = ce.search(qs, k)
    _, _, coarse_i, _ = cc.search_fp32(qs, k)
    _, _, ref_i, _ = cc.search_refined(ce, qs, k, 4.0)

    def recall(a):
        return np.mean([len(set(a[i]) & set(exact_i[i])) / k for i in range(q)])
    assert recall(ref_i) >= recall(coarse_i)
    assert recall(ref_i) > 0.97
