"""CPU tests: the oracle restatement is pinned against (1) the reference's own known-answer
vectors (tests/golden/kat.json, extracted by tests/golden/extract_kat.py) and (2) the
reference itself compiled from its sources (oracle/_ref), when that prebuilt library is there."""
import json
import os

import numpy as np
import pytest

import support as S

KAT = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "kat.json")))


def _encode(dtype, v):
    if dtype == "int8":
        return np.asarray(v, np.float64).astype(np.int8), S.DT_INT8
    if dtype == "fp32":
        return np.asarray(v, np.float32), S.DT_FP32
    return np.asarray(v, np.float32).astype(np.float16), S.DT_FP16


@pytest.mark.parametrize("i", range(len(KAT)))
def test_known_answer_vectors(i):
    c = KAT[i]
    a, dt = _encode(c["dtype"], c["a"])
    b, _ = _encode(c["dtype"], c["b"])
    metric = {"L2": S.M_L2, "L2SQ": S.M_L2SQ, "IP": S.M_IP, "IP_POS": S.M_IP}[c["metric"]]
    got = S.oracle().zo_distance(metric, dt, S.P(a), S.P(b), len(a))
    if c["metric"] == "IP_POS":
        got = -got
    exp = np.float32(c["expected"])
    if c["dtype"] == "int8" and c["metric"] != "L2":
        assert np.float32(got) == exp, c["source"]
    else:  # EXPECT_FLOAT_EQ = 4 ULP; fp16 operands are rounded by the test's own conversion
        tol = 4 * np.spacing(np.abs(exp)) if c["dtype"] != "fp16" else 2e-3 * max(abs(exp), 1.0)
        assert abs(np.float32(got) - exp) <= max(tol, 1e-6), (c["source"], got, exp)


def test_fp16_conversion_matches_numpy_rne():
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.standard_normal(20000).astype(np.float32) * s for s in (1e-8, 1e-4, 1, 100, 7e4)])
    x = np.concatenate([x, np.array([0.0, -0.0, 65504.0, 65520.0, 1e-7, 6.1e-5, 5.96e-8, 2.98e-8], np.float32)])
    out = np.zeros(x.size, np.uint16)
    S.oracle().zo_fp32_to_fp16(S.P(x), x.size, S.P(out))
    with np.errstate(over="ignore"):
        assert np.array_equal(out, x.astype(np.float16).view(np.uint16))
    back = np.zeros(x.size, np.float32)
    S.oracle().zo_fp16_to_fp32(S.P(out), x.size, S.P(back))
    assert np.array_equal(back, out.view(np.float16).astype(np.float32), equal_nan=True)


def test_heap_topk_semantics_small():
    # ties: the reference heap keeps earlier rows at the k-th score (strict '<' replace)
    rows = np.array([[3], [1], [1], [2], [1], [0]], np.int8)
    q = np.array([[0]], np.int8)
    _, sc, ix, cnt = S.oracle_flat_search(S.M_L2SQ, S.DT_INT8, rows, 1, q, 3)
    assert cnt[0] == 3 and list(sc[0]) == [0.0, 1.0, 1.0]
    assert ix[0][0] == 5 and set(ix[0][1:]) <= {1, 2, 4}
    # threshold gate and k > n
    _, sc, ix, cnt = S.oracle_flat_search(S.M_L2SQ, S.DT_INT8, rows, 1, q, 10, threshold=1.0)
    assert cnt[0] == 4 and np.isnan(sc[0][4:]).all() and (ix[0][4:] == 0xFFFFFFFF).all()


needs_ref = pytest.mark.skipif(S.ref() is None, reason="oracle/_ref not built (or CPU lacks the ISA)")


@needs_ref
@pytest.mark.parametrize("conv,metric,mid,dt", [
    ("", "SquaredEuclidean", S.M_L2SQ, S.DT_FP32),
    ("", "InnerProduct", S.M_IP, S.DT_FP32),
    ("HalfFloatConverter", "InnerProduct", S.M_IP, S.DT_FP16),
    ("HalfFloatConverter", "SquaredEuclidean", S.M_L2SQ, S.DT_FP16),
    ("CosineNormalizeConverter", "Cosine", S.M_COSINE, S.DT_FP32),
    ("Int8StreamingConverter", "SquaredEuclidean", S.M_QI8_L2SQ, S.DT_INT8),
    ("Int8StreamingConverter", "InnerProduct", S.M_QI8_IP, S.DT_INT8),
])
def test_restatement_matches_reference_search(conv, metric, mid, dt):
    rng = np.random.default_rng(42)
    n, d, q, k = 3000, 48, 6, 10
    x = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    qs = rng.uniform(-1, 1, (q, d)).astype(np.float32)
    ri = S.RefIndex(x, S.DT_FP32, d, metric, conv)
    assert ri.dtype == dt
    stored, tq = ri.stored(), ri.transform_queries(qs)
    rk, rs, rix, rc = ri.search(qs, k)
    ok, os_, oix, oc = S.oracle_flat_search(mid, dt, stored, ri.dim, tq, k)
    exact = dt == S.DT_INT8
    for i in range(q):
        alld = S.oracle_distances(mid, dt, stored, ri.dim, tq[i])
        # raw 1x1 distances of the reference metric vs the restatement
        rd = ri.distances(stored, tq[i])
        if exact:
            assert np.array_equal(rd.view(np.uint32), alld.view(np.uint32))
        else:
            assert np.allclose(rd, alld, rtol=2e-5, atol=1e-6)
        S.check_topk_parity(oix[i], os_[i], rix[i], rs[i], all_dist=alld, exact=exact,
                            what="%s/%s q%d" % (conv, metric, i))
    ri.close()


@needs_ref
def test_int8_quantizer_bytes_match_reference():
    rng = np.random.default_rng(7)
    x = rng.uniform(-1, 1, (2000, 96)).astype(np.float32)
    ri = S.RefIndex(x, S.DT_FP32, 96, "SquaredEuclidean", "Int8StreamingConverter")
    assert np.array_equal(ri.stored(), S.quantize_int8(x, True))
    assert np.array_equal(ri.transform_queries(x[:50]), S.quantize_int8(x[:50], True))
    ri.close()
    ri = S.RefIndex(x, S.DT_FP32, 96, "InnerProduct", "Int8StreamingConverter")
    assert np.array_equal(ri.stored(), S.quantize_int8(x, False))
    ri.close()


@needs_ref
def test_cosine_converter_layout_matches_reference():
    rng = np.random.default_rng(8)
    x = rng.standard_normal((500, 33)).astype(np.float32)
    ri = S.RefIndex(x, S.DT_FP32, 33, "Cosine", "CosineNormalizeConverter")
    st = ri.stored().view(np.float32).reshape(500, 34)
    mine = S.cosine_normalize_fp32(x)
    assert np.allclose(st, mine, rtol=3e-6, atol=1e-7)
    ri.close()
