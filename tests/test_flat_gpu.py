"""GPU parity tests: the CUDA scan (through the C ABI) against the oracle on the same seeded
inputs, and against the reference's own results where oracle/_ref is present."""
import numpy as np
import pytest

import support as S
import zvec_b200 as z

pytestmark = pytest.mark.gpu

ESZ = {S.DT_FP32: 4, S.DT_FP16: 2, S.DT_INT8: 1}


def make_stored(kind, n, d, seed, base=None):
    """-> (metric id, dtype, stored dim, stored rows bytes [n, elem], stored queries fn)"""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, d)).astype(np.float32) if base is None else base

    def conv(a):
        if kind in ("fp32_l2", "fp32_ip", "fp32_l2sqrt"):
            return a.astype(np.float32).view(np.uint8).reshape(len(a), -1)
        if kind in ("fp16_l2", "fp16_ip"):
            return a.astype(np.float16).view(np.uint8).reshape(len(a), -1)
        if kind == "fp32_cos":
            return S.cosine_normalize_fp32(a).view(np.uint8).reshape(len(a), -1)
        if kind == "fp16_cos":
            c = S.cosine_normalize_fp32(a)
            h = c[:, :-1].astype(np.float16).view(np.uint8).reshape(len(a), -1)
            return np.concatenate([h, c[:, -1:].copy().view(np.uint8).reshape(len(a), 4)], axis=1)
        if kind in ("i8_l2", "i8_ip"):
            return np.clip(np.round(a * 40), -127, 127).astype(np.int8).view(np.uint8).reshape(len(a), -1)
        if kind == "qi8_l2":
            return S.quantize_int8(a, True)
        if kind == "qi8_ip":
            return S.quantize_int8(a, False)
        if kind == "qi8_cos":
            c = S.cosine_normalize_fp32(a)
            qz = S.quantize_int8(c[:, :-1], False)
            return np.concatenate([qz, c[:, -1:].copy().view(np.uint8).reshape(len(a), 4)], axis=1)
        raise KeyError(kind)

    table = {
        "fp32_l2": (S.M_L2SQ, S.DT_FP32, d), "fp32_l2sqrt": (S.M_L2, S.DT_FP32, d), "fp32_ip": (S.M_IP, S.DT_FP32, d),
        "fp16_l2": (S.M_L2SQ, S.DT_FP16, d), "fp16_ip": (S.M_IP, S.DT_FP16, d),
        "fp32_cos": (S.M_COSINE, S.DT_FP32, d + 1), "fp16_cos": (S.M_COSINE, S.DT_FP16, d + 2),
        "i8_l2": (S.M_L2SQ, S.DT_INT8, d), "i8_ip": (S.M_IP, S.DT_INT8, d),
        "qi8_l2": (S.M_QI8_L2SQ, S.DT_INT8, d + 16), "qi8_ip": (S.M_QI8_IP, S.DT_INT8, d + 16),
        "qi8_cos": (S.M_QI8_COSINE, S.DT_INT8, d + 20),
    }
    metric, dt, sdim = table[kind]
    return metric, dt, sdim, np.ascontiguousarray(conv(x)), conv


KINDS = ["fp32_l2", "fp32_l2sqrt", "fp32_ip", "fp16_l2", "fp16_ip", "fp32_cos", "fp16_cos", "i8_l2", "i8_ip",
         "qi8_l2", "qi8_ip", "qi8_cos"]


def run_case(kind, n, d, q, k, seed, kernel=0, drop=None, keys=None, threshold=None, filt=None):
    metric, dt, sdim, stored, conv = make_stored(kind, n, d, seed)
    rng = np.random.default_rng(seed + 1000)
    queries = np.ascontiguousarray(conv(rng.standard_normal((q, d)).astype(np.float32)))
    st = z.Store(dt, metric, sdim)
    half = n // 2
    st.append(stored[:half], None if keys is None else keys[:half])   # two appends: growth path
    st.append(stored[half:], None if keys is None else keys[half:])
    st.seal()
    if drop is not None:
        for r in np.nonzero(drop)[0]:
            st.set_deleted(int(r), True)
    ctx = z.Context(st)
    ctx.set_kernel(kernel)
    fb = None
    odrop = None if drop is None else drop.astype(np.uint8).copy()
    if filt is not None:
        fb = np.zeros((n + 31) // 32, np.uint32)
        for r in np.nonzero(filt)[0]:
            fb[r >> 5] |= np.uint32(1 << (r & 31))
        odrop = filt.astype(np.uint8) if odrop is None else (odrop | filt.astype(np.uint8))
    th = float(np.finfo(np.float32).max) if threshold is None else threshold
    gk, gs, gi, gc = ctx.search(queries, k, threshold=th, filter_bitmap=fb)
    ek, es, ei, ec = S.oracle_flat_search(metric, dt, stored, sdim, queries, k, threshold=threshold, keys=keys,
                                          drop=odrop)
    exact = dt == S.DT_INT8
    for i in range(q):
        alld = S.oracle_distances(metric, dt, stored, sdim, queries[i])
        S.check_topk_parity(gi[i], gs[i], ei[i], es[i], all_dist=alld, exact=exact,
                            what="%s n=%d d=%d q%d k=%d [%s]" % (kind, n, d, i, k, ctx.last_kernel()))
        assert gc[i] == ec[i]
        nv = int(gc[i])
        exp_keys = gi[i][:nv].astype(np.uint64) if keys is None else np.asarray(keys, np.uint64)[gi[i][:nv]]
        assert np.array_equal(gk[i][:nv], exp_keys)
        assert (gk[i][nv:] == np.uint64(0xFFFFFFFFFFFFFFFF)).all() and np.isnan(gs[i][nv:]).all()
    name = ctx.last_kernel()
    ctx.close(); st.close()
    return name


STREAMING = {0: ("scan_gemv", "scan_mma"), 1: ("scan_gemv",), 4: ("scan_stream",)}


@pytest.mark.parametrize("kernel", [0, 1, 4])      # auto, LDG scan, bulk-async scan
@pytest.mark.parametrize("kind", KINDS)
def test_batch1_parity_all_metrics(kind, kernel):
    assert run_case(kind, 5000, 96, 1, 10, seed=11, kernel=kernel) in STREAMING[kernel]


@pytest.mark.parametrize("kernel", [0, 1, 4])
@pytest.mark.parametrize("kind", KINDS)
def test_small_batch_parity(kind, kernel):
    run_case(kind, 3000, 64, 7, 10, seed=12, kernel=kernel)


@pytest.mark.parametrize("kernel", [1, 4])
@pytest.mark.parametrize("d", [1, 3, 4, 7, 17, 33, 100, 128, 257, 768, 1536, 4100])
@pytest.mark.parametrize("kind", ["fp32_l2", "fp16_ip", "qi8_l2"])
def test_ragged_dimensions(kind, d, kernel):
    run_case(kind, 1500, d, 2, 5, seed=13 + d, kernel=kernel)


@pytest.mark.parametrize("kernel", [1, 4])
@pytest.mark.parametrize("k", [1, 2, 31, 32, 33, 100, 257, 1024])
def test_topk_sizes(k, kernel):
    run_case("fp32_l2", 4000, 32, 2, k, seed=14, kernel=kernel)
    run_case("i8_l2", 4000, 32, 1, k, seed=15, kernel=kernel)   # heavy ties


@pytest.mark.parametrize("n", [1, 7, 8, 9, 255, 257, 70001])
def test_stream_kernel_ragged_row_counts(n):
    run_case("fp32_l2", n, 40, 3, 10, seed=16 + n, kernel=4)
    run_case("qi8_ip", n, 40, 1, 10, seed=17 + n, kernel=4)


def test_k_larger_than_n_and_tiny_stores():
    run_case("fp32_l2", 2, 8, 1, 10, seed=16)
    run_case("fp32_ip", 37, 8, 3, 100, seed=17)
    run_case("qi8_l2", 33, 8, 1, 64, seed=18)


def test_deleted_rows_keys_and_holes():
    n = 3000
    rng = np.random.default_rng(3)
    drop = rng.random(n) < 0.3
    keys = (np.arange(n, dtype=np.uint64) * 7 + 5)
    holes = rng.random(n) < 0.1
    keys[holes] = np.uint64(0xFFFFFFFFFFFFFFFF)      # kInvalidKey holes are skipped by the scan
    run_case("fp32_l2", n, 24, 3, 10, seed=19, drop=drop, keys=keys)


def test_filter_bitmap_and_threshold():
    n = 4000
    rng = np.random.default_rng(4)
    filt = rng.random(n) < 0.5
    run_case("fp16_ip", n, 40, 2, 20, seed=20, filt=filt)
    run_case("fp32_l2", n, 16, 2, 50, seed=21, threshold=12.0)     # RNN radius: fewer than k results
    run_case("fp32_l2", n, 16, 1, 10, seed=22, filt=np.ones(n, bool))  # everything filtered


def test_exact_duplicates_tie_break_by_row():
    # 64 identical rows: every score ties; we must return rows 0..k-1 in order
    st = z.Store(S.DT_FP32, S.M_L2SQ, 8)
    st.append(np.ones((64, 8), np.float32)); st.seal()
    ctx = z.Context(st)
    _, sc, ix, cnt = ctx.search(np.zeros((1, 8), np.float32), 10)
    assert list(ix[0]) == list(range(10)) and (sc[0] == 8.0).all() and cnt[0] == 10


@pytest.mark.skipif(S.ref() is None, reason="oracle/_ref not available on this host")
@pytest.mark.parametrize("conv,metric,origin", [
    ("", "SquaredEuclidean", None), ("HalfFloatConverter", "InnerProduct", None),
    ("CosineNormalizeConverter", "Cosine", None), ("Int8StreamingConverter", "SquaredEuclidean", "SquaredEuclidean"),
    ("Int8StreamingConverter", "InnerProduct", "InnerProduct"),
])
def test_against_reference_itself(conv, metric, origin):
    """Same inputs through the reference's FlatSearcher (oracle/_ref) and through the GPU path,
    fed with the reference's own stored bytes and reformer-transformed queries."""
    rng = np.random.default_rng(77)
    n, d, q, k = 20000, 96, 8, 10
    x = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    qs = rng.uniform(-1, 1, (q, d)).astype(np.float32)
    ri = S.RefIndex(x, S.DT_FP32, d, metric, conv)
    stored, tq = ri.stored(), ri.transform_queries(qs)
    mid = z.metric_from_names(ri.metric_name, origin)
    st = z.Store(ri.dtype, mid, ri.dim)
    st.append(stored); st.seal()
    ctx = z.Context(st)
    _, gs, gi, _ = ctx.search(tq, k)
    _, rs, rix, _ = ri.search(qs, k)
    for i in range(q):
        alld = S.oracle_distances(mid, ri.dtype, stored, ri.dim, tq[i])
        S.check_topk_parity(gi[i], gs[i], rix[i], rs[i], all_dist=alld, exact=ri.dtype == S.DT_INT8,
                            what="vs reference %s/%s q%d" % (conv, metric, i))
    ri.close()


def test_distance_matrix_contract():
    for kind in ("fp32_l2", "fp16_ip", "qi8_l2", "i8_l2", "fp32_cos"):
        metric, dt, sdim, stored, conv = make_stored(kind, 257, 50, 5)
        queries = np.ascontiguousarray(conv(np.random.default_rng(6).standard_normal((3, 50)).astype(np.float32)))
        got = z.distance_matrix(metric, dt, sdim, stored, queries)
        for j in range(3):
            exp = S.oracle_distances(metric, dt, stored, sdim, queries[j])
            if dt == S.DT_INT8:
                assert np.array_equal(got[j].view(np.uint32), exp.view(np.uint32)), kind
            else:
                assert np.allclose(got[j], exp, rtol=1e-5, atol=1e-6), kind


def test_get_vector_round_trip():
    metric, dt, sdim, stored, _ = make_stored("qi8_l2", 100, 24, 9)
    st = z.Store(dt, metric, sdim); st.append(stored); st.seal()
    for r in (0, 57, 99):
        assert np.array_equal(st.get_vector(r), stored[r])


# ---- K2c tiled kernel (batched) ---------------------------------------------------------
@pytest.mark.parametrize("kind", KINDS)
def test_tile_kernel_parity_all_metrics(kind):
    assert run_case(kind, 4000, 96, 70, 10, seed=31, kernel=2) == "scan_tile"


@pytest.mark.parametrize("d", [1, 5, 16, 17, 63, 64, 65, 200, 768])
@pytest.mark.parametrize("kind", ["fp32_l2", "fp16_ip", "qi8_l2", "i8_l2"])
def test_tile_kernel_ragged_dimensions(kind, d):
    run_case(kind, 1000, d, 9, 7, seed=32 + d, kernel=2)


@pytest.mark.parametrize("k", [1, 33, 100, 128])
def test_tile_kernel_topk_sizes(k):
    run_case("fp32_l2", 3000, 24, 65, k, seed=33, kernel=2)
    run_case("i8_ip", 3000, 24, 5, k, seed=34, kernel=2)     # heavy ties


def test_tile_kernel_auto_selected_for_batches_and_filters():
    n = 3000
    rng = np.random.default_rng(35)
    assert run_case("i8_l2", 600, 1100, 40, 10, seed=35) == "scan_tile"     # int8 beyond d = 1040: exact CUDA-core tile
    drop = rng.random(n) < 0.3
    filt = rng.random(n) < 0.3
    run_case("fp16_l2", n, 32, 40, 10, seed=36, kernel=2, drop=drop, filt=filt)
    run_case("fp32_l2", 50, 8, 33, 100, seed=37, kernel=2)            # k > n
    run_case("fp32_l2", n, 16, 33, 20, seed=38, kernel=2, threshold=10.0)
    assert run_case("fp32_l2", n, 16, 33, 300, seed=39) == "scan_gemv"  # k beyond the tile kernel's lists


def test_tile_and_gemv_agree_bitwise_on_int8():
    metric, dt, sdim, stored, conv = make_stored("qi8_l2", 6000, 96, 40)
    queries = np.ascontiguousarray(conv(np.random.default_rng(41).standard_normal((48, 96)).astype(np.float32)))
    st = z.Store(dt, metric, sdim); st.append(stored); st.seal()
    ctx = z.Context(st)
    ctx.set_kernel(1); _, s1, i1, _ = ctx.search(queries, 100)
    ctx.set_kernel(2); _, s2, i2, _ = ctx.search(queries, 100)
    assert np.array_equal(s1.view(np.uint32), s2.view(np.uint32)) and np.array_equal(i1, i2)


# ---- K2 tcgen05 / TMEM kernel: fp16 InnerProduct / Cosine straight from the accumulator; fp16 L2 and
# every fp32 metric (kind::tf32) with the accumulator as filter + exact re-score; int8 and QuantizedInteger
# (kind::i8, exact int32 accumulators, the reference's float epilogue per column: bit-exact) ------------
MMA_KINDS = ["fp16_ip", "fp16_cos", "fp16_l2", "fp32_l2", "fp32_l2sqrt", "fp32_ip", "fp32_cos",
             "i8_l2", "i8_ip", "qi8_l2", "qi8_ip", "qi8_cos"]


@pytest.mark.parametrize("kind", MMA_KINDS)
@pytest.mark.parametrize("q", [32, 128, 130, 300])
def test_mma_kernel_parity(kind, q):
    assert run_case(kind, 5000, 96, q, 10, seed=51 + q, kernel=3) == "scan_mma"


@pytest.mark.parametrize("kind", ["fp16_ip", "fp16_l2", "fp32_l2", "fp32_cos", "qi8_l2", "i8_l2"])
@pytest.mark.parametrize("d", [1, 8, 33, 64, 65, 100, 200, 768])
def test_mma_kernel_ragged_dimensions(kind, d):
    run_case(kind, 3000, d, 40, 7, seed=52 + d, kernel=3)


@pytest.mark.parametrize("kind", ["fp16_ip", "fp16_l2", "fp32_l2", "qi8_l2"])
@pytest.mark.parametrize("n", [1, 37, 255, 256, 257, 40000])
def test_mma_kernel_ragged_row_counts(kind, n):
    run_case(kind, n, 48, 33, 10, seed=53 + n, kernel=3)


@pytest.mark.parametrize("kind", ["fp16_ip", "fp32_l2", "qi8_l2"])
@pytest.mark.parametrize("k", [1, 10, 32, 33, 80, 81, 100, 128])
def test_mma_kernel_topk_sizes(kind, k):
    if kind != "fp16_ip" and k == 128:
        k = 112     # these modes keep 2-8 KB of shared memory for column constants
    run_case(kind, 6000, 32, 70, k, seed=54 + k, kernel=3)


def test_mma_kernel_filters_threshold_and_auto_selection():
    n = 9000
    rng = np.random.default_rng(55)
    drop = rng.random(n) < 0.3
    filt = rng.random(n) < 0.3
    run_case("fp16_ip", n, 64, 64, 10, seed=55, kernel=3, drop=drop, filt=filt)
    run_case("fp16_cos", n, 64, 64, 20, seed=56, kernel=3, threshold=0.8)
    run_case("fp16_ip", n, 64, 40, 10, seed=57, kernel=3, filt=np.ones(n, bool))
    assert run_case("fp16_ip", n, 64, 256, 10, seed=58) == "scan_mma"      # auto: batched fp16 IP -> tensor cores
    assert run_case("fp16_l2", n, 64, 256, 10, seed=59) == "scan_mma"      # L2: tensor-core filter + exact re-score
    assert run_case("fp32_l2", n, 64, 256, 10, seed=62) == "scan_mma"      # fp32: kind::tf32 filter + exact re-score
    assert run_case("fp32_l2", n, 64, 3, 10, seed=63) == "scan_gemv"       # tiny batches stay on the streaming scan
    assert run_case("qi8_l2", n, 64, 256, 10, seed=64) == "scan_mma"       # int8 batched: kind::i8, exact
    assert run_case("i8_ip", n, 64, 256, 10, seed=68) == "scan_mma"
    run_case("qi8_l2", n, 64, 64, 10, seed=69, kernel=3, drop=drop, filt=filt)
    run_case("qi8_ip", n, 64, 64, 20, seed=70, kernel=3, threshold=-0.5)
    run_case("fp32_l2", n, 64, 64, 10, seed=65, kernel=3, drop=drop, filt=filt)
    run_case("fp32_cos", n, 64, 64, 20, seed=66, kernel=3, threshold=0.8)
    run_case("fp16_l2", n, 64, 64, 20, seed=67, kernel=3, threshold=100.0, drop=drop)


@pytest.mark.parametrize("kind", ["fp16_l2", "fp32_l2", "fp32_ip", "qi8_l2"])
@pytest.mark.parametrize("shape", ["offset", "norm_spread", "near_duplicates"])
def test_mma_rescore_filter_is_sound_on_hostile_data(kind, shape):
    """The tensor-core accumulator only nominates rows; these inputs maximise its error against the exact
    score (cancellation in |x|^2 - 2xq + |q|^2, tf32 truncation) — the results must still equal the oracle's."""
    n, d, q, k = 20000, 64, 48, 10
    rng = np.random.default_rng(70)
    base = rng.standard_normal((n, d)).astype(np.float32)
    qs = rng.standard_normal((q, d)).astype(np.float32)
    if shape == "offset":                       # large common component: distances << norms
        base = base * 0.05 + 30.0
        qs = qs * 0.05 + 30.0
    elif shape == "norm_spread":                # norms over 4 decades
        base *= (10.0 ** rng.uniform(-2, 2, (n, 1))).astype(np.float32)
    else:                                       # the queries are (almost) stored rows
        qs = base[rng.integers(0, n, q)] + 1e-3 * rng.standard_normal((q, d)).astype(np.float32)
    metric, dt, sdim, stored, conv = make_stored(kind, n, d, 71, base=base)
    queries = np.ascontiguousarray(conv(qs))
    st = z.Store(dt, metric, sdim); st.append(stored); st.seal()
    ctx = z.Context(st); ctx.set_kernel(3)
    gk, gs, gi, gc = ctx.search(queries, k)
    assert ctx.last_kernel() == "scan_mma"
    ctx.set_kernel(1)
    ek, es, ei, ec = ctx.search(queries, k)      # the streaming scan: exact arithmetic, oracle-checked elsewhere
    for i in range(q):
        alld = S.oracle_distances(metric, dt, stored, sdim, queries[i])
        S.check_topk_parity(gi[i], gs[i], ei[i], es[i], all_dist=alld, exact=kind == "qi8_l2",
                            what="%s %s q%d" % (kind, shape, i))
    ctx.close(); st.close()


# ---- QuantizedInteger L2: two tensor-core forms.  Kernel 3 contracts the store's bf16 shadow operands
# (kind::f16, no per-column constants), kernel 5 the int8 codes themselves (kind::i8); both only nominate
# with a sound margin and re-score from the int8 codes, so both must be bit-exact -------------------------
@pytest.mark.parametrize("kernel,mma_kind", [(3, 3), (5, 2)])
@pytest.mark.parametrize("d,q,k", [(96, 300, 100), (96, 64, 10), (7, 33, 10), (200, 130, 32), (1000, 40, 16)])
def test_qi8_l2_tensor_core_forms_are_bit_exact(kernel, mma_kind, d, q, k):
    metric, dt, sdim, stored, conv = make_stored("qi8_l2", 30000, d, 90 + d)
    queries = np.ascontiguousarray(conv(np.random.default_rng(91 + d).standard_normal((q, d)).astype(np.float32)))
    st = z.Store(dt, metric, sdim); st.append(stored); st.seal()
    ctx = z.Context(st); ctx.set_kernel(kernel)
    gk, gs, gi, gc = ctx.search(queries, k)
    # the shadow operands are kept only up to 320 dimensions (beyond, bf16's half rate loses to kind::i8)
    assert ctx.last_kernel() == "scan_mma" and ctx.last_mma_kind() == (mma_kind if d <= 320 else 2)
    ek, es, ei, ec = S.oracle_flat_search(metric, dt, stored, sdim, queries, k)
    for i in range(q):
        alld = S.oracle_distances(metric, dt, stored, sdim, queries[i])
        S.check_topk_parity(gi[i], gs[i], ei[i], es[i], all_dist=alld, exact=True, what="qi8_l2 kernel %d q%d" % (kernel, i))
    ctx.close(); st.close()


@pytest.mark.parametrize("kernel", [3, 5])
@pytest.mark.parametrize("shape", ["offset", "norm_spread", "near_duplicates", "constant_rows", "huge_values"])
def test_qi8_l2_tensor_core_filter_is_sound_on_hostile_data(kernel, shape):
    """Per-record scales over many decades, large common offsets (mb >> ma * codes), constant records (the
    quantiser's epsilon scale), values far beyond fp16's range: the bf16 / int8 filters must still nominate
    every true neighbour, and the results must equal the streaming scan's bit for bit."""
    n, d, q, k = 20000, 96, 48, 10
    rng = np.random.default_rng(92)
    base = rng.standard_normal((n, d)).astype(np.float32)
    qs = rng.standard_normal((q, d)).astype(np.float32)
    if shape == "offset":
        base = base * 0.05 + 30.0; qs = qs * 0.05 + 30.0
    elif shape == "norm_spread":
        base *= (10.0 ** rng.uniform(-3, 3, (n, 1))).astype(np.float32)
        qs *= (10.0 ** rng.uniform(-3, 3, (q, 1))).astype(np.float32)
    elif shape == "near_duplicates":
        qs = base[rng.integers(0, n, q)] + 1e-3 * rng.standard_normal((q, d)).astype(np.float32)
    elif shape == "constant_rows":
        base[::7] = rng.standard_normal((len(base[::7]), 1)).astype(np.float32)      # max == min inside a record
        qs[::5] = 0.25
    else:
        base *= 1.0e4; qs *= 1.0e4
    metric, dt, sdim, stored, conv = make_stored("qi8_l2", n, d, 93, base=base)
    queries = np.ascontiguousarray(conv(qs))
    st = z.Store(dt, metric, sdim); st.append(stored); st.seal()
    ctx = z.Context(st); ctx.set_kernel(kernel)
    gk, gs, gi, gc = ctx.search(queries, k)
    assert ctx.last_kernel() == "scan_mma"
    ctx.set_kernel(1)
    ek, es, ei, ec = ctx.search(queries, k)
    for i in range(q):
        alld = S.oracle_distances(metric, dt, stored, sdim, queries[i])
        S.check_topk_parity(gi[i], gs[i], ei[i], es[i], all_dist=alld, exact=True, what="qi8_l2 %s q%d" % (shape, i))
    ctx.close(); st.close()


def test_qi8_l2_shadow_follows_appends_and_overwrites():
    """rows added after a seal and rows overwritten in place (FlatStreamer add / add_with_id) must reach the
    shadow operands too"""
    d, k = 64, 10
    metric, dt, sdim, stored, conv = make_stored("qi8_l2", 6000, d, 94)
    queries = np.ascontiguousarray(conv(np.random.default_rng(95).standard_normal((40, d)).astype(np.float32)))
    st = z.Store(dt, metric, sdim); st.append(stored[:2000]); st.seal()
    ctx = z.Context(st); ctx.set_kernel(3)
    ctx.search(queries, k)
    st.append(stored[2000:]); st.seal()
    cur = stored.copy()
    for r in (5, 2500, 5999):
        cur[r] = queries[r % 40]                      # a stored row equal to a query: distance 0 must win
        st.write(r, cur[r], r)
    gk, gs, gi, gc = ctx.search(queries, k)
    assert ctx.last_mma_kind() == 3
    ek, es, ei, ec = S.oracle_flat_search(metric, dt, cur, sdim, queries, k)
    for i in range(40):
        alld = S.oracle_distances(metric, dt, cur, sdim, queries[i])
        S.check_topk_parity(gi[i], gs[i], ei[i], es[i], all_dist=alld, exact=True, what="after appends q%d" % i)
    ctx.close(); st.close()


def test_mma_and_tile_agree_on_ids():
    metric, dt, sdim, stored, conv = make_stored("fp16_ip", 50000, 768, 60)
    queries = np.ascontiguousarray(conv(np.random.default_rng(61).standard_normal((256, 768)).astype(np.float32)))
    st = z.Store(dt, metric, sdim); st.append(stored); st.seal()
    ctx = z.Context(st)
    ctx.set_kernel(2); _, s2, i2, _ = ctx.search(queries, 10)
    ctx.set_kernel(3); _, s3, i3, _ = ctx.search(queries, 10)
    assert np.allclose(s2, s3, rtol=1e-4, atol=1e-6)
    assert (i2 == i3).mean() > 0.999        # order may differ only inside near-ties


@pytest.mark.parametrize("kind,k", [("fp16_ip", 100), ("fp16_l2", 10), ("fp32_l2", 100), ("qi8_l2", 100), ("i8_l2", 33)])
def test_mma_results_are_deterministic_across_runs(kind, k):
    """races would show as run-to-run differences: the cooperative pool (k > 16), the constants ring fed by warp 3
    and the union bound across CTAs only ever remove work — the lists must be bit-identical every time, and equal
    to the streaming scan's"""
    metric, dt, sdim, stored, conv = make_stored(kind, 300000, 64, 97)
    queries = np.ascontiguousarray(conv(np.random.default_rng(98).standard_normal((257, 64)).astype(np.float32)))
    st = z.Store(dt, metric, sdim); st.append(stored); st.seal()
    ctx = z.Context(st); ctx.set_kernel(3)
    _, s0, i0, c0 = ctx.search(queries, k)
    assert ctx.last_kernel() == "scan_mma"
    for _ in range(12):
        _, s1, i1, c1 = ctx.search(queries, k)
        assert np.array_equal(i0, i1) and np.array_equal(s0.view(np.uint32), s1.view(np.uint32)) and np.array_equal(c0, c1)
    ctx.set_kernel(1)
    _, s2, i2, _ = ctx.search(queries[:16], k)
    for q in range(16):
        alld = S.oracle_distances(metric, dt, stored, sdim, queries[q])
        S.check_topk_parity(i0[q], s0[q], i2[q], s2[q], all_dist=alld, exact=dt == S.DT_INT8, what="%s q%d" % (kind, q))
    ctx.close(); st.close()


def test_concurrent_contexts_on_one_store():
    """IndexSearcher::search_impl is const and called from many threads, one context each
    (src/core/interface/index.cc:24-45; flat_streamer_test.cc:509-730): eight threads, own context and stream each,
    different batch sizes / kernels at the same time — every result equals the serial one"""
    import threading
    metric, dt, sdim, stored, conv = make_stored("fp16_ip", 200000, 96, 99)
    queries = np.ascontiguousarray(conv(np.random.default_rng(100).standard_normal((192, 96)).astype(np.float32)))
    st = z.Store(dt, metric, sdim); st.append(stored); st.seal()
    plans = [(1, 10), (3, 10), (64, 10), (192, 100), (1, 100), (130, 16), (40, 33), (7, 10)]
    serial = []
    for q, k in plans:
        c = z.Context(st); serial.append(c.search(queries[:q], k)); c.close()
    errors = []

    def work(i):
        try:
            q, k = plans[i]
            c = z.Context(st)
            for _ in range(20):
                got = c.search(queries[:q], k)
                for a, b in zip(got, serial[i]):
                    assert np.array_equal(np.asarray(a).view(np.uint8), np.asarray(b).view(np.uint8))
            c.close()
        except BaseException as e:  # noqa: BLE001
            errors.append((i, repr(e)))

    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(plans))]
    [t.start() for t in ts]; [t.join() for t in ts]
    assert not errors, errors
    st.close()


@pytest.mark.parametrize("kernel", [1, 2, 3, 4])
@pytest.mark.parametrize("kind", ["qi8_ip", "qi8_cos", "qi8_l2"])
def test_operand_order_of_the_streamer_is_bit_exact(kind, kernel):
    """FlatStreamer / the IVF list scan hand the metric (query, feature), FlatSearcher (feature, query)
    (flat_streamer_entity.h:155-162 vs flat_distance_matrix.h:53-55); the QuantizedInteger -IP epilogue sums its middle
    terms in operand order.  With zvb200_store_set_operand_order(1) every score equals the restatement called with the
    operands swapped, bit for bit (L2 is symmetric, it must not move)."""
    n, d, q, k = 1200, 32, 6, 100
    metric, dt, sdim, stored, conv = make_stored(kind, n, d, seed=91)
    rng = np.random.default_rng(92)
    queries = np.ascontiguousarray(conv(rng.standard_normal((q, d)).astype(np.float32)))
    st = z.Store(dt, metric, sdim)
    st.append(stored); st.seal()
    st.set_operand_order(True)
    ctx = z.Context(st)
    ctx.set_kernel(kernel)
    _, gs, gi, _ = ctx.search(queries, k)
    moved = 0
    for i in range(q):
        # the restatement with the roles swapped: the query is the "stored row", each row the "query"
        swapped = np.array([S.oracle_distances(metric, dt, queries[i:i + 1], sdim, stored[r])[0] for r in range(n)],
                           np.float32)
        plain = S.oracle_distances(metric, dt, stored, sdim, queries[i])
        moved += int((swapped.view(np.uint32) != plain.view(np.uint32)).sum())
        order = np.lexsort((np.arange(n), swapped))[:k]
        assert np.array_equal(gs[i].view(np.uint32), swapped[order].view(np.uint32)), (kind, kernel, i)
        assert np.array_equal(gi[i], order.astype(np.uint32))
    assert (moved == 0) == (kind == "qi8_l2")     # the case tests something: -IP scores do move
    ctx.close(); st.close()


def test_replayed_searches_follow_the_store():
    """A small search is recorded as one CUDA graph on its third call and replayed from the fourth (api.cu,
    run_replayable).  The recording must retire with everything it was built from: rows appended, a row deleted or
    overwritten, the context's settings, a buffer that had to grow — and a replay must still count its launches."""
    rng = np.random.default_rng(91)
    n, d, k = 6000, 32, 5
    x = rng.standard_normal((n, d)).astype(np.float32)
    q = rng.standard_normal((1, d)).astype(np.float32)

    def exact(rows, alive):
        dist = ((rows - q) ** 2).sum(1).astype(np.float32)
        dist[~alive] = np.inf
        return np.argsort(dist, kind="stable")[:k]

    st = z.Store(S.DT_FP32, S.M_L2SQ, d)
    st.reserve(2 * n)
    st.append(x); st.seal()
    ctx = z.Context(st)
    alive = np.ones(n, bool)
    first = ctx.search(q, k)
    assert list(first[2][0]) == list(exact(x, alive))
    for _ in range(4):                                   # direct, direct, recorded, replayed
        before = z.lib().zvb200_kernel_launches()
        again = ctx.search(q, k)
        assert z.lib().zvb200_kernel_launches() > before
        assert np.array_equal(again[2], first[2]) and np.array_equal(again[1].view(np.uint32), first[1].view(np.uint32))
        assert ctx.last_kernel() == "scan_gemv" and ctx.last_scan_ms() > 0
    best = int(first[2][0][0])
    st.set_deleted(best, True); alive[best] = False      # a writer: the next search is direct again
    for _ in range(4):
        assert list(ctx.search(q, k)[2][0]) == list(exact(x, alive))
    closer = (q + 1e-3).astype(np.float32)               # appended rows must be seen
    st.append(closer); st.seal()
    x2 = np.concatenate([x, closer]); alive2 = np.concatenate([alive, [True]])
    for _ in range(4):
        assert list(ctx.search(q, k)[2][0]) == list(exact(x2, alive2))
    st.write(7, closer[0], 7)                            # overwritten in place
    x2[7] = closer[0]
    for _ in range(4):
        assert list(ctx.search(q, k)[2][0])[:2] == [7, n]
    for _ in range(4):                                   # other arguments: another recording
        assert list(ctx.search(q, k, threshold=1e-3)[2][0][:2]) == [7, n]
        assert ctx.search(q, k, threshold=1e-3)[3][0] == 2
    big = ctx.search(rng.standard_normal((700, d)).astype(np.float32), 50)   # grows the context's buffers
    assert big[3].min() == 50
    for _ in range(4):
        assert list(ctx.search(q, k)[2][0])[:2] == [7, n]
    qd = rng.standard_normal((2, d)).astype(np.float32)  # the device-pointer form, keyed by its pointers
    import torch
    dq = torch.from_numpy(qd).cuda(); os_ = torch.empty((2, k), dtype=torch.float32, device="cuda")
    oi = torch.empty((2, k), dtype=torch.int32, device="cuda")
    want = ctx.search(qd, k)[2]
    for _ in range(5):
        oi.zero_()
        ctx.search_device(dq.data_ptr(), 2, k, os_.data_ptr(), oi.data_ptr()); ctx.sync()
        assert np.array_equal(oi.cpu().numpy().astype(np.uint32), want)
    dq.copy_(torch.from_numpy(qd[::-1].copy()))          # same pointers, other queries: the replay reads them afresh
    ctx.search_device(dq.data_ptr(), 2, k, os_.data_ptr(), oi.data_ptr()); ctx.sync()
    assert np.array_equal(oi.cpu().numpy().astype(np.uint32), want[::-1])
    ctx.close(); st.close()


@pytest.mark.parametrize("q", [1, 6, 40])
def test_threshold_leaves_fewer_than_k_in_many_partial_lists(q):
    """A threshold that most rows fail leaves the per-CTA partial lists mostly padding; the select-merge
    (merge_select_kernel, aux_kernels.cu) must still return every surviving row — its k-th smallest key is then the
    padding value itself, thousands of copies of it (found by the 2-GPU sharded-store test: rows were lost)."""
    rng = np.random.default_rng(101 + q)
    n, d, k = 70001, 72, 10
    x = rng.standard_normal((n, d)).astype(np.float32)
    qs = rng.standard_normal((q, d)).astype(np.float32)
    st = z.Store(S.DT_FP32, S.M_L2SQ, d)
    st.append(x); st.seal()
    ctx = z.Context(st)
    full = ctx.search(qs, k)
    th = float(np.median(full[1][:, k // 2]))
    got = ctx.search(qs, k, threshold=th)
    for i in range(q):
        want = int((full[1][i] <= np.float32(th)).sum())
        assert got[3][i] == want, (i, got[3][i], want)
        assert np.array_equal(got[2][i][:want], full[2][i][:want])
        assert np.array_equal(got[1][i][:want].view(np.uint32), full[1][i][:want].view(np.uint32))
    none = ctx.search(qs, k, threshold=-1.0)
    assert (none[3] == 0).all()
    ctx.close(); st.close()


@pytest.mark.parametrize("kind", ["fp16_ip", "fp32_l2", "qi8_l2"])
@pytest.mark.parametrize("q,k", [(5, 100), (40, 50), (64, 100), (65, 100)])
def test_shared_memory_pools_of_half_and_full_query_tiles(kind, q, k):
    """k > 16 keeps the tensor-core scan's top-k pools in shared memory; a launch of at most 64 queries lays them out
    for 64 query slots (one more pipeline stage), 65 queries and more for 128 — both against the oracle."""
    assert run_case(kind, 20000, 96, q, k, seed=300 + q + k, kernel=3) == "scan_mma"
