"""Test-only helpers: oracle loaders (oracle/ is the CHECKER, never the product path),
seeded data generators and the parity comparison with the documented tie rule."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

DT_FP16, DT_FP32, DT_INT8 = 1, 2, 4
M_L2SQ, M_L2, M_IP, M_COSINE, M_QI8_L2SQ, M_QI8_IP, M_QI8_COSINE = range(7)


def P(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


_oracle = None


def oracle():
    """The scalar CPU restatement (oracle/zvec_oracle.cc), built on demand with g++."""
    global _oracle
    if _oracle is None:
        so = os.path.join(ORACLE_DIR, "libzvec_oracle.so")
        src = os.path.join(ORACLE_DIR, "zvec_oracle.cc")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "libzvec_oracle.so"], stdout=subprocess.DEVNULL)
        L = C.CDLL(so)
        L.zo_distance.restype = C.c_float
        L.zo_distance.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]
        L.zo_normalize_score.restype = C.c_float
        L.zo_normalize_score.argtypes = [C.c_int, C.c_float]
        L.zo_distances.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32,
                                   C.c_void_p, C.c_void_p]
        L.zo_flat_search.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_float,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.zo_quantize_record_int8.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
        L.zo_cosine_normalize_fp32.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        L.zo_fp32_to_fp16.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        L.zo_fp16_to_fp32.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        _oracle = L
    return _oracle


def cpu_has(flag):
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return flag in line.split()
    except OSError:
        pass
    return False


_ref = False


def ref():
    """The reference itself (oracle/_ref, compiled from /root/reference by oracle/build_ref.py);
    None when the prebuilt library is absent or the host CPU cannot run it."""
    global _ref
    if _ref is False:
        _ref = None
        cands = []
        if cpu_has("avx512f") and cpu_has("avx512bw") and cpu_has("avx512vl") and cpu_has("avx512dq"):
            cands.append("libzvec_ref_skylake-avx512.so")
        if cpu_has("avx2") and cpu_has("fma"):
            cands.append("libzvec_ref_haswell.so")
        for c in cands:
            p = os.path.join(ORACLE_DIR, "_ref", c)
            if os.path.exists(p):
                # RTLD_GLOBAL: a plugin dlopen'ed later resolves the framework symbols from here, as it
                # would from a zvec host binary
                L = C.CDLL(p, mode=C.RTLD_GLOBAL)
                L.zref_index_create.restype = C.c_void_p
                L.zref_index_create.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_int,
                                                C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_char_p]
                L.zref_last_error.restype = C.c_char_p
                L.zref_index_destroy.argtypes = [C.c_void_p]
                L.zref_index_info.argtypes = [C.c_void_p] + [C.c_void_p] * 5 + [C.c_uint32]
                L.zref_index_stored.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
                L.zref_index_transform_queries.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
                L.zref_index_search.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                                C.c_uint32, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
                L.zref_index_distances.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
                L.zref_metric_distances.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_uint32, C.c_void_p,
                                                    C.c_uint64, C.c_void_p, C.c_void_p]
                L.zref_load_plugin.argtypes = [C.c_char_p]
                L.zref_has_searcher.argtypes = [C.c_char_p]
                L.zref_set_filter.argtypes = [C.c_uint64, C.c_uint64]
                L.zref_set_threshold.argtypes = [C.c_int, C.c_float]
                L.zref_ivf_export.argtypes = [C.c_void_p] + [C.c_void_p] * 6 + [C.c_char_p, C.c_uint32] + [C.c_void_p] * 4
                L._name = c
                _ref = L
                break
    return _ref


PLUGIN = os.path.join(ROOT, "integration", "_build", "libzvec_b200_plugin.so")
_plugin_loaded = False


def load_plugin():
    """Load the GPU plugin through the reference's own IndexPluginBroker; False if unavailable."""
    global _plugin_loaded
    if _plugin_loaded:
        return True
    if ref() is None or not os.path.exists(PLUGIN):
        return False
    if ref().zref_load_plugin(PLUGIN.encode()) != 0:
        raise RuntimeError(ref().zref_last_error().decode())
    _plugin_loaded = True
    return True


class RefIndex:
    """A Flat/IVF index built and searched by the reference's own classes."""

    def __init__(self, rows_fp32_or_native, dtype, dim, metric, converter="", kind="Flat", column_major=False,
                 build_params="", search_params=""):
        L = ref()
        self.L = L
        self.rows = np.ascontiguousarray(rows_fp32_or_native)
        esz = {DT_FP32: 4, DT_FP16: 2, DT_INT8: 1}[dtype]
        n = self.rows.nbytes // (dim * esz)
        self.h = L.zref_index_create(kind.encode(), P(self.rows), n, dim, dtype, metric.encode(),
                                     converter.encode(), int(column_major), build_params.encode(),
                                     search_params.encode())
        if not self.h:
            raise RuntimeError("reference index build failed: %s" % L.zref_last_error().decode())
        dt, dm, el, nn = C.c_int(), C.c_uint32(), C.c_uint32(), C.c_uint64()
        name = C.create_string_buffer(64)
        L.zref_index_info(self.h, C.byref(dt), C.byref(dm), C.byref(el), C.byref(nn), name, 64)
        self.dtype, self.dim, self.elem, self.n = dt.value, dm.value, el.value, nn.value
        self.metric_name = name.value.decode()
        self.in_elem = dim * esz

    def stored(self):
        out = np.zeros((self.n, self.elem), np.uint8)
        assert self.L.zref_index_stored(self.h, P(out), out.nbytes) == 0
        return out

    def transform_queries(self, queries):
        queries = np.ascontiguousarray(queries)
        q = queries.nbytes // self.in_elem
        out = np.zeros((q, self.elem), np.uint8)
        r = self.L.zref_index_transform_queries(self.h, P(queries), q, P(out), out.nbytes)
        assert r == self.elem, r
        return out

    def search(self, queries, k, batch=1, threads=1, normalize=False):
        queries = np.ascontiguousarray(queries)
        q = queries.nbytes // self.in_elem
        keys = np.zeros((q, k), np.uint64)
        scores = np.zeros((q, k), np.float32)
        index = np.zeros((q, k), np.uint32)
        counts = np.zeros(q, np.uint32)
        r = self.L.zref_index_search(self.h, P(queries), q, k, batch, threads, int(normalize), P(keys),
                                     P(scores), P(index), P(counts))
        if r != 0:
            import zvec_b200
            raise AssertionError("search_impl returned %d (%s)" % (r, zvec_b200.lib().zvb200_last_error().decode()))
        return keys, scores, index, counts

    def use_searcher(self, name, search_params=""):
        """search the same index file with another registered searcher class"""
        self.L.zref_index_use_searcher.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        if self.L.zref_index_use_searcher(self.h, name.encode(), search_params.encode()) != 0:
            raise RuntimeError(self.L.zref_last_error().decode())

    def use_streamer(self, name, search_params=""):
        """search the same index file through a registered IndexStreamer class (IVFStreamer and the like);
        an empty name goes back to the searcher"""
        self.L.zref_index_use_streamer.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        if self.L.zref_index_use_streamer(self.h, name.encode(), search_params.encode()) != 0:
            raise RuntimeError(self.L.zref_last_error().decode())

    def distances(self, stored_rows, stored_query):
        stored_rows = np.ascontiguousarray(stored_rows)
        n = stored_rows.nbytes // self.elem
        out = np.zeros(n, np.float32)
        assert self.L.zref_index_distances(self.h, P(stored_rows), n, P(np.ascontiguousarray(stored_query)), P(out)) == 0
        return out

    def ivf_export(self):
        """-> dict(nlist, centroids [nlist, c_elem] u8, c_dtype, c_dim, c_metric, list_offsets u64[nlist+1],
        rows [n, elem] u8 in list order, keys u64[n]) read with the reference's own IVFEntity."""
        nl, n, el, ce, cdt, cdm = C.c_uint32(), C.c_uint64(), C.c_uint32(), C.c_uint32(), C.c_int(), C.c_uint32()
        name = C.create_string_buffer(64)
        r = self.L.zref_ivf_export(self.h, C.byref(nl), C.byref(n), C.byref(el), C.byref(ce), C.byref(cdt),
                                   C.byref(cdm), name, 64, None, None, None, None)
        assert r == 0, self.L.zref_last_error()
        cent = np.zeros((nl.value, ce.value), np.uint8)
        offs = np.zeros(nl.value + 1, np.uint64)
        rows = np.zeros((n.value, el.value), np.uint8)
        keys = np.zeros(n.value, np.uint64)
        r = self.L.zref_ivf_export(self.h, C.byref(nl), C.byref(n), C.byref(el), C.byref(ce), C.byref(cdt),
                                   C.byref(cdm), name, 64, P(cent), P(offs), P(rows), P(keys))
        assert r == 0, self.L.zref_last_error()
        return dict(nlist=nl.value, centroids=cent, c_dtype=cdt.value, c_dim=cdm.value,
                    c_metric=name.value.decode(), list_offsets=offs, rows=rows, keys=keys)

    def close(self):
        if self.h:
            self.L.zref_index_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CiIndex:
    """core_interface::Index (src/core/interface) driven like tests/core/interface/index_interface_test.cc:
    IndexFactory::CreateAndInitIndex -> Open -> Add -> (Train) -> Search.  Which streamer class answers is decided
    by NAME inside the reference (flat_index.cc:31, ivf_index.cc:38-39): the stock one, or — after a plugin's
    zvb200_install_overrides() — the GPU one."""

    def __init__(self, kind, dim, metric, quantizer="", nlist=32, path="/tmp/zref_ci.index", dtype=DT_FP32,
                 open_existing=False):
        L = ref()
        L.zref_ci_create.restype = C.c_void_p
        L.zref_ci_create.argtypes = [C.c_char_p, C.c_int, C.c_uint32, C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
        L.zref_ci_add.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32]
        L.zref_ci_train.argtypes = [C.c_void_p]
        L.zref_ci_count.argtypes = [C.c_void_p]
        L.zref_ci_count.restype = C.c_uint32
        L.zref_ci_fetch.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.zref_ci_destroy.argtypes = [C.c_void_p]
        L.zref_ci_search.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_uint32,
                                     C.c_int, C.c_float, C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p]
        L.zref_set_batch_fn.argtypes = [C.c_void_p]
        self.L, self.dim, self.dtype = L, dim, dtype
        self.in_es = dim * {DT_FP32: 4, DT_FP16: 2, DT_INT8: 1}[dtype]
        self.h = L.zref_ci_create(kind.encode(), dtype, dim, metric.encode(), quantizer.encode(), nlist, path.encode(),
                                  int(open_existing))
        if not self.h:
            raise RuntimeError("core_interface index: %s" % L.zref_last_error().decode())

    def add(self, rows, first_id=0):
        rows = np.ascontiguousarray(rows)
        n = rows.nbytes // self.in_es
        r = self.L.zref_ci_add(self.h, P(rows), n, first_id)
        if r != 0:
            raise RuntimeError("Index::Add -> %d" % r)

    def train(self):
        return self.L.zref_ci_train(self.h)

    def count(self):
        return self.L.zref_ci_count(self.h)

    def fetch(self, doc_id):
        out = np.zeros(self.in_es, np.uint8)
        r = self.L.zref_ci_fetch(self.h, doc_id, P(out), out.nbytes)
        assert r == self.in_es, (r, self.L.zref_last_error())
        return out

    def search(self, queries, k, batch=False, bf_pks=None, linear=False, radius=0.0, refiner=None, scale=0.0,
               fetch_vector=False):
        """-> (keys u64 [q,k], scores f32 [q,k] (user-facing: normalised), counts u32[q][, vectors u8 [q,k,in_es]])"""
        queries = np.ascontiguousarray(queries)
        q = queries.nbytes // self.in_es
        keys = np.zeros((q, k), np.uint64); scores = np.zeros((q, k), np.float32); counts = np.zeros(q, np.uint32)
        vec = np.zeros((q, k, self.in_es), np.uint8) if fetch_vector else None
        pk = None if bf_pks is None else np.ascontiguousarray(bf_pks, np.uint64)
        r = self.L.zref_ci_search(self.h, P(queries), q, k, int(batch), P(pk), 0 if pk is None else pk.size,
                                  int(linear), float(radius), refiner.h if refiner is not None else None, float(scale),
                                  int(fetch_vector), P(keys), P(scores), P(counts), P(vec))
        if r != 0:
            import zvec_b200
            raise RuntimeError("Index::Search -> %d (%s / %s)" % (r, self.L.zref_last_error().decode(),
                                                                  zvec_b200.lib().zvb200_last_error().decode()))
        return (keys, scores, counts, vec) if fetch_vector else (keys, scores, counts)

    def close(self):
        if self.h:
            self.L.zref_ci_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def plugin_handle():
    """ctypes handle of the loaded plugin (zvb200_install_overrides, zvb200_index_search_batch)"""
    assert load_plugin()
    plug = C.CDLL(PLUGIN)
    ref().zref_set_batch_fn.argtypes = [C.c_void_p]
    ref().zref_set_batch_fn(C.cast(plug.zvb200_index_search_batch, C.c_void_p))
    return plug


# ---------------------------------------------------------------------------------------
# oracle-side helpers
# ---------------------------------------------------------------------------------------
def oracle_flat_search(metric, dtype, stored_rows, dim, queries, k, threshold=None, keys=None, drop=None):
    O = oracle()
    stored_rows = np.ascontiguousarray(stored_rows)
    queries = np.ascontiguousarray(queries)
    elem = dim * {DT_FP32: 4, DT_FP16: 2, DT_INT8: 1}[dtype]
    n = stored_rows.nbytes // elem
    q = queries.nbytes // elem
    okeys = np.zeros((q, k), np.uint64)
    oscores = np.zeros((q, k), np.float32)
    oindex = np.zeros((q, k), np.uint32)
    ocounts = np.zeros(q, np.uint32)
    th = float(np.finfo(np.float32).max) if threshold is None else threshold
    kk = None if keys is None else np.ascontiguousarray(keys, dtype=np.uint64)
    dd = None if drop is None else np.ascontiguousarray(drop, dtype=np.uint8)
    O.zo_flat_search(metric, dtype, P(stored_rows), n, dim, elem, P(kk), P(dd), P(queries), q, k, th,
                     P(okeys), P(oscores), P(oindex), P(ocounts))
    return okeys, oscores, oindex, ocounts


def oracle_distances(metric, dtype, stored_rows, dim, query):
    O = oracle()
    stored_rows = np.ascontiguousarray(stored_rows)
    elem = dim * {DT_FP32: 4, DT_FP16: 2, DT_INT8: 1}[dtype]
    n = stored_rows.nbytes // elem
    out = np.zeros(n, np.float32)
    O.zo_distances(metric, dtype, P(stored_rows), n, dim, elem, P(np.ascontiguousarray(query)), P(out))
    return out


def quantize_int8(rows_fp32, is_euclidean=True):
    """Int8StreamingConverter / IntegerStreamingReformer layout via the oracle restatement."""
    O = oracle()
    rows = np.ascontiguousarray(rows_fp32, dtype=np.float32)
    n, d = rows.shape
    out = np.zeros((n, d + 16), np.uint8)
    for i in range(n):
        O.zo_quantize_record_int8(C.c_void_p(rows[i].ctypes.data), d, int(is_euclidean), C.c_void_p(out[i].ctypes.data))
    return out


def cosine_normalize_fp32(rows_fp32):
    O = oracle()
    rows = np.ascontiguousarray(rows_fp32, dtype=np.float32)
    n, d = rows.shape
    out = np.zeros((n, d + 1), np.float32)
    for i in range(n):
        O.zo_cosine_normalize_fp32(C.c_void_p(rows[i].ctypes.data), d, C.c_void_p(out[i].ctypes.data))
    return out


# ---------------------------------------------------------------------------------------
# parity rule (DESIGN.md "tie-break")
# ---------------------------------------------------------------------------------------
def check_topk_parity(got_index, got_scores, exp_index, exp_scores, all_dist=None, exact=False, rtol=1e-4,
                      what=""):
    """Compare one query's result lists.

    exact=True (integer paths): score sequences must be bit-identical; indices must match as
    SETS inside every run of equal scores, except the run that touches position k-1, whose
    members must only come from rows having exactly that score (checked with all_dist).
    exact=False (fp32/fp16): scores within rtol; index order may differ only inside groups of
    adjacent results whose scores are within rtol of each other (near-ties), and boundary rows
    must have a true distance within rtol of the k-th score.
    """
    got_index = np.asarray(got_index); exp_index = np.asarray(exp_index)
    got_scores = np.asarray(got_scores, np.float32); exp_scores = np.asarray(exp_scores, np.float32)
    valid_e = exp_index != 0xFFFFFFFF
    valid_g = got_index != 0xFFFFFFFF
    assert valid_e.sum() == valid_g.sum(), "%s: result count %d != %d" % (what, valid_g.sum(), valid_e.sum())
    n = int(valid_e.sum())
    if n == 0:
        return
    gs, es, gi, ei = got_scores[:n], exp_scores[:n], got_index[:n], exp_index[:n]
    assert np.all(np.diff(gs) >= 0), "%s: scores not ascending" % what
    if exact:
        assert np.array_equal(gs.view(np.uint32), es.view(np.uint32)), \
            "%s: scores differ bitwise\n got %s\n exp %s" % (what, gs, es)
        last = es[n - 1]
        for s in np.unique(es):
            a = set(gi[gs == s].tolist()); b = set(ei[es == s].tolist())
            if s == last:
                if all_dist is not None:
                    cand = set(np.nonzero(all_dist == s)[0].tolist())
                    assert a <= cand, "%s: boundary ids %s not at the k-th score" % (what, a - cand)
                else:
                    assert len(a) == len(b)
            else:
                assert a == b, "%s: ids differ at score %r: %s vs %s" % (what, s, a, b)
        # inside a run of equal scores our order is by row index (documented tie-break)
        for s in np.unique(gs):
            run = gi[gs == s]
            assert np.all(np.diff(run.astype(np.int64)) > 0), "%s: tie-break not by row index" % what
        return
    denom = np.maximum(np.abs(es), 1e-30)
    assert np.all(np.abs(gs - es) <= rtol * np.maximum(denom, 1e-6) + 1e-7), \
        "%s: scores beyond rtol\n got %s\n exp %s" % (what, gs, es)
    if np.array_equal(gi, ei):
        return
    # differences allowed only among near-ties
    tol = rtol * max(abs(float(es[n - 1])), 1e-6) + 1e-7
    for pos in range(n):
        if gi[pos] == ei[pos]:
            continue
        # the row we returned must be (near-)tied with the expected score at this position
        if all_dist is not None:
            assert abs(float(all_dist[gi[pos]]) - float(es[pos])) <= 2 * tol, \
                "%s: pos %d id %d (true %r) vs expected id %d (%r): not a near-tie" % (
                    what, pos, gi[pos], all_dist[gi[pos]], ei[pos], es[pos])
        else:
            assert abs(float(gs[pos]) - float(es[pos])) <= 2 * tol
