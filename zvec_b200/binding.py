"""ctypes binding over include/zvec_b200.h (no torch types cross this boundary)."""
import ctypes as C
import os

import numpy as np

DT_FP16, DT_FP32, DT_INT8 = 1, 2, 4
(METRIC_L2SQ, METRIC_L2, METRIC_IP, METRIC_COSINE,
 METRIC_QI8_L2SQ, METRIC_QI8_IP, METRIC_QI8_COSINE) = range(7)
INVALID_INDEX = 0xFFFFFFFF
INVALID_KEY = 0xFFFFFFFFFFFFFFFF
FLT_MAX = float(np.finfo(np.float32).max)

_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_path():
    return os.path.join(_HERE, "lib", "libzvec_b200.so")


class ZvecB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("zvec_b200 error %d: %s" % (code, msg))
        self.code = code


_lib = None

_SIGS = {
    "zvb200_version": (C.c_char_p, []),
    "zvb200_last_error": (C.c_char_p, []),
    "zvb200_device_count": (C.c_int, []),
    "zvb200_kernel_launches": (C.c_uint64, []),
    "zvb200_store_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "zvb200_store_create_sharded": (C.c_int, [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "zvb200_store_destroy": (C.c_int, [C.c_void_p]),
    "zvb200_store_reserve": (C.c_int, [C.c_void_p, C.c_uint64]),
    "zvb200_store_append": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "zvb200_store_append_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "zvb200_store_seal": (C.c_int, [C.c_void_p]),
    "zvb200_store_count": (C.c_uint64, [C.c_void_p]),
    "zvb200_store_set_deleted": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int]),
    "zvb200_store_set_operand_order": (C.c_int, [C.c_void_p, C.c_int]),
    "zvb200_ctx_set_operand_order": (C.c_int, [C.c_void_p, C.c_int]),
    "zvb200_store_write": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]),
    "zvb200_store_get_vector": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_size_t]),
    "zvb200_store_scan_bytes": (C.c_uint64, [C.c_void_p]),
    "zvb200_ctx_create": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "zvb200_ctx_destroy": (C.c_int, [C.c_void_p]),
    "zvb200_search": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_float, C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_search_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_float,
                                       C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_search_device_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_float,
                                               C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_ctx_sync": (C.c_int, [C.c_void_p]),
    "zvb200_ctx_stream": (C.c_void_p, [C.c_void_p]),
    "zvb200_ctx_last_scan_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "zvb200_ctx_set_kernel": (C.c_int, [C.c_void_p, C.c_int]),
    "zvb200_ctx_last_kernel": (C.c_char_p, [C.c_void_p]),
    "zvb200_ctx_set_row_limit": (C.c_int, [C.c_void_p, C.c_uint64]),
    "zvb200_comm_unique_id": (C.c_int, [C.c_void_p]),
    "zvb200_comm_init": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    "zvb200_comm_destroy": (C.c_int, [C.c_void_p]),
    "zvb200_comm_set_shard_bases": (C.c_int, [C.c_void_p, C.c_void_p]),
    "zvb200_comm_transport": (C.c_int, [C.c_void_p]),
    "zvb200_comm_status": (C.c_int, [C.c_void_p]),
    "zvb200_search_sharded_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_float,
                                               C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_ctx_last_mma_kind": (C.c_int, [C.c_void_p]),
    "zvb200_distance_matrix": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_void_p, C.c_uint64,
                                         C.c_void_p, C.c_uint32, C.c_void_p]),
    "zvb200_merge_topk_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                           C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]),
    "zvb200_fuse_rrf": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                  C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_fuse_weighted": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32,
                                       C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_void_p]),
    "zvb200_ivf_attach": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    "zvb200_ivf_set_budget_sizes": (C.c_int, [C.c_void_p, C.c_void_p]),
    "zvb200_ivf_attach_typed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.c_int, C.c_uint32]),
    "zvb200_ivf_search_split": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                          C.c_uint64, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p]),
    "zvb200_store_set_score_scale": (C.c_int, [C.c_void_p, C.c_float, C.c_void_p]),
    "zvb200_ivf_search": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.c_float,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_ivf_search_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64,
                                           C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_ivf_set_global_bases": (C.c_int, [C.c_void_p, C.c_void_p]),
    "zvb200_ivf_global_rows_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "zvb200_merge_topk_rows_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                                C.c_void_p, C.c_void_p]),
    "zvb200_reform_queries": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]),
    "zvb200_reform_queries_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]),
    "zvb200_search_fp32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_search_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_float,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_search_refined": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                        C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_kmeans": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32,
                                C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "zvb200_ivf_assign": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64,
                                    C.c_void_p]),
    "zvb200_ivf_pack": (C.c_int, [C.c_int, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]),
    "zvb200_kmeans_device": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint32,
                                       C.c_uint32, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
}


def exported_symbols():
    return sorted(_SIGS)


def lib():
    """Load libzvec_b200.so; fail loudly when it has not been built."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise ZvecB200Error(-1, "CUDA library %s is missing: run __graft_entry__.build() "
                                    "(make -C zvec_b200/csrc); there is no CPU fallback" % p)
        L = C.CDLL(p)
        for name, (res, args) in _SIGS.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise ZvecB200Error(rc, lib().zvb200_last_error().decode())


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    return a.ctypes.data_as(C.c_void_p)


def metric_from_names(metric, origin=None):
    """Reference metric name (+ origin metric of QuantizedInteger) -> ZVB200_METRIC_*."""
    plain = {"SquaredEuclidean": METRIC_L2SQ, "Euclidean": METRIC_L2, "InnerProduct": METRIC_IP,
             "Cosine": METRIC_COSINE}
    if metric == "QuantizedInteger":
        return {"SquaredEuclidean": METRIC_QI8_L2SQ, "InnerProduct": METRIC_QI8_IP,
                "Cosine": METRIC_QI8_COSINE}[origin]
    return plain[metric]


class Store:
    """Device-resident vector store on one GPU (zvb200_store_*)."""

    def __init__(self, dtype, metric, dim, device=0, devices=None):
        """devices=[0, 1, ...]: rows sharded over several GPUs of this process (zvb200_store_create_sharded);
        reserve(total rows) is then required before the first append."""
        h = C.c_void_p()
        if devices is not None and len(devices) > 1:
            devs = np.ascontiguousarray(devices, np.int32)
            _check(lib().zvb200_store_create_sharded(_ptr(devs), devs.size, dtype, metric, dim, C.byref(h)))
            device = int(devs[0])
        else:
            if devices is not None:
                device = int(devices[0])
            _check(lib().zvb200_store_create(device, dtype, metric, dim, C.byref(h)))
        self.h, self.dtype, self.metric, self.dim, self.device = h, dtype, metric, dim, device
        esz = {DT_FP32: 4, DT_FP16: 2, DT_INT8: 1}[dtype]
        self.elem_bytes = dim * esz

    def reserve(self, rows):
        _check(lib().zvb200_store_reserve(self.h, rows))

    def append(self, rows, keys=None):
        rows = np.ascontiguousarray(rows)
        n = rows.nbytes // self.elem_bytes
        assert n * self.elem_bytes == rows.nbytes, "rows are not whole elements"
        if keys is not None:
            keys = np.ascontiguousarray(keys, dtype=np.uint64)
            assert keys.size == n
        _check(lib().zvb200_store_append(self.h, _ptr(rows), n, _ptr(keys)))

    def append_device(self, dev_ptr, n):
        _check(lib().zvb200_store_append_device(self.h, C.c_void_p(dev_ptr), n))

    def seal(self):
        _check(lib().zvb200_store_seal(self.h))

    def count(self):
        return lib().zvb200_store_count(self.h)

    def scan_bytes(self):
        return lib().zvb200_store_scan_bytes(self.h)

    def set_operand_order(self, query_first):
        _check(lib().zvb200_store_set_operand_order(self.h, int(query_first)))

    def set_deleted(self, row, deleted=True):
        _check(lib().zvb200_store_set_deleted(self.h, row, int(deleted)))

    def get_vector(self, row):
        out = np.zeros(self.elem_bytes, np.uint8)
        _check(lib().zvb200_store_get_vector(self.h, row, _ptr(out), out.nbytes))
        return out

    def ivf_set_budget_sizes(self, list_sizes):
        ls = None if list_sizes is None else np.ascontiguousarray(list_sizes, np.uint64)
        _check(lib().zvb200_ivf_set_budget_sizes(self.h, _ptr(ls)))

    def ivf_set_global_bases(self, list_global_base):
        gb = None if list_global_base is None else np.ascontiguousarray(list_global_base, np.uint64)
        _check(lib().zvb200_ivf_set_global_bases(self.h, _ptr(gb)))

    def write(self, row, vec, key):
        """overwrite one row in place (FlatStreamerEntity::add_vector_with_id, id < count)"""
        vec = np.ascontiguousarray(vec)
        _check(lib().zvb200_store_write(self.h, row, _ptr(vec), key))

    def ivf_attach(self, centroids, list_offsets):
        """centroids: nlist elements in the store's layout; rows [list_offsets[l], list_offsets[l+1]) = list l."""
        centroids = np.ascontiguousarray(centroids)
        nlist = centroids.nbytes // self.elem_bytes
        offs = np.ascontiguousarray(list_offsets, dtype=np.uint64)
        assert offs.size == nlist + 1
        _check(lib().zvb200_ivf_attach(self.h, _ptr(centroids), nlist, _ptr(offs)))

    def ivf_attach_typed(self, centroids, list_offsets, c_dtype, c_metric, c_dim):
        """centroids in a layout of their own (IVFBuilder with a quantizer_class): [nlist, c_elem_bytes] bytes"""
        centroids = np.ascontiguousarray(centroids)
        offs = np.ascontiguousarray(list_offsets, dtype=np.uint64)
        nlist = offs.size - 1
        assert centroids.nbytes % nlist == 0
        _check(lib().zvb200_ivf_attach_typed(self.h, _ptr(centroids), nlist, _ptr(offs), c_dtype, c_metric, c_dim))

    def set_score_scale(self, scale=1.0, list_scales=None):
        """scores leave the scans multiplied by `scale`, or by list_scales[list of the row] (IVFEntity's norm_val)"""
        ls = None if list_scales is None else np.ascontiguousarray(list_scales, np.float32)
        _check(lib().zvb200_store_set_score_scale(self.h, scale, _ptr(ls)))

    def close(self):
        if self.h:
            lib().zvb200_store_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Context:
    """Per-thread search context (zvb200_ctx_*)."""

    def __init__(self, store):
        h = C.c_void_p()
        _check(lib().zvb200_ctx_create(store.h, C.byref(h)))
        self.h, self.store = h, store

    def set_kernel(self, which):
        _check(lib().zvb200_ctx_set_kernel(self.h, which))

    def last_kernel(self):
        return lib().zvb200_ctx_last_kernel(self.h).decode()

    def set_row_limit(self, rows):
        """searches see rows [0, rows) only (0 = all): a snapshot next to concurrent appends"""
        _check(lib().zvb200_ctx_set_row_limit(self.h, rows))

    def search_sharded_device(self, comm, q_ptr, q, k, out_scores_ptr, out_rows_ptr, threshold=FLT_MAX, filter_ptr=None):
        """collective over the communicator's ranks: local scan -> exchange -> merge, on this context's stream"""
        _check(lib().zvb200_search_sharded_device(self.h, comm.h, C.c_void_p(q_ptr), q, k, threshold,
                                                  C.c_void_p(filter_ptr) if filter_ptr else None,
                                                  C.c_void_p(out_scores_ptr), C.c_void_p(out_rows_ptr)))

    def last_mma_kind(self):
        return lib().zvb200_ctx_last_mma_kind(self.h)

    def last_scan_ms(self):
        ms = C.c_float()
        _check(lib().zvb200_ctx_last_scan_ms(self.h, C.byref(ms)))
        return ms.value

    def stream(self):
        return lib().zvb200_ctx_stream(self.h)

    def sync(self):
        _check(lib().zvb200_ctx_sync(self.h))

    def search(self, queries, k, threshold=FLT_MAX, filter_bitmap=None):
        """Host queries (stored layout) -> (keys u64, scores f32, index u32, counts u32)."""
        queries = np.ascontiguousarray(queries)
        q = queries.nbytes // self.store.elem_bytes
        assert q * self.store.elem_bytes == queries.nbytes
        keys = np.empty((q, k), np.uint64)
        scores = np.empty((q, k), np.float32)
        index = np.empty((q, k), np.uint32)
        counts = np.empty(q, np.uint32)
        if filter_bitmap is not None:
            filter_bitmap = np.ascontiguousarray(filter_bitmap, dtype=np.uint32)
        _check(lib().zvb200_search(self.h, _ptr(queries), q, k, threshold, _ptr(filter_bitmap),
                                   _ptr(keys), _ptr(scores), _ptr(index), _ptr(counts)))
        return keys, scores, index, counts

    def ivf_search(self, queries, k, nprobe, max_scan_count, threshold=FLT_MAX, filter_bitmap=None):
        """IVFSearcher::search_impl semantics -> (keys u64, scores f32, index u32, counts u32)."""
        queries = np.ascontiguousarray(queries)
        if filter_bitmap is not None:
            filter_bitmap = np.ascontiguousarray(filter_bitmap, np.uint32)
        q = queries.nbytes // self.store.elem_bytes
        keys = np.empty((q, k), np.uint64)
        scores = np.empty((q, k), np.float32)
        index = np.empty((q, k), np.uint32)
        counts = np.empty(q, np.uint32)
        _check(lib().zvb200_ivf_search(self.h, _ptr(queries), q, k, nprobe, max_scan_count, threshold,
                                       _ptr(filter_bitmap), _ptr(keys), _ptr(scores), _ptr(index), _ptr(counts)))
        return keys, scores, index, counts

    def ivf_search_split(self, coarse_queries, list_queries, k, nprobe, max_scan_count, threshold=FLT_MAX,
                         filter_bitmap=None):
        """zvb200_ivf_search for a store whose centroids have a layout of their own (ivf_attach_typed)."""
        coarse_queries = np.ascontiguousarray(coarse_queries)
        list_queries = np.ascontiguousarray(list_queries)
        if filter_bitmap is not None:
            filter_bitmap = np.ascontiguousarray(filter_bitmap, np.uint32)
        q = list_queries.nbytes // self.store.elem_bytes
        keys = np.empty((q, k), np.uint64)
        scores = np.empty((q, k), np.float32)
        index = np.empty((q, k), np.uint32)
        counts = np.empty(q, np.uint32)
        _check(lib().zvb200_ivf_search_split(self.h, _ptr(coarse_queries), _ptr(list_queries), q, k, nprobe,
                                             max_scan_count, threshold, _ptr(filter_bitmap), _ptr(keys), _ptr(scores),
                                             _ptr(index), _ptr(counts)))
        return keys, scores, index, counts

    def ivf_search_device(self, queries_ptr, q, k, nprobe, max_scan_count, out_scores_ptr, out_index_ptr,
                          threshold=FLT_MAX, filter_ptr=None):
        _check(lib().zvb200_ivf_search_device(self.h, C.c_void_p(queries_ptr), q, k, nprobe, max_scan_count, threshold,
                                              C.c_void_p(filter_ptr) if filter_ptr else None,
                                              C.c_void_p(out_scores_ptr), C.c_void_p(out_index_ptr)))

    def ivf_global_rows_device(self, index_ptr, count, out_rows_ptr):
        _check(lib().zvb200_ivf_global_rows_device(self.h, C.c_void_p(index_ptr), count, C.c_void_p(out_rows_ptr)))

    def merge_topk_rows_device(self, scores_ptr, rows_ptr, g, q, k, out_scores_ptr, out_rows_ptr):
        _check(lib().zvb200_merge_topk_rows_device(self.h, C.c_void_p(scores_ptr), C.c_void_p(rows_ptr), g, q, k,
                                                   C.c_void_p(out_scores_ptr), C.c_void_p(out_rows_ptr)))

    def fuse_rrf(self, keys, counts, topn, rank_constant=60):
        """RrfReRanker.rerank over F per-field lists: keys [F][q][k] u64, counts [F][q] -> (keys, fp64 scores, counts)."""
        keys = np.ascontiguousarray(keys, np.uint64)
        counts = np.ascontiguousarray(counts, np.uint32)
        F, q, k = keys.shape
        ok = np.empty((q, topn), np.uint64); os_ = np.empty((q, topn), np.float64); oc = np.empty(q, np.uint32)
        _check(lib().zvb200_fuse_rrf(self.h, _ptr(keys), _ptr(counts), F, q, k, topn, rank_constant,
                                     _ptr(ok), _ptr(os_), _ptr(oc)))
        return ok, os_, oc

    def fuse_weighted(self, keys, scores, counts, topn, fuse_metric, weights, negate=False):
        """WeightedReRanker.rerank: scores [F][q][k] f32 are Doc.score (negate=True: internal InnerProduct scores)."""
        keys = np.ascontiguousarray(keys, np.uint64)
        scores = np.ascontiguousarray(scores, np.float32)
        counts = np.ascontiguousarray(counts, np.uint32)
        weights = np.ascontiguousarray(weights, np.float64)
        F, q, k = keys.shape
        ok = np.empty((q, topn), np.uint64); os_ = np.empty((q, topn), np.float64); oc = np.empty(q, np.uint32)
        _check(lib().zvb200_fuse_weighted(self.h, _ptr(keys), _ptr(scores), _ptr(counts), F, q, k, topn, fuse_metric,
                                          _ptr(weights), 1 if negate else 0, _ptr(ok), _ptr(os_), _ptr(oc)))
        return ok, os_, oc

    def _outs(self, q, k):
        return (np.empty((q, k), np.uint64), np.empty((q, k), np.float32), np.empty((q, k), np.uint32),
                np.empty(q, np.uint32))

    def reform_queries(self, queries_fp32):
        """IndexReformer::transform of the store's type on the device: [q, dim] fp32 -> [q, elem_bytes] u8."""
        queries_fp32 = np.ascontiguousarray(queries_fp32, np.float32)
        q, dim = queries_fp32.shape
        out = np.empty((q, self.store.elem_bytes), np.uint8)
        _check(lib().zvb200_reform_queries(self.h, _ptr(queries_fp32), q, dim, _ptr(out)))
        return out

    def search_fp32(self, queries_fp32, k, threshold=FLT_MAX, filter_bitmap=None):
        """Original fp32 queries: reformer + scan on the device."""
        queries_fp32 = np.ascontiguousarray(queries_fp32, np.float32)
        q, dim = queries_fp32.shape
        if filter_bitmap is not None:
            filter_bitmap = np.ascontiguousarray(filter_bitmap, np.uint32)
        keys, scores, index, counts = self._outs(q, k)
        _check(lib().zvb200_search_fp32(self.h, _ptr(queries_fp32), q, dim, k, threshold, _ptr(filter_bitmap),
                                        _ptr(keys), _ptr(scores), _ptr(index), _ptr(counts)))
        return keys, scores, index, counts

    def search_rows(self, queries, rows, k, threshold=FLT_MAX, filter_bitmap=None):
        """FlatStreamer::search_bf_by_p_keys_impl with stored rows for keys: rows [q, m] u32."""
        queries = np.ascontiguousarray(queries)
        rows = np.ascontiguousarray(rows, np.uint32)
        q, m = rows.shape
        assert q * self.store.elem_bytes == queries.nbytes
        if filter_bitmap is not None:
            filter_bitmap = np.ascontiguousarray(filter_bitmap, np.uint32)
        keys, scores, index, counts = self._outs(q, k)
        _check(lib().zvb200_search_rows(self.h, _ptr(queries), q, _ptr(rows), m, k, threshold, _ptr(filter_bitmap),
                                        _ptr(keys), _ptr(scores), _ptr(index), _ptr(counts)))
        return keys, scores, index, counts

    def search_refined(self, exact_ctx, queries_fp32, k, scale_factor, filter_bitmap=None):
        """The refiner flow (self = the quantised index's context, exact_ctx = the reference Flat index's)."""
        queries_fp32 = np.ascontiguousarray(queries_fp32, np.float32)
        q, dim = queries_fp32.shape
        if filter_bitmap is not None:
            filter_bitmap = np.ascontiguousarray(filter_bitmap, np.uint32)
        keys, scores, index, counts = self._outs(q, k)
        _check(lib().zvb200_search_refined(self.h, exact_ctx.h, _ptr(queries_fp32), q, dim, k, scale_factor,
                                           _ptr(filter_bitmap), _ptr(keys), _ptr(scores), _ptr(index), _ptr(counts)))
        return keys, scores, index, counts

    def search_device_noalloc(self, q_ptr, q, k, out_scores_ptr, out_index_ptr, threshold=FLT_MAX, filter_ptr=None):
        """zvb200_search_device with every scratch buffer in the library's static device array (no allocation)"""
        _check(lib().zvb200_search_device_noalloc(self.h, C.c_void_p(q_ptr), q, k, threshold,
                                                  C.c_void_p(filter_ptr) if filter_ptr else None,
                                                  C.c_void_p(out_scores_ptr), C.c_void_p(out_index_ptr)))

    def search_device(self, q_ptr, q, k, out_scores_ptr, out_index_ptr, threshold=FLT_MAX, filter_ptr=None):
        _check(lib().zvb200_search_device(self.h, C.c_void_p(q_ptr), q, k, threshold,
                                          C.c_void_p(filter_ptr) if filter_ptr else None,
                                          C.c_void_p(out_scores_ptr), C.c_void_p(out_index_ptr)))

    def merge_topk_device(self, scores_ptr, index_ptr, g, shard_base, q, k, out_scores_ptr, out_rows_ptr):
        sb = np.ascontiguousarray(shard_base, dtype=np.uint64)
        _check(lib().zvb200_merge_topk_device(self.h, C.c_void_p(scores_ptr), C.c_void_p(index_ptr), g,
                                              _ptr(sb), q, k, C.c_void_p(out_scores_ptr),
                                              C.c_void_p(out_rows_ptr)))

    def close(self):
        if self.h:
            lib().zvb200_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


COMM_ID_BYTES = 128


def comm_unique_id():
    """128 opaque bytes made on one rank (ncclGetUniqueId); the caller carries them to the other ranks"""
    buf = np.zeros(COMM_ID_BYTES, np.uint8)
    _check(lib().zvb200_comm_unique_id(_ptr(buf)))
    return buf


class Comm:
    """One rank of a multi-process row-sharded search (zvb200_comm_*), bound to a context's stream."""

    def __init__(self, ctx, rank, world, unique_id, shard_bases):
        h = C.c_void_p()
        uid = np.ascontiguousarray(unique_id, np.uint8)
        assert uid.size == COMM_ID_BYTES
        _check(lib().zvb200_comm_init(ctx.h, rank, world, _ptr(uid), C.byref(h)))
        self.h, self.rank, self.world = h, rank, world
        bases = np.ascontiguousarray(shard_bases, np.uint64)
        assert bases.size == world
        _check(lib().zvb200_comm_set_shard_bases(self.h, _ptr(bases)))

    def transport(self):
        return "peer mailboxes (CUDA IPC over NVLink)" if lib().zvb200_comm_transport(self.h) == 1 else "ncclAllGather"

    def status(self):
        _check(lib().zvb200_comm_status(self.h))

    def close(self):
        if self.h:
            lib().zvb200_comm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def distance_matrix(metric, dtype, dim, rows, queries, device=0):
    """out[j, i] = dist(rows[i], queries[j]) — the ailego <T,1,1>::Compute contract."""
    rows = np.ascontiguousarray(rows)
    queries = np.ascontiguousarray(queries)
    esz = {DT_FP32: 4, DT_FP16: 2, DT_INT8: 1}[dtype]
    n = rows.nbytes // (dim * esz)
    nq = queries.nbytes // (dim * esz)
    out = np.empty((nq, n), np.float32)
    _check(lib().zvb200_distance_matrix(device, metric, dtype, dim, _ptr(rows), n, _ptr(queries), nq, _ptr(out)))
    return out


def ivf_assign(rows, centroids, dtype, metric, dim, device=0):
    """nearest centroid of every row (IVFBuilder::label): rows / centroids in the stored layout -> labels u32[n]"""
    rows, centroids = np.ascontiguousarray(rows), np.ascontiguousarray(centroids)
    esz = dim * {DT_FP32: 4, DT_FP16: 2, DT_INT8: 1}[dtype]
    n, k = rows.nbytes // esz, centroids.nbytes // esz
    labels = np.zeros(n, np.uint32)
    _check(lib().zvb200_ivf_assign(device, dtype, metric, dim, _ptr(centroids), k, _ptr(rows), n, _ptr(labels)))
    return labels


def ivf_pack(labels, nlist, device=0):
    """rows grouped by posting list, ascending inside a list (IVFBuilder::dump_index's order) -> (order u32[n],
    list_offsets u64[nlist + 1])"""
    labels = np.ascontiguousarray(labels, np.uint32)
    order = np.zeros(labels.size, np.uint32)
    offs = np.zeros(nlist + 1, np.uint64)
    _check(lib().zvb200_ivf_pack(device, _ptr(labels), labels.size, nlist, _ptr(order), _ptr(offs)))
    return order, offs


def kmeans(rows, centroids, metric=None, max_iterations=20, epsilon=float(np.finfo(np.float32).eps), device=0,
           rows_device_ptr=None, n=None):
    """Lloyd iterations on the device (zvb200_kmeans): rows [n, dim] fp32 (or a device pointer + n), centroids [k, dim]
    fp32 = the initial centroids.  -> (centroids, labels u32[n], follows u64[k], sse, iterations)"""
    cent = np.ascontiguousarray(centroids, np.float32).copy()
    k, dim = cent.shape
    metric = METRIC_L2SQ if metric is None else metric
    if rows_device_ptr is None:
        rows = np.ascontiguousarray(rows, np.float32)
        n = rows.shape[0]
    labels = np.zeros(n, np.uint32)
    follows = np.zeros(k, np.uint64)
    sse = C.c_double()
    iters = C.c_uint32()
    if rows_device_ptr is None:
        _check(lib().zvb200_kmeans(device, metric, _ptr(rows), n, dim, _ptr(cent), k, max_iterations, epsilon,
                                   _ptr(labels), _ptr(follows), C.addressof(sse), C.addressof(iters)))
    else:
        _check(lib().zvb200_kmeans_device(device, metric, rows_device_ptr, n, dim, _ptr(cent), k, max_iterations, epsilon,
                                          _ptr(labels), _ptr(follows), C.addressof(sse), C.addressof(iters)))
    return cent, labels, follows, sse.value, iters.value
