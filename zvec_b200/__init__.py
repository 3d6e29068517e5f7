"""zvec_b200 — B200-native exhaustive-scan path (distance + fused top-k) for zvec.

The product is the C-ABI shared library ``zvec_b200/lib/libzvec_b200.so``
(``include/zvec_b200.h``).  This package is the thin ctypes binding used by the
tests and bench; it never computes anything itself and raises if the CUDA
library is missing (there is no CPU fallback).
"""
from .binding import (  # noqa: F401
    DT_FP16, DT_FP32, DT_INT8,
    METRIC_L2SQ, METRIC_L2, METRIC_IP, METRIC_COSINE,
    METRIC_QI8_L2SQ, METRIC_QI8_IP, METRIC_QI8_COSINE,
    ZvecB200Error, lib, lib_path, Store, Context, Comm, comm_unique_id, distance_matrix, metric_from_names, kmeans, ivf_assign, ivf_pack,
)
