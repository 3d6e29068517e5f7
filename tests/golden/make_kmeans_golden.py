#!/usr/bin/env python3
"""Golden vectors for k-means, produced by the REFERENCE's own KmeansCluster (compiled from its sources into
oracle/_ref, driven through oracle/ref_shim.cc::zref_kmeans = IndexFactory::CreateCluster("KmeansCluster") ->
init / mount / cluster / label, src/core/algorithm/cluster/kmeans_cluster.cc:469-535,571-610).
Inputs are small integers / 8 so every sum is exact in double and the file is tiny.
Run in the build container only:  python tests/golden/make_kmeans_golden.py > tests/golden/kmeans.json"""
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import support as S  # noqa: E402

L = S.ref()
assert L is not None, "build oracle/_ref first (python oracle/build_ref.py)"
L.zref_kmeans.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32,
                          C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
cases = []
for seed, (n, d, k, iters, far) in enumerate([(240, 6, 5, 1, False), (240, 6, 5, 4, False), (150, 3, 4, 2, True)]):
    rng = np.random.default_rng(100 + seed)
    centres = rng.integers(-40, 41, (k, d))
    x = (centres[rng.integers(0, k, n)] + rng.integers(-9, 10, (n, d))).astype(np.float32) / 8.0
    init = x[rng.choice(n, k, replace=False)].copy()
    if far:
        init[k - 1] = 1.0e4                       # nobody's nearest centroid: it must come back as NaN
    cent = init.copy()
    lab = np.zeros(n, np.uint32); fol = np.zeros(k, np.uint64); sse = C.c_double()
    rc = L.zref_kmeans(b"SquaredEuclidean", x.ctypes.data, n, d, cent.ctypes.data, k, iters, 2, lab.ctypes.data,
                       fol.ctypes.data, C.byref(sse))
    assert rc == 0, L.zref_last_error()
    cases.append({"n": n, "dim": d, "k": k, "iterations": iters,
                  "rows_times_8": (x * 8).astype(int).ravel().tolist(),
                  "init_times_8": np.where(np.abs(init) > 1e3, 80000, (init * 8)).astype(int).ravel().tolist(),
                  "centroids_bits": cent.view(np.uint32).ravel().tolist(),
                  "labels": lab.tolist(), "follows": fol.astype(int).tolist(), "sse": sse.value})
json.dump({"source": "KmeansCluster via oracle/_ref (zref_kmeans), metric SquaredEuclidean", "cases": cases}, sys.stdout)
