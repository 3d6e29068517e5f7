"""CPU tests of the C-ABI boundary: the library loads, exports every symbol include/zvec_b200.h
declares, and fails loudly (no CPU fallback) without a device."""
import ctypes as C
import os
import re

import pytest

import zvec_b200 as z
from zvec_b200 import binding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "zvec_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(zvb200_\w+)\s*\(", txt)))


def test_library_loads_and_exports_every_declared_symbol():
    L = C.CDLL(z.lib_path())
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(L, s), "missing export %s" % s
    # the binding covers the whole header too
    assert set(syms) == set(binding.exported_symbols())
    assert b"sm_100a" in z.lib().zvb200_version()


def test_argument_errors_use_reference_codes():
    L = z.lib()
    h = C.c_void_p()
    assert L.zvb200_store_create(0, 3, z.METRIC_L2SQ, 8, C.byref(h)) == -12       # DT_FP64 unsupported
    assert L.zvb200_store_create(0, z.DT_FP32, 99, 8, C.byref(h)) == -12           # unknown metric
    assert L.zvb200_store_create(0, z.DT_FP32, z.METRIC_QI8_L2SQ, 32, C.byref(h)) == -12
    assert L.zvb200_store_create(0, z.DT_INT8, z.METRIC_COSINE, 32, C.byref(h)) == -12
    assert L.zvb200_store_create(0, z.DT_INT8, z.METRIC_QI8_L2SQ, 16, C.byref(h)) == -31  # no room for codes
    assert L.zvb200_store_create(0, z.DT_FP32, z.METRIC_L2SQ, 8, None) == -31
    assert L.zvb200_last_error()


def test_no_cpu_fallback_without_device():
    if z.lib().zvb200_device_count() > 0:
        pytest.skip("a device is present")
    with pytest.raises(z.ZvecB200Error) as e:
        z.Store(z.DT_FP32, z.METRIC_L2SQ, 128)
    assert e.value.code == -1 and "no CPU fallback" in str(e.value)


def test_product_never_touches_the_oracle():
    bad = []
    for d in ("zvec_b200", "include"):
        for root, _, files in os.walk(os.path.join(ROOT, d)):
            if "build" in root or "/lib" in root:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp")):
                    t = open(os.path.join(root, f), errors="ignore").read()
                    if re.search(r"oracle/|zvec_oracle|zvec_ref|zo_\w+\(|zref_", t):
                        bad.append(os.path.join(root, f))
    assert not bad, bad
