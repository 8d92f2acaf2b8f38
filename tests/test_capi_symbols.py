"""The C-ABI library builds for sm_100a, loads without a GPU and exports every symbol
declared in include/mfhe.h.  No compute call is made here."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    out = []
    for fn in sorted(os.listdir(os.path.join(ROOT, "include"))):
        if fn.endswith(".h"):
            src = open(os.path.join(ROOT, "include", fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            out += re.findall(r"\b(mfhe_[a-z0-9_]+)\s*\(", src)
    return sorted(set(out))


def test_library_exports_every_declared_symbol():
    import matrix_fhe_gpu_b200 as m
    lib = ctypes.CDLL(m.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 15
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_version_and_error_plumbing_without_gpu():
    import matrix_fhe_gpu_b200 as m
    L = m.lib()
    assert b"sm_100a" in L.mfhe_version()
    h = ctypes.c_void_p()
    # invalid size is rejected on the host before any CUDA call
    rc = L.mfhe_ntt_plan_create(3, 1, (ctypes.c_uint64 * 1)(97), 0, ctypes.byref(h))
    assert rc == -4 and b"logn" in L.mfhe_last_error()
    rc = L.mfhe_ntt_plan_create(6, 1, (ctypes.c_uint64 * 1)(1 << 61), 0, ctypes.byref(h))
    assert rc == -1 and b"2^60" in L.mfhe_last_error()


def test_no_cpu_fallback_in_product_package():
    # the product package must not import or reference the oracle
    pkg = os.path.join(ROOT, "matrix_fhe_gpu_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "liboracle" not in txt, f
