"""Builds the reference's OWN drivers (src/main.cu, test/*.cu), unmodified, against this repo's
headers and libraries -- the drop-in check of SURVEY 8(b).

Nothing of the reference is copied: each driver is reached through a symlink placed under
dropin/_build/{src,test}/ so that its `#include "../include/core/..."` resolves to THIS repo's
include/ (dropin/_build/include is a symlink to it).  Outputs (shim library + binaries) land in
git-ignored directories and travel to the GPU box with the snapshot.

  python dropin/build.py            # needs /root/reference; a no-op elsewhere
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("MFHE_REFERENCE_DIR", "/root/reference")
PKG = os.path.join(ROOT, "matrix_fhe_gpu_b200")
OUT = os.path.join(ROOT, "dropin", "_build")
SHIM_SRC = os.path.join(PKG, "shim", "matrix_fhe_shim.cu")
SHIM_LIB = os.path.join(PKG, "libmatrix_fhe_shim.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]

# CMake-built drivers of the reference first, then the stale ones that still only need this API
DRIVERS = ["src/main.cu", "test/test_custom_ntt_roundtrip.cu", "test/test_wcrt_roundtrip.cu",
           "test/test_encode_decode_wcrt.cu", "test/test_encode_encrypt_decrypt_decode_wcrt.cu",
           "test/test_encode_decode_loop.cu", "test/tmp.cu"]
# test/phantom_ntt_roundtrip.cu drives Phantom's own C++ classes (DNTTTable, nwt_2d_radix8_*): not part
# of the reference's API surface, and the reference itself cannot run it (get_ntt_table() exits).


def nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def build_shim(force=False):
    sys.path.insert(0, ROOT)
    from matrix_fhe_gpu_b200 import build as b
    b.build()
    deps = [SHIM_SRC] + [os.path.join(dp, f) for dp, _, fs in os.walk(os.path.join(ROOT, "include")) for f in fs]
    if not force and os.path.exists(SHIM_LIB) and all(os.path.getmtime(d) < os.path.getmtime(SHIM_LIB) for d in deps):
        return SHIM_LIB
    subprocess.check_call([nvcc(), *ARCH, "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-shared", SHIM_SRC,
                           "-o", SHIM_LIB, "-L" + PKG, "-lmfhe_b200", "-Xlinker", "-rpath=$ORIGIN"])
    return SHIM_LIB


def build_drivers():
    if not os.path.isdir(os.path.join(REF, "test")):
        print("reference not present at", REF, "- skipping drivers")
        return []
    os.makedirs(os.path.join(OUT, "bin"), exist_ok=True)
    inc = os.path.join(OUT, "include")
    if not os.path.islink(inc):
        os.symlink(os.path.join(ROOT, "include"), inc)
    built = []
    for rel in DRIVERS:
        d = os.path.join(OUT, os.path.dirname(rel))
        os.makedirs(d, exist_ok=True)
        link = os.path.join(OUT, rel)
        if os.path.lexists(link):
            os.remove(link)
        os.symlink(os.path.join(REF, rel), link)
        exe = os.path.join(OUT, "bin", os.path.splitext(os.path.basename(rel))[0])
        cmd = [nvcc(), *ARCH, "-O3", "-std=c++17", "-w", link, "-o", exe, "-L" + PKG, "-lmatrix_fhe_shim", "-lmfhe_b200",
               "-Xlinker", "-rpath=$ORIGIN/../../../matrix_fhe_gpu_b200"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("drop-in build failed for " + rel + "\n" + r.stdout)
        built.append(exe)
    # our own driver for the SURVEY 8(f) rank-1 row (same style, same link line; it includes ../include/...)
    own = os.path.join(ROOT, "dropin", "drivers", "test_encrypted_matmul.cu")
    d = os.path.join(OUT, "test")
    os.makedirs(d, exist_ok=True)
    link = os.path.join(d, "test_encrypted_matmul.cu")
    if os.path.lexists(link):
        os.remove(link)
    os.symlink(own, link)
    exe = os.path.join(OUT, "bin", "test_encrypted_matmul")
    cmd = [nvcc(), *ARCH, "-O3", "-std=c++17", "-w", link, "-o", exe, "-L" + PKG, "-lmatrix_fhe_shim", "-lmfhe_b200",
           "-Xlinker", "-rpath=$ORIGIN/../../../matrix_fhe_gpu_b200"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("drop-in build failed for test_encrypted_matmul.cu\n" + r.stdout)
    built.append(exe)
    return built


if __name__ == "__main__":
    print(build_shim(force="--force" in sys.argv))
    for e in build_drivers():
        print(e)
