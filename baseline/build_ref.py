"""Test infrastructure, not product: builds the UNMODIFIED reference (Shaibk/Matrix-FHE-GPU, src/core/*.cu) and its
vendored Phantom-FHE for sm_100a from where the sources lie under /root/reference, with separable compilation
(-rdc=true; without it the reference's cross-TU `extern __constant__` tables are silently zero), and
links the two tools of this directory against them:

  baseline/_ref/ref_dump[_stock]   -- calls the reference's own API / Phantom's kernels on fixed inputs and writes hashes and
                              leading words of the outputs (the second parity pin of SURVEY 8(c)); run on the B200 by
                              tests/golden/make_golden_refcuda.sh, its output is committed under tests/golden/
  baseline/_ref/ref_time[_stock]   -- times the reference's pipeline stages and Phantom's NTT / bConv / encoder / symmetric
                              encryption with CUDA events; bench.py reports them as `baselines.{reference_cuda,phantom}`

The flags are the reference's own (CMakeLists.txt:11,14,95-99: -O3 --use_fast_math, PHANTOM_USE_CUDA_PTX) with the
architecture changed from 80 to 100a. Nothing of the reference is copied: objects and binaries go to baseline/_ref/
(git-ignored, travels to the GPU box with the snapshot). Without /root/reference this script is a no-op.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MFHE_REFERENCE_DIR", "/root/reference")
PH = os.path.join(REF, "extern", "phantom-fhe")
OUT = os.path.join(HERE, "_ref")
OBJ = os.path.join(OUT, "obj")
NVCC = "/usr/local/cuda/bin/nvcc"

# the source list of CMakeLists.txt:27-66 (PHANTOM_SOURCES) and :71-77 (src/*.cu minus main.cu)
PHANTOM_SOURCES = [
    "src/context.cu", "src/ckks.cu", "src/rns.cu", "src/polymath.cu", "src/secretkey.cu", "src/galois.cu",
    "src/rns_base.cu", "src/rns_bconv.cu", "src/scalingvariant.cu", "src/evaluate.cu", "src/eval_key_switch.cu",
    "src/fft.cu", "src/prng.cu", "src/batchencoder.cu",
    "src/host/modulus.cu", "src/host/rns.cu", "src/host/ntt.cu", "src/host/uintarith.cu", "src/host/uintarithmod.cu",
    "src/host/uintarithsmallmod.cu", "src/host/globals.cu", "src/host/hash.cu", "src/host/numth.cu",
    "src/host/blake2b.cu", "src/host/blake2xb.cu",
    "src/ntt/ntt_1d.cu", "src/ntt/fntt_2d.cu", "src/ntt/intt_2d.cu", "src/ntt/ntt_modup.cu", "src/ntt/ntt_moddown.cu",
    "src/ntt/ntt_keyswitch_old.cu",
]
REF_SOURCES = ["src/core/HE.cu", "src/core/ntt_core.cu", "src/core/encoder.cu", "src/core/batched_encoder.cu",
               "src/core/trace.cu", "src/core/batched_trace.cu"]
BASE_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "--use_fast_math",
              "-DPHANTOM_USE_CUDA_PTX", "-w",
              "-I", os.path.join(REF, "include"), "-I", os.path.join(PH, "include"), "-I", os.path.join(PH, "src")]
TOOLS = ["ref_dump", "ref_time"]
# flavour -> (extra compile flags, executable suffix).  "rdc": separable compilation, ONE instance of the cross-TU
# __constant__ tables (what a correct build of the reference needs).  "stock": the reference as its
# CMakeLists.txt builds it (no separable compilation), only the architecture changed.
FLAVOURS = {"rdc": (["-rdc=true", "-DREF_RDC=1"], ""), "stock": (["-DREF_RDC=0"], "_stock")}
HEADERS = [os.path.join(HERE, "ref_common.h")]


def _compile(src: str, obj: str, flags) -> str:
    deps = [src] + (HEADERS if src.startswith(HERE) else [])
    if os.path.exists(obj) and all(os.path.getmtime(obj) >= os.path.getmtime(d) for d in deps):
        return obj
    r = subprocess.run([NVCC, *BASE_FLAGS, *flags, "-c", src, "-o", obj], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}\n{r.stdout}")
    return obj


def build(verbose: bool = False, flavours=("rdc", "stock")) -> bool:
    if not os.path.isdir(REF):
        return False
    for fl in flavours:
        flags, suffix = FLAVOURS[fl]
        objdir = OBJ + "_" + fl
        os.makedirs(objdir, exist_ok=True)
        jobs = []
        for s in PHANTOM_SOURCES:
            jobs.append((os.path.join(PH, s), os.path.join(objdir, "ph_" + s.replace("/", "_")[:-3] + ".o")))
        for s in REF_SOURCES:
            jobs.append((os.path.join(REF, s), os.path.join(objdir, "ref_" + os.path.basename(s)[:-3] + ".o")))
        for t in TOOLS:
            jobs.append((os.path.join(HERE, t + ".cu"), os.path.join(objdir, "tool_" + t + ".o")))
        with ThreadPoolExecutor(max_workers=max(1, (os.cpu_count() or 2))) as ex:
            objs = list(ex.map(lambda j: _compile(j[0], j[1], flags), jobs))
        lib_objs = [o for o in objs if not os.path.basename(o).startswith("tool_")]
        for t in TOOLS:
            exe = os.path.join(OUT, t + suffix)
            tool_obj = os.path.join(objdir, "tool_" + t + ".o")
            if os.path.exists(exe) and all(os.path.getmtime(exe) >= os.path.getmtime(o) for o in lib_objs + [tool_obj]):
                continue
            link = ["-rdc=true"] if fl == "rdc" else []
            cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", *link, "-o", exe, tool_obj, *lib_objs, "-lm"]
            r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"link failed for {t}{suffix}\n{r.stdout}")
            if verbose:
                print("built", exe)
    return True


if __name__ == "__main__":
    print("reference built" if build(verbose=True) else "no reference tree: nothing to do")
