"""CPU oracle for the Matrix-FHE hot path (ctypes view of oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(matrix_fhe_gpu_b200) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_REF_PATH = os.path.join(_HERE, "_ref", "libphantom_host.so")

u64 = C.c_uint64
_p64 = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_p32 = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")


def build(force: bool = False) -> str:
    """Compile oracle/liboracle.so (and oracle/_ref when the reference is present)."""
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h"))]
    stale = force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-s", "-C", _HERE, "all"])
    if os.path.isdir("/root/reference/extern/phantom-fhe/src/host") and (
            force or not os.path.exists(_REF_PATH)):
        subprocess.call(["make", "-s", "-C", _HERE, "ref"])
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        for name in ("mfo_addmod", "mfo_submod", "mfo_mulmod", "mfo_powmod"):
            getattr(L, name).restype = u64
            getattr(L, name).argtypes = [u64, u64, u64]
        L.mfo_invmod.restype = u64; L.mfo_invmod.argtypes = [u64, u64]
        L.mfo_shoup.restype = u64; L.mfo_shoup.argtypes = [u64, u64]
        L.mfo_mulmod_shoup.restype = u64; L.mfo_mulmod_shoup.argtypes = [u64, u64, u64, u64]
        L.mfo_is_prime.restype = C.c_int; L.mfo_is_prime.argtypes = [u64]
        L.mfo_get_primes.restype = C.c_int
        L.mfo_get_primes.argtypes = [u64, C.c_int, C.c_size_t, _p64]
        L.mfo_coeff_modulus_create.restype = C.c_int
        L.mfo_coeff_modulus_create.argtypes = [u64, C.c_int, C.c_size_t, _p64]
        L.mfo_min_primitive_root.restype = u64
        L.mfo_min_primitive_root.argtypes = [u64, u64]
        L.mfo_phantom_tables.restype = C.c_int
        L.mfo_phantom_tables.argtypes = [C.c_int, u64, _p64, _p64, _p64, _p64,
                                         C.POINTER(u64), C.POINTER(u64)]
        L.mfo_negacyclic_forward.restype = None
        L.mfo_negacyclic_forward.argtypes = [_p64, C.c_int, u64, _p64, _p64]
        L.mfo_negacyclic_inverse.restype = None
        L.mfo_negacyclic_inverse.argtypes = [_p64, C.c_int, u64, _p64, _p64, u64, u64]
        L.mfo_negacyclic_forward_def.restype = None
        L.mfo_negacyclic_forward_def.argtypes = [_p64, _p64, C.c_int, u64, u64]
        L.mfo_schoolbook_mul.restype = None
        L.mfo_schoolbook_mul.argtypes = [_p64, _p64, _p64, C.c_size_t, u64, u64]
        L.mfo_negacyclic_forward_batch.restype = None
        L.mfo_negacyclic_forward_batch.argtypes = [_p64, C.c_size_t, C.c_size_t, C.c_int,
                                                   _p64, _p64, _p64]
        L.mfo_negacyclic_inverse_batch.restype = None
        L.mfo_negacyclic_inverse_batch.argtypes = [_p64, C.c_size_t, C.c_size_t, C.c_int,
                                                   _p64, _p64, _p64, _p64, _p64]
        L.mfo_num_threads.restype = C.c_int
        L.mfo_gl_get_psi.restype = u64; L.mfo_gl_get_psi.argtypes = [u64, C.c_int]
        for name in ("mfo_gl_forward", "mfo_gl_backward", "mfo_cyclic_forward", "mfo_cyclic_backward"):
            getattr(L, name).restype = None
            getattr(L, name).argtypes = [_p64, C.c_int, u64]
        L.mfo_gl_perm.restype = None; L.mfo_gl_perm.argtypes = [C.c_int, _p32, _p32]
        _lib = L
    return _lib


_ref = None


def ref():
    """The real Phantom host code (oracle/_ref), or None when it was not built."""
    global _ref
    if _ref is None and os.path.exists(_REF_PATH):
        R = C.CDLL(_REF_PATH)
        R.ref_get_primes.restype = C.c_int
        R.ref_get_primes.argtypes = [u64, C.c_int, C.c_size_t, _p64]
        R.ref_coeff_modulus_create.restype = C.c_int
        R.ref_coeff_modulus_create.argtypes = [u64, C.c_int, C.c_size_t, _p64]
        R.ref_ntt_tables.restype = C.c_int
        R.ref_ntt_tables.argtypes = [C.c_int, u64, C.POINTER(u64), _p64, _p64, _p64, _p64,
                                     C.POINTER(u64), C.POINTER(u64)]
        R.ref_is_prime.restype = C.c_int; R.ref_is_prime.argtypes = [u64]
        _ref = R
    return _ref


# ---------------------------------------------------------------- helpers ----
def get_primes(ntt_size: int, bits: int, count: int) -> np.ndarray:
    out = np.zeros(count, dtype=np.uint64)
    if lib().mfo_get_primes(ntt_size, bits, count, out) != 0:
        raise ValueError("not enough primes")
    return out


def coeff_modulus_create(n: int, bits: int, count: int) -> np.ndarray:
    out = np.zeros(count, dtype=np.uint64)
    if lib().mfo_coeff_modulus_create(n, bits, count, out) != 0:
        raise ValueError("not enough primes")
    return out


class PhantomTables:
    """Per-limb Phantom-convention tables, arrays of shape [L][N]."""

    def __init__(self, logn: int, moduli):
        self.logn = logn
        self.n = 1 << logn
        self.moduli = np.ascontiguousarray(np.asarray(moduli, dtype=np.uint64))
        L = len(self.moduli)
        self.tw = np.zeros((L, self.n), dtype=np.uint64)
        self.tws = np.zeros_like(self.tw)
        self.itw = np.zeros_like(self.tw)
        self.itws = np.zeros_like(self.tw)
        self.n_inv = np.zeros(L, dtype=np.uint64)
        self.n_inv_shoup = np.zeros(L, dtype=np.uint64)
        for l, q in enumerate(self.moduli):
            a, b = u64(), u64()
            rc = lib().mfo_phantom_tables(logn, int(q), self.tw[l], self.tws[l], self.itw[l],
                                          self.itws[l], C.byref(a), C.byref(b))
            if rc != 0:
                raise ValueError(f"modulus {q} has no 2N-th root for N=2^{logn}")
            self.n_inv[l], self.n_inv_shoup[l] = a.value, b.value

    def psi(self, l: int) -> int:
        # tw[bitrev(1)] = tw[N/2] = psi^1
        return int(self.tw[l, self.n // 2])

    def forward(self, data: np.ndarray) -> np.ndarray:
        """data [..., L, N] -> NTT (bit-reversed order), OpenMP over limb-polys."""
        d = np.ascontiguousarray(data, dtype=np.uint64).copy()
        L = len(self.moduli)
        batch = d.size // (L * self.n)
        lib().mfo_negacyclic_forward_batch(d.reshape(-1), batch, L, self.logn, self.moduli,
                                           self.tw.reshape(-1), self.tws.reshape(-1))
        return d

    def inverse(self, data: np.ndarray) -> np.ndarray:
        d = np.ascontiguousarray(data, dtype=np.uint64).copy()
        L = len(self.moduli)
        batch = d.size // (L * self.n)
        lib().mfo_negacyclic_inverse_batch(d.reshape(-1), batch, L, self.logn, self.moduli,
                                           self.itw.reshape(-1), self.itws.reshape(-1),
                                           self.n_inv, self.n_inv_shoup)
        return d


def splitmix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on a uint64 array (the synthetic-input generator of SURVEY 8d)."""
    with np.errstate(over="ignore"):
        z = x.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def synth_polys(seed: int, batch: int, moduli, n: int) -> np.ndarray:
    """x[b][l][j] = splitmix64(seed + flat_idx) mod q_l  (BASELINE config 2 inputs)."""
    L = len(moduli)
    idx = np.arange(batch * L * n, dtype=np.uint64) + np.uint64(seed)
    v = splitmix64(idx).reshape(batch, L, n)
    q = np.asarray(moduli, dtype=np.uint64).reshape(1, L, 1)
    return v % q
