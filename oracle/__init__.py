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
        L.mfo_set_num_threads.restype = None; L.mfo_set_num_threads.argtypes = [C.c_int]
        L.mfo_philox4x32_10.restype = None; L.mfo_philox4x32_10.argtypes = [_p32, C.c_uint32, C.c_uint32]
        L.mfo_sound_cdt.restype = None; L.mfo_sound_cdt.argtypes = [_p64]
        L.mfo_sound_sample.restype = None
        L.mfo_sound_sample.argtypes = [C.c_int, _p64, _p64, C.c_int, C.c_int, C.c_int, u64, u64, u64]
        L.mfo_gl_get_psi.restype = u64; L.mfo_gl_get_psi.argtypes = [u64, C.c_int]
        for name in ("mfo_gl_forward", "mfo_gl_backward", "mfo_cyclic_forward", "mfo_cyclic_backward"):
            getattr(L, name).restype = None
            getattr(L, name).argtypes = [_p64, C.c_int, u64]
        L.mfo_gl_perm.restype = None; L.mfo_gl_perm.argtypes = [C.c_int, _p32, _p32]
        _pd = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
        L.mfo_fp64_modmul_check.restype = C.c_long
        L.mfo_fp64_modmul_check.argtypes = [u64, C.c_long, u64, C.c_int, C.POINTER(C.c_double)]
        L.mfo_fp64_negacyclic_forward.restype = C.c_int
        L.mfo_fp64_negacyclic_forward.argtypes = [_p64, C.c_int, u64, _p64, _pd]
        L.mfo_fp64_negacyclic_inverse.restype = C.c_int
        L.mfo_fp64_negacyclic_inverse.argtypes = [_p64, C.c_int, u64, _p64, u64, _pd]
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
def use_all_cores() -> int:
    """Sets the oracle's OpenMP thread count to the cores this process may run on (torch.distributed.run exports
    OMP_NUM_THREADS=1, which would otherwise make the CPU arm single-threaded); returns the count."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    lib().mfo_set_num_threads(n)
    return lib().mfo_num_threads()


def philox4x32_10(ctr, key):
    c = np.asarray(ctr, dtype=np.uint32).copy()
    lib().mfo_philox4x32_10(c, int(key[0]), int(key[1]))
    return [int(v) for v in c]


def sound_cdt():
    t = np.zeros(83, dtype=np.uint64); lib().mfo_sound_cdt(t); return t


def sound_sample(kind, moduli, n, phi, seed, draw):
    """kind 0 uniform / 1 gaussian: [phi][L][n][n]; 2 ternary: [phi][L][n]  (sampler_oracle.c)."""
    mods = np.ascontiguousarray(np.asarray(moduli, dtype=np.uint64))
    out = np.zeros(phi * len(mods) * (n if kind == 2 else n * n), dtype=np.uint64)
    lib().mfo_sound_sample(kind, out, mods, n, len(mods), phi, int(seed[0]), int(seed[1]), int(draw))
    return out


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


def fp64_modmul_check(q: int, count: int, seed: int, bound16: int):
    """(failures, max |t| / q) of the FP64-pipe modular product over `count` operand pairs with |y| <= bound16/16 * q."""
    mx = C.c_double()
    bad = lib().mfo_fp64_modmul_check(int(q), count, seed, bound16, C.byref(mx))
    return int(bad), float(mx.value)


def fp64_ntt(poly: np.ndarray, logn: int, q: int, table: np.ndarray, inverse: bool = False, n_inv: int = 0):
    """One limb-poly through the FP64-pipe arithmetic of the CUDA kernels, emulated on the CPU with the kernels' fold
    schedule.  Returns (result, rc, largest |v| / q entering every stage); rc < 0 names a stage that broke its bound."""
    a = np.ascontiguousarray(poly, dtype=np.uint64).copy()
    st = np.zeros(logn, dtype=np.float64)
    tab = np.ascontiguousarray(table, dtype=np.uint64)
    if inverse:
        rc = lib().mfo_fp64_negacyclic_inverse(a, logn, int(q), tab, int(n_inv), st)
    else:
        rc = lib().mfo_fp64_negacyclic_forward(a, logn, int(q), tab, st)
    return a, int(rc), st


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


# ------------------------------------------------------------------ matrix pipeline (family B) ----
REF_MODULI = [17592186435073, 17182765057, 17184541441, 17186120449, 17186515201, 17186909953,
              17188883713, 17190462721, 17190857473, 17191844353, 17192831233]      # config.h:31-43
REF_DELTA = 34359738368.0                                                             # config.h:25

_pcx = np.ctypeslib.ndpointer(dtype=np.complex128, flags="C_CONTIGUOUS")
_pi64 = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_pf64 = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_pu8 = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_scheme_ready = False


def _scheme_lib():
    global _scheme_ready
    L = lib()
    if _scheme_ready:
        return L
    vp, ci = C.c_void_p, C.c_int
    L.mfo_ctx_create.restype = vp
    L.mfo_ctx_create.argtypes = [ci, ci, ci, _p64, C.c_double, ci]
    L.mfo_ctx_destroy.argtypes = [vp]
    L.mfo_ctx_wntt.restype = C.POINTER(u64); L.mfo_ctx_wntt.argtypes = [vp, ci]
    L.mfo_ctx_wexp.restype = C.POINTER(C.c_uint16); L.mfo_ctx_wexp.argtypes = [vp]
    L.mfo_ctx_eta.restype = u64; L.mfo_ctx_eta.argtypes = [vp, ci]
    L.mfo_ctx_big_words.restype = ci; L.mfo_ctx_big_words.argtypes = [vp]
    L.mfo_ctx_crt.restype = C.POINTER(u64); L.mfo_ctx_crt.argtypes = [vp, ci]
    L.mfo_ctx_cx.restype = C.POINTER(C.c_double); L.mfo_ctx_cx.argtypes = [vp, ci]
    L.mfo_find_eta.restype = u64; L.mfo_find_eta.argtypes = [u64, ci]
    L.mfo_matrix_inverse_mod.restype = ci; L.mfo_matrix_inverse_mod.argtypes = [_p64, _p64, ci, u64]
    sig = {
        "mfo_wntt_forward_matrix": [vp, _p64, _p64], "mfo_wntt_inverse_matrix": [vp, _p64, _p64],
        "mfo_wntt_forward_vector": [vp, _p64, _p64],
        "mfo_wntt_forward_centered": [vp, _pi64, _pi64], "mfo_wntt_inverse_centered": [vp, _pi64, _pi64],
        "mfo_matrix_to_poly": [vp, _p64, _p64], "mfo_poly_to_matrix": [vp, _p64, _p64],
        "mfo_wdft_forward_complex": [vp, _pcx, _pcx], "mfo_w_idft": [vp, _pcx, _pcx],
        "mfo_wdft_forward_pair": [vp, _pi64, _pi64, _pf64, _pf64],
        "mfo_wdft_inverse_pair": [vp, _pf64, _pf64, _pf64, _pf64],
        "mfo_xy_idft2": [vp, _pcx, _pcx, ci], "mfo_xy_dft2": [vp, _pcx, _pcx, ci],
        "mfo_quantize": [vp, _pcx, _p64, _p64, ci],
        "mfo_crt_compose_big": [vp, _p64, _p64, _pu8, ci, ci],
        "mfo_crt_compose_i64": [vp, _p64, _pi64, ci, ci],
        "mfo_dequantize_exact": [vp, _p64, _p64, _pcx, ci, ci],
        "mfo_big_pair_to_complex": [vp, _p64, _pu8, _p64, _pu8, _pcx, C.c_size_t],
        "mfo_encode": [vp, _pcx, _p64, _p64], "mfo_encode_lane": [vp, _pcx, _p64, _p64],
        "mfo_decode_eval_pair": [vp, _p64, _p64, _pcx],
        "mfo_decode_lane_from_rns_eval": [vp, _p64, _p64, _pcx],
        "mfo_sample_uniform": [vp, _p64], "mfo_sample_gaussian": [vp, _p64], "mfo_sample_ternary": [vp, _p64],
        "mfo_keygen": [vp, _p64],
        "mfo_xy_ntt_forward": [vp, _p64, C.c_size_t], "mfo_xy_ntt_backward": [vp, _p64, C.c_size_t],
        "mfo_encrypt_pair_with": [vp, _p64, _p64, _p64, _p64, _p64, _p64, _p64, _p64],
        "mfo_encrypt_pair": [vp, _p64, _p64, _p64, _p64, _p64],
        "mfo_encrypt": [vp, _p64, _p64, _p64],
        "mfo_decrypt_to_eval": [vp, _p64, _p64, _p64],
        "mfo_decrypt_and_decode": [vp, _p64, _p64, _p64, _pcx],
        "mfo_ct_add": [vp, _p64, _p64, _p64], "mfo_ct_mul_raw": [vp, _p64, _p64, _p64, _p64, _p64],
        "mfo_trace_map": [vp, _p64, _p64, _p64, _p64, ci],
        "mfo_trace_gemm": [vp, _p64, _p64, _p64, _p64, _p64, _p64, ci],
        "mfo_trace_rescale": [vp, _p64, _p64, ci, u64, u64, u64],
        "mfo_rescale_last": [vp, _p64, _p64],
        "mfo_trace_map_x": [vp, _p64, _p64, _p64, _p64, ci],
        "mfo_trace_gemm_ty": [vp, _p64, _p64, _p64, _p64, _p64, _p64, ci],
        "mfo_ct_matmul": [vp, _p64, _p64, _p64, _p64, _p64],
        "mfo_ct_matmul_decrypt": [vp, _p64, _p64, _p64, _p64],
        "mfo_bconv": [_p64, _p64, C.c_size_t, ci, _p64, ci, _p64],
    }
    for name, args in sig.items():
        getattr(L, name).restype = None
        getattr(L, name).argtypes = args
    _scheme_ready = True
    return L


def primes_congruent_one(modulus: int, bits: int, count: int, start_above: int | None = None):
    """First `count` primes q = 1 (mod modulus) above 2^(bits-1) (the reference's prime rule)."""
    v = (start_above if start_above is not None else (1 << (bits - 1)))
    v += (1 - v) % modulus
    out = []
    L = lib()
    while len(out) < count:
        if L.mfo_is_prime(v):
            out.append(v)
        v += modulus
    return out


class SchemeCtx:
    """Oracle context for the ring Z_Q[X,Y,W]/(X^n - i, Y^n - i, Phi_{3f}(W)).

    Native shape of the reference: SchemeCtx() == n=64, f=257 (p=771, phi=512), L=11, delta=2^35."""

    def __init__(self, n=64, f=257, moduli=None, delta=REF_DELTA, gauss_jordan=False):
        self._L = _scheme_lib()
        self.moduli = np.asarray(REF_MODULI if moduli is None else moduli, dtype=np.uint64)
        self.n, self.f, self.p, self.phi, self.L, self.delta = n, f, 3 * f, 2 * (f - 1), len(self.moduli), float(delta)
        self._h = self._L.mfo_ctx_create(n, f, self.L, self.moduli, self.delta, int(gauss_jordan))
        if not self._h:
            raise ValueError("mfo_ctx_create failed (moduli must be = 1 mod lcm(2n, 3f))")
        self.n2 = n * n
        self.T = self.phi * self.L * self.n2
        self.big_words = self._L.mfo_ctx_big_words(self._h)

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.mfo_ctx_destroy(self._h)
            self._h = None

    # tables ------------------------------------------------------------------------------------
    def wexp(self):
        return np.ctypeslib.as_array(self._L.mfo_ctx_wexp(self._h), shape=(self.phi,)).copy()

    def eta(self, l):
        return int(self._L.mfo_ctx_eta(self._h, l))

    def wntt(self, inverse=False):
        return np.ctypeslib.as_array(self._L.mfo_ctx_wntt(self._h, int(inverse)),
                                     shape=(self.L, self.phi, self.phi)).copy()

    def crt(self):
        W = self.big_words
        g = lambda k, shape: np.ctypeslib.as_array(self._L.mfo_ctx_crt(self._h, k), shape=shape).copy()
        return {"Q": g(0, (W,)), "Qhalf": g(1, (W,)), "M": g(2, (self.L, W)), "inv": g(3, (self.L,))}

    def cx_table(self, which):
        k = {"wdft": 0, "wdft_inv": 1, "V": 2, "VT": 3, "Vinv": 4, "VinvT": 5}[which]
        d = self.phi if k < 2 else self.n
        a = np.ctypeslib.as_array(self._L.mfo_ctx_cx(self._h, k), shape=(d, d, 2)).copy()
        return a.view(np.complex128).reshape(d, d)

    # generic call helper: allocates outputs -------------------------------------------------------
    def _u(self, *shape):
        return np.zeros(shape, dtype=np.uint64)

    def _c(self, a, dt):
        return np.ascontiguousarray(a, dtype=dt)

    def wntt_forward_matrix(self, x):
        o = self._u(self.T); self._L.mfo_wntt_forward_matrix(self._h, self._c(x, np.uint64).reshape(-1), o); return o

    def wntt_inverse_matrix(self, x):
        o = self._u(self.T); self._L.mfo_wntt_inverse_matrix(self._h, self._c(x, np.uint64).reshape(-1), o); return o

    def wntt_forward_vector(self, x):
        o = self._u(self.phi * self.L * self.n)
        self._L.mfo_wntt_forward_vector(self._h, self._c(x, np.uint64).reshape(-1), o); return o

    def wntt_forward_centered(self, x):
        o = np.zeros(self.phi * self.n2, dtype=np.int64)
        self._L.mfo_wntt_forward_centered(self._h, self._c(x, np.int64).reshape(-1), o); return o

    def wntt_inverse_centered(self, x):
        o = np.zeros(self.phi * self.n2, dtype=np.int64)
        self._L.mfo_wntt_inverse_centered(self._h, self._c(x, np.int64).reshape(-1), o); return o

    def matrix_to_poly(self, x):
        o = self._u(self.T); self._L.mfo_matrix_to_poly(self._h, self._c(x, np.uint64).reshape(-1), o); return o

    def poly_to_matrix(self, x):
        o = self._u(self.T); self._L.mfo_poly_to_matrix(self._h, self._c(x, np.uint64).reshape(-1), o); return o

    def _cx(self, fn, x, count=None):
        x = self._c(x, np.complex128).reshape(-1)
        o = np.zeros(x.size if count is None else count, dtype=np.complex128)
        fn(self._h, x, o); return o

    def wdft_forward_complex(self, x):
        return self._cx(self._L.mfo_wdft_forward_complex, x)

    def w_idft(self, x):
        return self._cx(self._L.mfo_w_idft, x)

    def wdft_forward_pair(self, re, im):
        ore = np.zeros(self.phi * self.n2); oim = np.zeros(self.phi * self.n2)
        self._L.mfo_wdft_forward_pair(self._h, self._c(re, np.int64).reshape(-1), self._c(im, np.int64).reshape(-1), ore, oim)
        return ore, oim

    def wdft_inverse_pair(self, re, im):
        ore = np.zeros(self.phi * self.n2); oim = np.zeros(self.phi * self.n2)
        self._L.mfo_wdft_inverse_pair(self._h, self._c(re, np.float64).reshape(-1), self._c(im, np.float64).reshape(-1), ore, oim)
        return ore, oim

    def xy_idft2(self, x, lanes):
        x = self._c(x, np.complex128).reshape(-1); o = np.zeros_like(x)
        self._L.mfo_xy_idft2(self._h, x, o, lanes); return o

    def xy_dft2(self, x, lanes):
        x = self._c(x, np.complex128).reshape(-1); o = np.zeros_like(x)
        self._L.mfo_xy_dft2(self._h, x, o, lanes); return o

    def quantize(self, x, lanes):
        re = self._u(lanes * self.L * self.n2); im = self._u(lanes * self.L * self.n2)
        self._L.mfo_quantize(self._h, self._c(x, np.complex128).reshape(-1), re, im, lanes); return re, im

    def crt_compose_big(self, x, n2=None, limbs=None):
        n2 = self.n2 if n2 is None else n2; limbs = self.L if limbs is None else limbs
        mag = self._u(n2 * self.big_words); neg = np.zeros(n2, dtype=np.uint8)
        self._L.mfo_crt_compose_big(self._h, self._c(x, np.uint64).reshape(-1), mag, neg, n2, limbs); return mag, neg

    def crt_compose_i64(self, x, n2=None, limbs=None):
        n2 = self.n2 if n2 is None else n2; limbs = self.L if limbs is None else limbs
        o = np.zeros(n2, dtype=np.int64)
        self._L.mfo_crt_compose_i64(self._h, self._c(x, np.uint64).reshape(-1), o, n2, limbs); return o

    def dequantize_exact(self, re, im, n2=None, limbs=None):
        n2 = self.n2 if n2 is None else n2; limbs = self.L if limbs is None else limbs
        o = np.zeros(n2, dtype=np.complex128)
        self._L.mfo_dequantize_exact(self._h, self._c(re, np.uint64).reshape(-1), self._c(im, np.uint64).reshape(-1), o, n2, limbs)
        return o

    def big_pair_to_complex(self, rm, rn, im, in_):
        total = rn.size; o = np.zeros(total, dtype=np.complex128)
        self._L.mfo_big_pair_to_complex(self._h, rm, rn, im, in_, o, total); return o

    def encode(self, msg):
        re = self._u(self.T); im = self._u(self.T)
        self._L.mfo_encode(self._h, self._c(msg, np.complex128).reshape(-1), re, im); return re, im

    def encode_lane(self, msg):
        re = self._u(self.L * self.n2); im = self._u(self.L * self.n2)
        self._L.mfo_encode_lane(self._h, self._c(msg, np.complex128).reshape(-1), re, im); return re, im

    def decode_eval_pair(self, er, ei):
        o = np.zeros(self.phi * self.n2, dtype=np.complex128)
        self._L.mfo_decode_eval_pair(self._h, self._c(er, np.uint64).reshape(-1), self._c(ei, np.uint64).reshape(-1), o); return o

    def decode_lane_from_rns_eval(self, re, im):
        o = np.zeros(self.n2, dtype=np.complex128)
        self._L.mfo_decode_lane_from_rns_eval(self._h, self._c(re, np.uint64).reshape(-1), self._c(im, np.uint64).reshape(-1), o)
        return o

    def sample_uniform(self):
        o = self._u(self.T); self._L.mfo_sample_uniform(self._h, o); return o

    def sample_gaussian(self):
        o = self._u(self.T); self._L.mfo_sample_gaussian(self._h, o); return o

    def sample_ternary(self):
        o = self._u(self.phi * self.L * self.n); self._L.mfo_sample_ternary(self._h, o); return o

    def keygen(self):
        o = self._u(self.phi * self.L * self.n); self._L.mfo_keygen(self._h, o); return o

    def xy_ntt_forward(self, x, cnt):
        d = self._c(x, np.uint64).reshape(-1).copy(); self._L.mfo_xy_ntt_forward(self._h, d, cnt); return d

    def xy_ntt_backward(self, x, cnt):
        d = self._c(x, np.uint64).reshape(-1).copy(); self._L.mfo_xy_ntt_backward(self._h, d, cnt); return d

    def encrypt_pair_with(self, m_re, m_im, sk, a, e_re, e_im):
        cr = self._u(2 * self.T); ci = self._u(2 * self.T)
        self._L.mfo_encrypt_pair_with(self._h, m_re, m_im, sk, a, e_re, e_im, cr, ci); return cr, ci

    def encrypt_pair(self, m_re, m_im, sk):
        cr = self._u(2 * self.T); ci = self._u(2 * self.T)
        self._L.mfo_encrypt_pair(self._h, m_re, m_im, sk, cr, ci); return cr, ci

    def encrypt(self, m, sk):
        ct = self._u(2 * self.T); self._L.mfo_encrypt(self._h, m, sk, ct); return ct

    def decrypt_to_eval(self, ct, sk):
        o = self._u(self.T); self._L.mfo_decrypt_to_eval(self._h, ct, sk, o); return o

    def decrypt_and_decode(self, cr, ci, sk):
        o = np.zeros(self.phi * self.n2, dtype=np.complex128)
        self._L.mfo_decrypt_and_decode(self._h, cr, ci, sk, o); return o

    def ct_add(self, a, b):
        o = self._u(2 * self.T); self._L.mfo_ct_add(self._h, a, b, o); return o

    def ct_mul_raw(self, a, b):
        d = [self._u(self.T) for _ in range(3)]; self._L.mfo_ct_mul_raw(self._h, a, b, *d); return d

    def trace_map(self, Br, Bi, batch):
        o = [np.zeros_like(Br), np.zeros_like(Bi)]; self._L.mfo_trace_map(self._h, Br, Bi, o[0], o[1], batch); return o

    def trace_gemm(self, Ar, Ai, Bpr, Bpi, batch):
        o = [np.zeros_like(Ar), np.zeros_like(Ai)]
        self._L.mfo_trace_gemm(self._h, Ar, Ai, Bpr, Bpi, o[0], o[1], batch); return o

    # -- encrypted matrix product --------------------------------------------------
    def trace_map_x(self, Br, Bi, batch):
        o = [np.zeros_like(Br), np.zeros_like(Bi)]; self._L.mfo_trace_map_x(self._h, Br, Bi, o[0], o[1], batch); return o

    def trace_gemm_ty(self, Ur, Ui, Vr, Vi, batch):
        o = [np.zeros_like(Ur), np.zeros_like(Ui)]
        self._L.mfo_trace_gemm_ty(self._h, Ur, Ui, Vr, Vi, o[0], o[1], batch); return o

    def ct_matmul(self, cu_re, cu_im, cv_re, cv_im):
        o = self._u(8 * self.T); self._L.mfo_ct_matmul(self._h, cu_re, cu_im, cv_re, cv_im, o); return o

    def ct_matmul_decrypt(self, prod, sk):
        o = [self._u(self.T), self._u(self.T)]; self._L.mfo_ct_matmul_decrypt(self._h, prod, sk, o[0], o[1]); return o

    def rescale_last(self, x):
        o = self._u(self.phi * (self.L - 1) * self.n2); self._L.mfo_rescale_last(self._h, self._c(x, np.uint64).reshape(-1), o); return o

    def conj_lane(self):
        """Involution l -> l' of the W lanes with exp[l'] = -exp[l] mod p: coefficient-wise conjugation of an encoded
        plaintext conjugates the slot values of lane l' into lane l (the W evaluation points are complex)."""
        e = [int(v) for v in self.wexp()]
        pos = {v: i for i, v in enumerate(e)}
        return np.asarray([pos[(self.p - v) % self.p] for v in e], dtype=np.int64)

    def trace_rescale(self, Cr, Ci, batch, inv0, inv1, inv2):
        Cr, Ci = Cr.copy(), Ci.copy(); self._L.mfo_trace_rescale(self._h, Cr, Ci, batch, inv0, inv1, inv2); return Cr, Ci


def bconv(x, q_in, p_out):
    """Phantom bConv_BEHZ on x [L_in][N] -> [L_out][N]."""
    L = _scheme_lib()
    q_in = np.asarray(q_in, dtype=np.uint64); p_out = np.asarray(p_out, dtype=np.uint64)
    x = np.ascontiguousarray(x, dtype=np.uint64)
    N = x.size // len(q_in)
    out = np.zeros(len(p_out) * N, dtype=np.uint64)
    L.mfo_bconv(x.reshape(-1), out, N, len(q_in), q_in, len(p_out), p_out)
    return out.reshape(len(p_out), N)
