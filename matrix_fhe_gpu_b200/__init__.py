"""matrix_fhe_gpu_b200 -- B200-native (sm_100a) hot path of Shaibk/Matrix-FHE-GPU.

The product is the CUDA shared library ``libmfhe_b200.so`` (C ABI in ``include/mfhe.h``).
This package is the thin Python host side above that ABI: it loads the library with
ctypes, passes raw device pointers / stream handles of torch tensors, and mirrors the
reference's operator names.  There is no CPU fallback: importing works without a GPU
(for build and symbol checks), but every compute call requires the CUDA library and a
CUDA device and raises otherwise.
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

__all__ = ["lib", "NttPlan", "MfheError", "PHANTOM_NEGACYCLIC", "GL_XN_MINUS_I", "CYCLIC_NATURAL",
           "LIB_PATH"]

LIB_PATH = _build.LIB
PHANTOM_NEGACYCLIC, GL_XN_MINUS_I, CYCLIC_NATURAL = 0, 1, 2

_u64 = C.c_uint64
_vp = C.c_void_p


class MfheError(RuntimeError):
    pass


_lib = None


def lib() -> C.CDLL:
    """The C-ABI library.  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MfheError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                            "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.mfhe_last_error.restype = C.c_char_p
        L.mfhe_version.restype = C.c_char_p
        L.mfhe_ntt_plan_create.restype = C.c_int
        L.mfhe_ntt_plan_create.argtypes = [C.c_int, C.c_int, C.POINTER(_u64), C.c_int, C.POINTER(_vp)]
        L.mfhe_ntt_plan_destroy.restype = None
        L.mfhe_ntt_plan_destroy.argtypes = [_vp]
        L.mfhe_ntt_plan_logn.argtypes = [_vp]
        L.mfhe_ntt_plan_limbs.argtypes = [_vp]
        L.mfhe_ntt_plan_root.restype = _u64
        L.mfhe_ntt_plan_root.argtypes = [_vp, C.c_int]
        L.mfhe_ntt_plan_get_table.argtypes = [_vp, C.c_int, C.c_int, _vp, _vp]
        L.mfhe_ntt_plan_set_chunk.argtypes = [_vp, C.c_size_t]
        L.mfhe_ntt_plan_arith.argtypes = [_vp]
        L.mfhe_ntt_plan_set_arith.argtypes = [_vp, C.c_int]
        for name in ("mfhe_ntt_forward", "mfhe_ntt_inverse"):
            getattr(L, name).restype = C.c_int
            getattr(L, name).argtypes = [_vp, _vp, C.c_size_t, _vp]
        for name in ("mfhe_ntt_forward_limbs", "mfhe_ntt_inverse_limbs"):
            getattr(L, name).restype = C.c_int
            getattr(L, name).argtypes = [_vp, _vp, C.c_size_t, C.c_int, C.c_int, _vp]
        for name in ("mfhe_ntt_forward_host", "mfhe_ntt_inverse_host"):
            getattr(L, name).restype = C.c_int
            getattr(L, name).argtypes = [_vp, _vp, C.c_size_t]
        for name in ("mfhe_poly_mul", "mfhe_poly_add", "mfhe_poly_sub"):
            getattr(L, name).restype = C.c_int
            getattr(L, name).argtypes = [_vp, _vp, _vp, _vp, C.c_size_t, _vp]
        L.mfhe_ntt_mul_intt.restype = C.c_int
        L.mfhe_ntt_mul_intt.argtypes = [_vp, _vp, _vp, C.c_int, C.c_size_t, _vp]
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise MfheError(f"mfhe error {rc}: {lib().mfhe_last_error().decode()}")


def _stream_handle(stream=None):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return _vp(s.cuda_stream)


def _dev_ptr(t):
    import torch
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.is_contiguous()):
        raise MfheError("expected a contiguous CUDA tensor (no CPU path exists)")
    if t.element_size() != 8:
        raise MfheError("expected 64-bit elements")
    return _vp(t.data_ptr())


class NttPlan:
    """Per-(N, moduli, convention) twiddle tables on the current device.

    Mirrors what the reference gets from ``PhantomContext::gpu_rns_tables()``
    (src/core/HE.cu:327-336, 429-435) and ``init_ntt_tables_manual`` /
    ``init_gl_twist_tables`` (src/core/ntt_core.cu:75-198).
    """

    def __init__(self, logn: int, moduli, convention: int = PHANTOM_NEGACYCLIC):
        mods = [int(m) for m in moduli]
        arr = (_u64 * len(mods))(*mods)
        h = _vp()
        check(lib().mfhe_ntt_plan_create(logn, len(mods), arr, convention, C.byref(h)))
        self._h = h
        self.logn, self.n, self.moduli, self.convention = logn, 1 << logn, mods, convention

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and _lib is not None:
            _lib.mfhe_ntt_plan_destroy(h)

    @property
    def limbs(self) -> int:
        return len(self.moduli)

    def root(self, limb: int) -> int:
        return int(lib().mfhe_ntt_plan_root(self._h, limb))

    def table(self, limb: int, inverse: bool = False):
        import numpy as np
        w = np.zeros(self.n, dtype=np.uint64)
        ws = np.zeros(self.n, dtype=np.uint64)
        check(lib().mfhe_ntt_plan_get_table(self._h, limb, int(inverse), w.ctypes.data_as(_vp),
                                            ws.ctypes.data_as(_vp)))
        return w, ws

    def set_chunk(self, limb_polys: int) -> None:
        check(lib().mfhe_ntt_plan_set_chunk(self._h, limb_polys))

    @property
    def arith(self) -> int:
        """0: integer butterflies; 1: FP64-pipe butterflies (flat plans with every modulus below 2^50, then the default)."""
        return lib().mfhe_ntt_plan_arith(self._h)

    def set_arith(self, arith: int) -> None:
        check(lib().mfhe_ntt_plan_set_arith(self._h, arith))

    def _batch(self, t, limbs=None) -> int:
        per = (limbs or self.limbs) * self.n
        if t.numel() % per:
            raise MfheError(f"tensor of {t.numel()} elements is not [batch][{limbs or self.limbs}][{self.n}]")
        return t.numel() // per

    # device tensors (torch int64/uint64 viewed as u64), in place ------------------------
    def forward_(self, t, stream=None):
        check(lib().mfhe_ntt_forward(self._h, _dev_ptr(t), self._batch(t), _stream_handle(stream)))
        return t

    def inverse_(self, t, stream=None):
        check(lib().mfhe_ntt_inverse(self._h, _dev_ptr(t), self._batch(t), _stream_handle(stream)))
        return t

    def forward_limbs_(self, t, limb_start: int, limb_count: int, stream=None):
        check(lib().mfhe_ntt_forward_limbs(self._h, _dev_ptr(t), self._batch(t, limb_count), limb_start,
                                           limb_count, _stream_handle(stream)))
        return t

    def inverse_limbs_(self, t, limb_start: int, limb_count: int, stream=None):
        check(lib().mfhe_ntt_inverse_limbs(self._h, _dev_ptr(t), self._batch(t, limb_count), limb_start,
                                           limb_count, _stream_handle(stream)))
        return t

    # host numpy arrays (uint64, C-contiguous), in place, synchronous --------------------
    def forward_host_(self, a):
        check(lib().mfhe_ntt_forward_host(self._h, a.ctypes.data_as(_vp), a.size // (self.limbs * self.n)))
        return a

    def inverse_host_(self, a):
        check(lib().mfhe_ntt_inverse_host(self._h, a.ctypes.data_as(_vp), a.size // (self.limbs * self.n)))
        return a

    def _pw(self, fn, a, b, out, stream):
        out = a if out is None else out
        check(fn(self._h, _dev_ptr(a), _dev_ptr(b), _dev_ptr(out), self._batch(a), _stream_handle(stream)))
        return out

    def mul(self, a, b, out=None, stream=None):
        return self._pw(lib().mfhe_poly_mul, a, b, out, stream)

    def mul_intt_(self, a, b, b_is_ntt=False, stream=None):
        """a <- INTT(NTT(a) . NTT(b)) in place; b is left in NTT form (mfhe_ntt_mul_intt)."""
        check(lib().mfhe_ntt_mul_intt(self._h, _dev_ptr(a), _dev_ptr(b), int(bool(b_is_ntt)), self._batch(a),
                                      _stream_handle(stream)))
        return a

    def add(self, a, b, out=None, stream=None):
        return self._pw(lib().mfhe_poly_add, a, b, out, stream)

    def sub(self, a, b, out=None, stream=None):
        return self._pw(lib().mfhe_poly_sub, a, b, out, stream)
