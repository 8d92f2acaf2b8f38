"""Host-side mirror of the reference's matrix-pipeline API (include/core/HE.cuh, encoder.cuh,
batched_encoder.cuh, trace.cuh, batched_trace.cuh) above the C ABI of libmfhe_b200.so.

Every method takes and returns torch CUDA tensors (uint64 data is carried as int64 storage,
complex as complex128) and enqueues work on the current torch stream.  There is no CPU path.
Method names follow the reference's functions; docstrings cite the function each one stands for.
"""
from __future__ import annotations

import ctypes as C

from . import MfheError, _dev_ptr, _stream_handle, check, lib

MATRIX, POLY = 0, 1
REF_MODULI = [17592186435073, 17182765057, 17184541441, 17186120449, 17186515201, 17186909953,
              17188883713, 17190462721, 17190857473, 17191844353, 17192831233]       # config.h:31-43
REF_DELTA = 34359738368.0                                                              # config.h:25

_vp, _u64, _ci = C.c_void_p, C.c_uint64, C.c_int


class _Params(C.Structure):
    _fields_ = [("n", _ci), ("f", _ci), ("num_limbs", _ci), ("moduli", C.POINTER(_u64)),
                ("delta", C.c_double), ("arena_bytes", C.c_size_t)]


_ready = False


def _L():
    global _ready
    L = lib()
    if _ready:
        return L
    L.mfhe_ctx_create.argtypes = [C.POINTER(_Params), C.POINTER(_vp)]
    L.mfhe_ctx_destroy.argtypes = [_vp]; L.mfhe_ctx_destroy.restype = None
    L.mfhe_ctx_info.argtypes = [_vp, C.POINTER(_ci)]
    L.mfhe_ctx_get_moduli.argtypes = [_vp, _vp, _ci]
    L.mfhe_ctx_eta.argtypes = [_vp, _ci]; L.mfhe_ctx_eta.restype = _u64
    L.mfhe_ctx_get_wcrt_table.argtypes = [_vp, _ci, _ci, _vp]
    L.mfhe_ctx_get_crt.argtypes = [_vp, _ci, _vp]
    L.mfhe_ctx_xy_plan.argtypes = [_vp]; L.mfhe_ctx_xy_plan.restype = _vp
    sig = {
        "mfhe_wcrt_forward": [_vp, _vp, _vp, _ci, _ci, _vp], "mfhe_wcrt_inverse": [_vp, _vp, _vp, _ci, _ci, _vp],
        "mfhe_wcrt_forward_vector": [_vp, _vp, _vp, _vp],
        "mfhe_wcrt_forward_centered": [_vp, _vp, _vp, _vp], "mfhe_wcrt_inverse_centered": [_vp, _vp, _vp, _vp],
        "mfhe_layout_convert": [_vp, _vp, _vp, _ci, _ci, _vp],
        "mfhe_wdft_forward_complex": [_vp, _vp, _vp, _vp], "mfhe_wdft_inverse_complex": [_vp, _vp, _vp, _vp],
        "mfhe_wdft_forward_centered_pair": [_vp, _vp, _vp, _vp, _vp, _vp],
        "mfhe_wdft_inverse_pair": [_vp, _vp, _vp, _vp, _vp, _vp],
        "mfhe_xy_idft2": [_vp, _vp, _vp, _ci, _vp], "mfhe_xy_dft2": [_vp, _vp, _vp, _ci, _vp],
        "mfhe_matmul_complex": [_vp, _vp, _vp, _ci, _vp],
        "mfhe_quantize": [_vp, _vp, _vp, _vp, _ci, _vp],
        "mfhe_crt_compose_big": [_vp, _vp, _vp, _vp, _ci, _ci, _ci, _vp],
        "mfhe_crt_compose_i64": [_vp, _vp, _vp, _ci, _ci, _vp],
        "mfhe_dequantize_exact": [_vp, _vp, _vp, _vp, _ci, _ci, _vp],
        "mfhe_big_pair_to_complex": [_vp, _vp, _vp, _vp, _vp, _vp, C.c_size_t, _vp],
        "mfhe_encode": [_vp, _vp, _vp, _vp, _vp], "mfhe_encode_lane": [_vp, _vp, _vp, _vp, _vp],
        "mfhe_decode_eval_pair": [_vp, _vp, _vp, _ci, _vp, _vp],
        "mfhe_decode_lane_from_rns_eval": [_vp, _vp, _vp, _vp, _vp],
        "mfhe_sample_uniform": [_vp, _vp, _vp], "mfhe_sample_gaussian": [_vp, _vp, _vp],
        "mfhe_sample_ternary": [_vp, _vp, _vp], "mfhe_keygen": [_vp, _vp, _vp],
        "mfhe_encrypt_pair": [_vp, _vp, _vp, _vp, _vp, _vp, _vp],
        "mfhe_encrypt_pair_with": [_vp] * 10,
        "mfhe_encode_encrypt_pair": [_vp] * 6,
        "mfhe_encrypt": [_vp, _vp, _vp, _vp, _vp],
        "mfhe_decrypt_to_eval": [_vp, _vp, _vp, _vp, _ci, _vp],
        "mfhe_decrypt_decode": [_vp, _vp, _vp, _vp, _vp, _vp],
        "mfhe_ct_add": [_vp, _vp, _vp, _vp, _ci, _vp],
        "mfhe_ct_mul_raw": [_vp, _vp, _vp, _vp, _vp, _vp, _ci, _vp],
        "mfhe_trace_map": [_vp, _vp, _vp, _vp, _vp, _ci, _vp],
        "mfhe_trace_gemm": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _ci, _vp],
        "mfhe_trace_rescale": [_vp, _vp, _vp, _ci, _u64, _u64, _u64, _vp],
        "mfhe_rescale_last": [_vp, _vp, _vp, _vp],
        "mfhe_modup": [_vp, _vp, _vp, _vp, C.c_size_t, _vp],
        "mfhe_moddown": [_vp, _vp, _vp, _vp, _vp, C.c_size_t, _vp],
        "mfhe_ctx_set_sampler": [_vp, _ci, C.POINTER(_u64)],
        "mfhe_sampler_cdt": [C.POINTER(_u64), _ci],
        "mfhe_trace_map_x": [_vp, _vp, _vp, _vp, _vp, _ci, _vp],
        "mfhe_trace_gemm_ty": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _ci, _vp],
        "mfhe_transpose": [_vp, _vp, _vp, C.c_size_t, _vp],
        "mfhe_trace_rescale_limbs": [_vp, _vp, _vp, _ci, C.POINTER(_u64), _vp],
        "mfhe_ct_matmul": [_vp, _vp, _vp, _vp, _vp, _vp, _vp],
        "mfhe_ct_matmul_decrypt": [_vp, _vp, _vp, _vp, _vp, _vp],
        "mfhe_gl_perm_tables": [_ci, _vp, _vp],
        "mfhe_apply_gl_perm": [_vp, _vp, _ci, _ci, _ci, _ci, _vp],
        "mfhe_bconv_plan_create": [_ci, C.POINTER(_u64), _ci, C.POINTER(_u64), C.POINTER(_vp)],
        "mfhe_bconv": [_vp, _vp, _vp, C.c_size_t, C.c_size_t, _vp],
        "mfhe_bconv_partial": [_vp, _vp, _ci, _ci, _vp, C.c_size_t, C.c_size_t, _vp],
        "mfhe_bconv_partial_scatter": [_vp, _vp, _ci, _ci, C.POINTER(_vp), C.POINTER(C.c_size_t), C.c_size_t, C.c_size_t, _vp],
        "mfhe_bconv_reduce": [_vp, _vp, _ci, _ci, _ci, _vp, C.c_size_t, C.c_size_t, _vp],
    }
    for name, args in sig.items():
        fn = getattr(L, name)
        fn.restype = _ci
        fn.argtypes = args
    L.mfhe_bconv_plan_destroy.argtypes = [_vp]; L.mfhe_bconv_plan_destroy.restype = None
    _ready = True
    return L


def _any_ptr(t):
    """Device pointer of any contiguous CUDA tensor (complex128 / float64 / int64 / uint8)."""
    import torch
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.is_contiguous()):
        raise MfheError("expected a contiguous CUDA tensor (no CPU path exists)")
    return _vp(t.data_ptr())


class Context:
    """init_he_backend (src/core/HE.cu:318-408) as an object: all tables and the scratch arena."""

    def __init__(self, n=64, f=257, moduli=None, delta=REF_DELTA, arena_bytes=0):
        mods = [int(m) for m in (REF_MODULI if moduli is None else moduli)]
        self._mods = (_u64 * len(mods))(*mods)
        prm = _Params(n, f, len(mods), self._mods, float(delta), arena_bytes)
        h = _vp()
        check(_L().mfhe_ctx_create(C.byref(prm), C.byref(h)))
        self._h = h
        info = (_ci * 6)()
        check(_L().mfhe_ctx_info(h, info))
        self.n, self.f, self.p, self.phi, self.L, self.big_words = list(info)
        self.n2 = self.n * self.n
        self.T = self.phi * self.L * self.n2
        self.moduli, self.delta = mods, float(delta)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _L().mfhe_ctx_destroy(h)
            except Exception:
                pass

    # -- helpers ---------------------------------------------------------------------------------
    def _new(self, numel, dtype=None, like=None):
        import torch
        return torch.empty(numel, dtype=dtype or torch.int64, device=like.device if like is not None else "cuda")

    def set_row_arith(self, arith: int) -> None:
        """Arithmetic of the fused encrypt / decrypt row kernel (X transforms + key product): 0 integer, 1 FP64 pipe
        (the default when every modulus is below 2^50).  Same words either way; for A/B runs and parity tests."""
        import ctypes as C
        L = lib()
        L.mfhe_ntt_plan_set_arith.argtypes = [_vp, C.c_int]
        check(L.mfhe_ntt_plan_set_arith(L.mfhe_ctx_xy_plan(self._h), arith))

    @property
    def row_arith(self) -> int:
        L = lib()
        L.mfhe_ntt_plan_arith.argtypes = [_vp]
        return L.mfhe_ntt_plan_arith(L.mfhe_ctx_xy_plan(self._h))

    def eta(self, limb):
        return int(_L().mfhe_ctx_eta(self._h, limb))

    def copy_device_moduli(self):
        """copy_device_moduli (HE.cu:410-422)."""
        buf = (_u64 * self.L)()
        check(_L().mfhe_ctx_get_moduli(self._h, buf, self.L))
        return list(buf)

    def wcrt_table(self, limb, inverse=False):
        import numpy as np
        a = np.zeros((self.phi, self.phi), dtype=np.uint64)
        check(_L().mfhe_ctx_get_wcrt_table(self._h, limb, int(inverse), a.ctypes.data_as(_vp)))
        return a

    def crt_tables(self):
        import numpy as np
        W, L = self.big_words, self.L
        out = {}
        for k, (name, shape) in enumerate([("Q", (W,)), ("Qhalf", (W,)), ("M", (L, W)), ("inv", (L,))]):
            a = np.zeros(shape, dtype=np.uint64)
            check(_L().mfhe_ctx_get_crt(self._h, k, a.ctypes.data_as(_vp)))
            out[name] = a
        return out

    # -- W-CRT -------------------------------------------------------------------------------------
    def wntt_forward_matrix(self, x, in_layout=MATRIX, out_layout=POLY):
        """wntt_forward_matrix (HE.cu:437-442): matrix-major W-coeff -> poly-major W-eval."""
        o = self._new(self.T, like=x)
        check(_L().mfhe_wcrt_forward(self._h, _dev_ptr(x), _dev_ptr(o), in_layout, out_layout, _stream_handle()))
        return o

    def wntt_inverse_matrix(self, x, in_layout=POLY, out_layout=MATRIX):
        """wntt_inverse_matrix (HE.cu:444-452)."""
        o = self._new(self.T, like=x)
        check(_L().mfhe_wcrt_inverse(self._h, _dev_ptr(x), _dev_ptr(o), in_layout, out_layout, _stream_handle()))
        return o

    def wntt_forward_vector(self, x):
        o = self._new(self.phi * self.L * self.n, like=x)
        check(_L().mfhe_wcrt_forward_vector(self._h, _dev_ptr(x), _dev_ptr(o), _stream_handle()))
        return o

    def wntt_forward_centered(self, x):
        o = self._new(self.phi * self.n2, like=x)
        check(_L().mfhe_wcrt_forward_centered(self._h, _dev_ptr(x), _dev_ptr(o), _stream_handle()))
        return o

    def wntt_inverse_centered(self, x):
        o = self._new(self.phi * self.n2, like=x)
        check(_L().mfhe_wcrt_inverse_centered(self._h, _dev_ptr(x), _dev_ptr(o), _stream_handle()))
        return o

    def layout_convert(self, x, in_layout, out_layout):
        o = self._new(self.T, like=x)
        check(_L().mfhe_layout_convert(self._h, _dev_ptr(x), _dev_ptr(o), in_layout, out_layout, _stream_handle()))
        return o

    # -- complex W-DFT --------------------------------------------------------------------------------
    def _cx(self, fn, x):
        import torch
        o = torch.empty_like(x)
        check(fn(self._h, _any_ptr(x), _any_ptr(o), _stream_handle()))
        return o

    def wdft_forward_complex(self, x):
        return self._cx(_L().mfhe_wdft_forward_complex, x)

    def wdft_inverse_complex(self, x):
        """w_idft_kernel (batched_encoder.cu:104-123)."""
        return self._cx(_L().mfhe_wdft_inverse_complex, x)

    def wdft_forward_centered_pair(self, re, im):
        import torch
        ore = torch.empty(re.numel(), dtype=torch.float64, device=re.device); oim = torch.empty_like(ore)
        check(_L().mfhe_wdft_forward_centered_pair(self._h, _any_ptr(re), _any_ptr(im), _any_ptr(ore), _any_ptr(oim), _stream_handle()))
        return ore, oim

    def wdft_inverse_pair(self, re, im):
        import torch
        ore = torch.empty_like(re); oim = torch.empty_like(im)
        check(_L().mfhe_wdft_inverse_pair(self._h, _any_ptr(re), _any_ptr(im), _any_ptr(ore), _any_ptr(oim), _stream_handle()))
        return ore, oim

    # -- XY embedding / quantise / CRT ------------------------------------------------------------------
    def idft2(self, x, lanes):
        """Encoder::idft2 (encoder.cu:460-467) for `lanes` matrices."""
        import torch
        o = torch.empty_like(x)
        check(_L().mfhe_xy_idft2(self._h, _any_ptr(x), _any_ptr(o), lanes, _stream_handle()))
        return o

    def dft2(self, x, lanes):
        """Encoder::decode_from_eval_complex (encoder.cu:492-501) for `lanes` matrices."""
        import torch
        o = torch.empty_like(x)
        check(_L().mfhe_xy_dft2(self._h, _any_ptr(x), _any_ptr(o), lanes, _stream_handle()))
        return o

    def quantize(self, x, lanes):
        re = self._new(lanes * self.L * self.n2, like=x); im = self._new(lanes * self.L * self.n2, like=x)
        check(_L().mfhe_quantize(self._h, _any_ptr(x), _dev_ptr(re), _dev_ptr(im), lanes, _stream_handle()))
        return re, im

    def crt_compose_centerlift_big(self, x, n2=None, limbs=None, lanes=1):
        import torch
        n2 = self.n2 if n2 is None else n2; limbs = self.L if limbs is None else limbs
        mag = self._new(lanes * n2 * self.big_words, like=x)
        neg = torch.empty(lanes * n2, dtype=torch.uint8, device=x.device)
        check(_L().mfhe_crt_compose_big(self._h, _dev_ptr(x), _dev_ptr(mag), _any_ptr(neg), n2, limbs, lanes, _stream_handle()))
        return mag, neg

    def crt_compose_centerlift(self, x, n2=None, limbs=None):
        n2 = self.n2 if n2 is None else n2; limbs = self.L if limbs is None else limbs
        o = self._new(n2, like=x)
        check(_L().mfhe_crt_compose_i64(self._h, _dev_ptr(x), _dev_ptr(o), n2, limbs, _stream_handle()))
        return o

    def dequantize_exact(self, re, im, n2=None, limbs=None):
        import torch
        n2 = self.n2 if n2 is None else n2; limbs = self.L if limbs is None else limbs
        o = torch.empty(n2, dtype=torch.complex128, device=re.device)
        check(_L().mfhe_dequantize_exact(self._h, _dev_ptr(re), _dev_ptr(im), _any_ptr(o), n2, limbs, _stream_handle()))
        return o

    def big_pair_to_complex(self, rm, rn, im, in_):
        import torch
        o = torch.empty(rn.numel(), dtype=torch.complex128, device=rm.device)
        check(_L().mfhe_big_pair_to_complex(self._h, _dev_ptr(rm), _any_ptr(rn), _dev_ptr(im), _any_ptr(in_), _any_ptr(o),
                                            rn.numel(), _stream_handle()))
        return o

    def encode_to_wntt_eval(self, msg):
        """BatchedEncoder::encode_to_wntt_eval (batched_encoder.cu:161-228)."""
        re = self._new(self.T, like=msg); im = self._new(self.T, like=msg)
        check(_L().mfhe_encode(self._h, _any_ptr(msg), _dev_ptr(re), _dev_ptr(im), _stream_handle()))
        return re, im

    def encode_lane(self, msg):
        """Encoder::encode (encoder.cu:446-456)."""
        re = self._new(self.L * self.n2, like=msg); im = self._new(self.L * self.n2, like=msg)
        check(_L().mfhe_encode_lane(self._h, _any_ptr(msg), _dev_ptr(re), _dev_ptr(im), _stream_handle()))
        return re, im

    def decode_eval_pair(self, er, ei, layout=POLY):
        import torch
        o = torch.empty(self.phi * self.n2, dtype=torch.complex128, device=er.device)
        check(_L().mfhe_decode_eval_pair(self._h, _dev_ptr(er), _dev_ptr(ei), layout, _any_ptr(o), _stream_handle()))
        return o

    def decode_lane_from_rns_eval(self, re, im):
        import torch
        o = torch.empty(self.n2, dtype=torch.complex128, device=re.device)
        check(_L().mfhe_decode_lane_from_rns_eval(self._h, _dev_ptr(re), _dev_ptr(im), _any_ptr(o), _stream_handle()))
        return o

    def set_sampler(self, mode, seed=(0, 0)):
        """mode 0: the reference's index hashes (parity); 1: Philox4x32-10 keyed by the 128-bit seed (mfhe_ctx_set_sampler)."""
        arr = (_u64 * 2)(int(seed[0]), int(seed[1]))
        check(_L().mfhe_ctx_set_sampler(self._h, int(mode), arr))

    # -- samplers / keys / encrypt / decrypt ----------------------------------------------------------------
    def sample_uniform(self):
        o = self._new(self.T); check(_L().mfhe_sample_uniform(self._h, _dev_ptr(o), _stream_handle())); return o

    def sample_gaussian(self):
        o = self._new(self.T); check(_L().mfhe_sample_gaussian(self._h, _dev_ptr(o), _stream_handle())); return o

    def sample_ternary(self):
        o = self._new(self.phi * self.L * self.n); check(_L().mfhe_sample_ternary(self._h, _dev_ptr(o), _stream_handle())); return o

    def generate_secret_key(self):
        """generate_secret_key (HE.cu:1272-1307)."""
        o = self._new(self.phi * self.L * self.n); check(_L().mfhe_keygen(self._h, _dev_ptr(o), _stream_handle())); return o

    def allocate_ciphertext(self):
        """allocate_ciphertext (HE.cu:675-682): zero-filled [b | a]."""
        import torch
        return torch.zeros(2 * self.T, dtype=torch.int64, device="cuda")

    def encrypt_pair(self, m_re, m_im, sk, ct_re=None, ct_im=None):
        ct_re = self._new(2 * self.T, like=m_re) if ct_re is None else ct_re
        ct_im = self._new(2 * self.T, like=m_re) if ct_im is None else ct_im
        check(_L().mfhe_encrypt_pair(self._h, _dev_ptr(m_re), _dev_ptr(m_im), _dev_ptr(sk), _dev_ptr(ct_re), _dev_ptr(ct_im),
                                     _stream_handle()))
        return ct_re, ct_im

    def encrypt_pair_with(self, m_re, m_im, sk, a, e_re, e_im):
        ct_re = self._new(2 * self.T, like=m_re); ct_im = self._new(2 * self.T, like=m_re)
        check(_L().mfhe_encrypt_pair_with(self._h, _dev_ptr(m_re), _dev_ptr(m_im), _dev_ptr(sk), _dev_ptr(a), _dev_ptr(e_re),
                                          _dev_ptr(e_im), _dev_ptr(ct_re), _dev_ptr(ct_im), _stream_handle()))
        return ct_re, ct_im

    def encode_encrypt_pair(self, msg, sk, ct_re=None, ct_im=None):
        """encode_to_wntt_eval + encrypt_pair fused (same ciphertext words; message and error share a W-CRT)."""
        ct_re = self._new(2 * self.T) if ct_re is None else ct_re
        ct_im = self._new(2 * self.T) if ct_im is None else ct_im
        check(_L().mfhe_encode_encrypt_pair(self._h, _any_ptr(msg), _dev_ptr(sk), _dev_ptr(ct_re), _dev_ptr(ct_im),
                                            _stream_handle()))
        return ct_re, ct_im

    def encrypt(self, m, sk):
        ct = self._new(2 * self.T, like=m)
        check(_L().mfhe_encrypt(self._h, _dev_ptr(m), _dev_ptr(sk), _dev_ptr(ct), _stream_handle()))
        return ct

    def decrypt_to_eval(self, ct, sk, layout=POLY):
        o = self._new(self.T, like=ct)
        check(_L().mfhe_decrypt_to_eval(self._h, _dev_ptr(ct), _dev_ptr(sk), _dev_ptr(o), layout, _stream_handle()))
        return o

    def decrypt_and_decode(self, ct_re, ct_im, sk, out=None):
        import torch
        o = torch.empty(self.phi * self.n2, dtype=torch.complex128, device=ct_re.device) if out is None else out
        check(_L().mfhe_decrypt_decode(self._h, _dev_ptr(ct_re), _dev_ptr(ct_im), _dev_ptr(sk), _any_ptr(o), _stream_handle()))
        return o

    def add_ciphertexts(self, c1, c2, reference_indexing=True):
        o = self._new(2 * self.T, like=c1)
        check(_L().mfhe_ct_add(self._h, _dev_ptr(c1), _dev_ptr(c2), _dev_ptr(o), int(reference_indexing), _stream_handle()))
        return o

    def multiply_ciphertexts_raw(self, c1, c2, reference_indexing=True):
        d = [self._new(self.T, like=c1) for _ in range(3)]
        check(_L().mfhe_ct_mul_raw(self._h, _dev_ptr(c1), _dev_ptr(c2), *[_dev_ptr(x) for x in d], int(reference_indexing),
                                   _stream_handle()))
        return d

    # -- trace step -----------------------------------------------------------------------------------------
    def map_B_to_Bprime(self, Br, Bi, batch=1):
        import torch
        o = [torch.empty_like(Br), torch.empty_like(Bi)]
        check(_L().mfhe_trace_map(self._h, _dev_ptr(Br), _dev_ptr(Bi), _dev_ptr(o[0]), _dev_ptr(o[1]), batch, _stream_handle()))
        return o

    def trace_gemm(self, Ar, Ai, Bpr, Bpi, batch=1):
        import torch
        o = [torch.empty_like(Ar), torch.empty_like(Ai)]
        check(_L().mfhe_trace_gemm(self._h, _dev_ptr(Ar), _dev_ptr(Ai), _dev_ptr(Bpr), _dev_ptr(Bpi), _dev_ptr(o[0]),
                                   _dev_ptr(o[1]), batch, _stream_handle()))
        return o

    def rescale_by_delta(self, Cr, Ci, inv0, inv1, inv2, batch=1):
        check(_L().mfhe_trace_rescale(self._h, _dev_ptr(Cr), _dev_ptr(Ci), batch, inv0, inv1, inv2, _stream_handle()))
        return Cr, Ci


    # -- encrypted matrix product (include/mfhe.h) -----------------------------------------------
    def trace_map_x(self, Br, Bi, batch=None):
        import torch
        batch = self.phi if batch is None else batch
        o = [torch.empty_like(Br), torch.empty_like(Bi)]
        check(_L().mfhe_trace_map_x(self._h, _dev_ptr(Br), _dev_ptr(Bi), _dev_ptr(o[0]), _dev_ptr(o[1]), batch, _stream_handle()))
        return o

    def trace_gemm_ty(self, Ur, Ui, Vr, Vi, batch=None):
        import torch
        batch = self.phi if batch is None else batch
        o = [torch.empty_like(Ur), torch.empty_like(Ui)]
        check(_L().mfhe_trace_gemm_ty(self._h, _dev_ptr(Ur), _dev_ptr(Ui), _dev_ptr(Vr), _dev_ptr(Vi), _dev_ptr(o[0]),
                                      _dev_ptr(o[1]), batch, _stream_handle()))
        return o

    def transpose(self, x):
        import torch
        o = torch.empty_like(x)
        check(_L().mfhe_transpose(self._h, _dev_ptr(x), _dev_ptr(o), x.numel() // self.n2, _stream_handle()))
        return o

    def rescale_limbs(self, Cr, Ci, inv, batch=None):
        batch = self.phi if batch is None else batch
        arr = (_u64 * self.L)(*[int(v) for v in inv])
        check(_L().mfhe_trace_rescale_limbs(self._h, _dev_ptr(Cr), _dev_ptr(Ci), batch, arr, _stream_handle()))
        return Cr, Ci

    def ct_matmul(self, cu_re, cu_im, cv_re, cv_im):
        """The four component GEMMs of ct_U x ct_V' (V' = trace_map_x of the encoded V, then encrypted): 8 poly sets."""
        o = self._new(8 * self.T, like=cu_re)
        check(_L().mfhe_ct_matmul(self._h, _dev_ptr(cu_re), _dev_ptr(cu_im), _dev_ptr(cv_re), _dev_ptr(cv_im), _dev_ptr(o),
                                  _stream_handle()))
        return o

    def ct_matmul_decrypt(self, prod, sk):
        """Consumes `prod`; returns evals (re, im) [w][l][z][x] of n * sum_y U[y][x] V'[y][z]."""
        o = [self._new(self.T, like=prod), self._new(self.T, like=prod)]
        check(_L().mfhe_ct_matmul_decrypt(self._h, _dev_ptr(prod), _dev_ptr(sk), _dev_ptr(o[0]), _dev_ptr(o[1]), _stream_handle()))
        return o

    def rescale_last(self, x):
        """mfhe_rescale_last: [w][L][n^2] -> [w][L-1][n^2], divided by the last modulus."""
        o = self._new(self.phi * (self.L - 1) * self.n2, like=x)
        check(_L().mfhe_rescale_last(self._h, _dev_ptr(x), _dev_ptr(o), _stream_handle()))
        return o

    def conj_lane(self):
        """l -> l' with exp[l'] = -exp[l] mod p (see encrypted_matmul)."""
        import numpy as np
        e = [((a * self.f + 3 * b) % self.p) for a in (1, 2) for b in range(1, self.f)]      # k_wntt_exp, HE.cu:72-105
        pos = {v: i for i, v in enumerate(e)}
        return np.asarray([pos[(self.p - v) % self.p] for v in e], dtype=np.int64)

    def encrypted_matmul(self, U, V, sk):
        """Encrypted matrix product driver: per lane l, R_l = V_l^H . U_l for complex n x n matrices U, V ([phi][n][n]
        complex128 device tensors): encode -> (V: lane permutation + X twist) -> encrypt both -> four GEMMs on the
        ciphertext components (tensor cores) -> 4-component decrypt -> decode.  Returns R [phi][n][n] complex128."""
        import torch
        cl = torch.from_numpy(self.conj_lane()).to(U.device)
        ur, ui = self.encode_to_wntt_eval(U.reshape(-1))
        vr, vi = self.encode_to_wntt_eval(V.reshape(self.phi, -1)[cl].reshape(-1))
        vpr, vpi = self.trace_map_x(vr, vi)
        cu = self.encrypt_pair(ur, ui, sk)
        cv = self.encrypt_pair(vpr, vpi, sk)
        prod = self.ct_matmul(cu[0], cu[1], cv[0], cv[1])
        pr, pi = self.ct_matmul_decrypt(prod, sk)
        out = self.decode_eval_pair(pr, pi, MATRIX)            # divides by delta once; the product carries delta^2
        return (out / self.delta).reshape(self.phi, self.n, self.n)


def modup(plan_qp, q_to_p, x, batch):
    """mfhe_modup: x [batch][LQ][N] coefficient form -> [batch][LQ+LP][N] NTT form (plan_qp: NttPlan over Q then P)."""
    import torch
    o = torch.empty(batch * plan_qp.limbs * plan_qp.n, dtype=torch.int64, device=x.device)
    check(_L().mfhe_modup(plan_qp._h, q_to_p._h, _dev_ptr(x), _dev_ptr(o), batch, _stream_handle()))
    return o


def moddown(plan_qp, p_to_q, x, batch):
    """mfhe_moddown: x [batch][LQ+LP][N] NTT form -> [batch][LQ][N] NTT form, divided by P."""
    import torch
    lq, lp = len(p_to_q.p), len(p_to_q.q)
    o = torch.empty(batch * lq * plan_qp.n, dtype=torch.int64, device=x.device)
    scratch = torch.empty(batch * lp * plan_qp.n, dtype=torch.int64, device=x.device)
    check(_L().mfhe_moddown(plan_qp._h, p_to_q._h, _dev_ptr(x), _dev_ptr(o), _dev_ptr(scratch), batch, _stream_handle()))
    return o


def sampler_cdt():
    import numpy as np
    arr = (_u64 * 83)()
    check(_L().mfhe_sampler_cdt(arr, 83))
    return np.asarray(list(arr), dtype=np.uint64)


def gl_perm_tables(n):
    import numpy as np
    p = np.zeros(n, dtype=np.uint32); ip = np.zeros(n, dtype=np.uint32)
    check(_L().mfhe_gl_perm_tables(n, p.ctypes.data_as(_vp), ip.ctypes.data_as(_vp)))
    return p, ip


def apply_gl_perm(x, limbs, batch_count, n, inverse=False):
    """apply_gl_perm (src/core/ntt_core.cu:433-441)."""
    import torch
    o = torch.empty_like(x)
    check(_L().mfhe_apply_gl_perm(_dev_ptr(x), _dev_ptr(o), limbs, batch_count, n, int(inverse), _stream_handle()))
    return o


class BconvPlan:
    """Phantom DBaseConverter (ph/src/rns_bconv.cu:212-229) for one (input base, output base) pair."""

    def __init__(self, q_in, p_out):
        self.q = [int(v) for v in q_in]; self.p = [int(v) for v in p_out]
        qa = (_u64 * len(self.q))(*self.q); pa = (_u64 * len(self.p))(*self.p)
        h = _vp()
        check(_L().mfhe_bconv_plan_create(len(self.q), qa, len(self.p), pa, C.byref(h)))
        self._h = h

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _L().mfhe_bconv_plan_destroy(h)
            except Exception:
                pass

    def convert(self, x, N, batch=1):
        import torch
        o = torch.empty(batch * len(self.p) * N, dtype=torch.int64, device=x.device)
        check(_L().mfhe_bconv(self._h, _dev_ptr(x), _dev_ptr(o), N, batch, _stream_handle()))
        return o

    def partial(self, x_local, in_start, in_count, N, batch=1):
        import torch
        o = torch.empty(batch * len(self.p) * N, dtype=torch.int64, device=x_local.device)
        check(_L().mfhe_bconv_partial(self._h, _dev_ptr(x_local), in_start, in_count, _dev_ptr(o), N, batch, _stream_handle()))
        return o

    def partial_scatter(self, x_local, in_start, in_count, dst_ptrs, dst_batch_strides, N, batch=1):
        """mfhe_bconv_partial_scatter: output limb j of item b is stored at dst_ptrs[j] + b * dst_batch_strides[j]
        words; dst_ptrs are raw device addresses (own or peer memory), one per output limb."""
        if len(dst_ptrs) != len(self.p) or len(dst_batch_strides) != len(self.p):
            raise ValueError("partial_scatter: one destination and stride per output limb")
        pa = (_vp * len(self.p))(*[_vp(int(v)) for v in dst_ptrs])
        sa = (C.c_size_t * len(self.p))(*[int(v) for v in dst_batch_strides])
        check(_L().mfhe_bconv_partial_scatter(self._h, _dev_ptr(x_local), in_start, in_count, pa, sa, N, batch, _stream_handle()))

    def reduce(self, parts, nparts, out_start, out_count, N, batch=1):
        import torch
        o = torch.empty(batch * out_count * N, dtype=torch.int64, device=parts.device)
        check(_L().mfhe_bconv_reduce(self._h, _dev_ptr(parts), nparts, out_start, out_count, _dev_ptr(o), N, batch, _stream_handle()))
        return o
