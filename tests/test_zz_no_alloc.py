"""mfhe_moddown_static: the mod-down entry point that allocates nothing (its scratch is a static device array inside the
library, no counterpart in the reference).  Called through the C ABI with
ctypes and compared word for word (integer path: bit-exact, the tolerance of every integer test here) with mfhe_moddown
given caller scratch, on two calls in a row: the second call uses a smaller, different input, so stale contents of the
shared array from the first call would show.  The size limit is checked too.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu

STATIC_WORDS = 1 << 21                                     # MFHE_MODDOWN_STATIC_WORDS (include/mfhe.h)


def test_moddown_static_equals_moddown_twice_and_rejects_oversize(cuda_device):
    import torch
    import matrix_fhe_gpu_b200 as m
    import matrix_fhe_gpu_b200.scheme as S
    from matrix_fhe_gpu_b200 import _dev_ptr, _stream_handle, check, lib

    L = lib()
    vp = C.c_void_p
    L.mfhe_moddown.argtypes = [vp, vp, vp, vp, vp, C.c_size_t, vp]
    L.mfhe_moddown.restype = C.c_int
    L.mfhe_moddown_static.argtypes = [vp, vp, vp, vp, C.c_size_t, vp]
    L.mfhe_moddown_static.restype = C.c_int

    logn, lq, lp, bits = 13, 4, 2, 55
    N = 1 << logn
    primes = [int(v) for v in O.get_primes(2 * N, bits, lq + lp)]
    q, p = primes[:lq], primes[lq:]
    plan = m.NttPlan(logn, q + p)
    conv = S.BconvPlan(p, q)

    def to_dev(x):
        return torch.from_numpy(np.ascontiguousarray(x).view(np.int64)).to(cuda_device)

    def both(seed, batch):
        y = O.synth_polys(seed, batch, q + p, N)            # arbitrary NTT-form element over Q u P
        x = to_dev(y)
        ref = torch.zeros(batch * lq * N, dtype=torch.int64, device=cuda_device)
        scratch = torch.empty(batch * lp * N, dtype=torch.int64, device=cuda_device)
        check(L.mfhe_moddown(plan._h, conv._h, _dev_ptr(x), _dev_ptr(ref), _dev_ptr(scratch), batch, _stream_handle()))
        out = torch.full((batch * lq * N,), -1, dtype=torch.int64, device=cuda_device)
        # everything the new entry point needs exists before the call; the call itself gets only inputs and outputs
        check(L.mfhe_moddown_static(plan._h, conv._h, _dev_ptr(x), _dev_ptr(out), batch, _stream_handle()))
        torch.cuda.synchronize()
        assert torch.equal(out, ref), f"seed {seed} batch {batch}"
        assert (x.cpu().numpy().view(np.uint64).reshape(y.shape) == y).all()      # input untouched
        return out.cpu().numpy().view(np.uint64).reshape(batch, lq, N), y

    out1, y1 = both(5, 3)                                   # first call: fills 3 * lp * N words of the shared array
    out2, _ = both(9, 1)                                    # second call, smaller and different: stale words beyond it
    out3, _ = both(5, 3)                                    # and the first input again: same words as the first call
    assert (out3 == out1).all()

    # one poly of the first call against the oracle (INTT -> bConv_{P->Q} -> NTT -> (x - t) / P)
    tq, tqp = O.PhantomTables(logn, q), O.PhantomTables(logn, q + p)
    coeff = tqp.inverse(y1[:1])
    t = tq.forward(O.bconv(coeff[0, lq:], p, q)[None])[0]
    P = 1
    for v in p:
        P *= v
    for i, qi in enumerate(q):
        want = ((y1[0, i].astype(object) - t[i].astype(object)) * pow(P % qi, -1, qi) % qi).astype(np.uint64)
        assert (out1[0, i] == want).all()

    # the limit: batch * LP * N beyond the static array is refused with a clear message, nothing is launched
    too_many = STATIC_WORDS // (lp * N) + 1
    dummy = torch.zeros(8, dtype=torch.int64, device=cuda_device)
    rc = L.mfhe_moddown_static(plan._h, conv._h, _dev_ptr(dummy), _dev_ptr(dummy), too_many, _stream_handle())
    assert rc != 0
    msg = L.mfhe_last_error().decode()
    assert "exceeds the static scratch" in msg and str(STATIC_WORDS) in msg
    # exactly at the limit the size check passes (NULL data is what is refused then)
    rc = L.mfhe_moddown_static(plan._h, conv._h, None, None, STATIC_WORDS // (lp * N), _stream_handle())
    assert rc != 0 and "NULL" in L.mfhe_last_error().decode()
    assert L.mfhe_moddown_static(plan._h, conv._h, None, None, 0, _stream_handle()) == 0      # empty batch: no-op
