"""compute-sanitizer is closed on this pool (profiles/r2_03_compute_sanitizer_closed.txt), so the two things it would have
shown are checked with our own means:

  * out-of-bounds WRITES: every output buffer of every kernel family sits between two canary regions that must be
    untouched afterwards (inputs sit between canaries too, so an in-place kernel cannot hide a stray store);
  * shared-memory / TMEM / mbarrier RACES: every kernel family runs 12 times on the same input and must give the same
    words each time (a hazard in the tcgen05 W-CRT's ring, the warp-private exchanges of the NTT row kernels or the staged
    digit tiles of the trace GEMMs shows up as run-to-run differences), on top of the bit-exact oracle comparisons.
"""
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
PAD = 4096
CANARY = -0x0123456789ABCDEF


class Guarded:
    """A device buffer of `words` int64 between two canary regions."""

    def __init__(self, words, device, fill=None):
        import torch
        self.raw = torch.full((words + 2 * PAD,), CANARY, dtype=torch.int64, device=device)
        self.view = self.raw[PAD:PAD + words]
        if fill is not None:
            self.view.copy_(fill)

    def intact(self):
        return bool((self.raw[:PAD] == CANARY).all()) and bool((self.raw[-PAD:] == CANARY).all())


def dev(x, device):
    import torch
    x = np.ascontiguousarray(x)
    return torch.from_numpy(x.view(np.int64) if x.dtype == np.uint64 else x).to(device)


def test_canaries_and_determinism_matrix_family(cuda_device, monkeypatch):
    import torch
    import matrix_fhe_gpu_b200 as m
    import matrix_fhe_gpu_b200.scheme as S
    from matrix_fhe_gpu_b200 import check, _dev_ptr, _stream_handle
    L_ = S._L()
    mods = O.REF_MODULI[:3]                       # 45-bit (6 digits) + two 35-bit (5 digits)
    g = S.Context(16, 257, mods, 2.0 ** 35)       # phi = 512: the tcgen05 W-CRT runs
    T = g.T
    a = g.sample_uniform()
    for path in ("umma", "mma", "cuda"):
        monkeypatch.setenv("MFHE_WCRT_PATH", path)
        ref = None
        for lin, lout in ((S.MATRIX, S.POLY), (S.POLY, S.MATRIX)):
            gi, go = Guarded(T, cuda_device, a), Guarded(T, cuda_device)
            outs = []
            for _ in range(12):
                go.view.fill_(0)
                check(L_.mfhe_wcrt_forward(g._h, _dev_ptr(gi.view), _dev_ptr(go.view), lin, lout, _stream_handle()))
                outs.append(go.view.clone())
            assert gi.intact() and go.intact(), f"W-CRT {path} wrote outside its output"
            assert all(torch.equal(outs[0], o) for o in outs[1:]), f"W-CRT {path} is not deterministic"
    monkeypatch.delenv("MFHE_WCRT_PATH")
    # fused row kernel (encrypt, decrypt pair), CRT + fp64 decode, trace GEMM (tensor cores need n = 64: second context)
    sk = g.generate_secret_key()
    ell, i = np.meshgrid(np.arange(g.phi, dtype=np.float64), np.arange(g.n2, dtype=np.float64), indexing="ij")
    msg = dev(((ell + 0.001 * i) * (1 - 1j)).reshape(-1), cuda_device)
    re, im = g.encode_to_wntt_eval(msg)
    cr, ci = Guarded(2 * T, cuda_device), Guarded(2 * T, cuda_device)
    dec = Guarded(2 * g.phi * g.n2, cuda_device)
    first = None
    for _ in range(12):
        check(L_.mfhe_encrypt_pair(g._h, _dev_ptr(re), _dev_ptr(im), _dev_ptr(sk), _dev_ptr(cr.view), _dev_ptr(ci.view), _stream_handle()))
        check(L_.mfhe_decrypt_decode(g._h, _dev_ptr(cr.view), _dev_ptr(ci.view), _dev_ptr(sk), _dev_ptr(dec.view), _stream_handle()))
        cur = (cr.view.clone(), ci.view.clone(), dec.view.clone())
        first = first or cur
        assert all(torch.equal(x, y) for x, y in zip(first, cur)), "encrypt / decrypt pipeline is not deterministic"
    assert cr.intact() and ci.intact() and dec.intact()
    g64 = S.Context(64, 5, O.primes_congruent_one(2 * 64 * 15, 45, 2), 2.0 ** 30)
    T64 = g64.T
    u = [g64.sample_uniform() for _ in range(2)]
    outs = [Guarded(T64, cuda_device), Guarded(T64, cuda_device)]
    first = None
    for _ in range(12):
        check(L_.mfhe_trace_gemm_ty(g64._h, _dev_ptr(u[0]), _dev_ptr(u[1]), _dev_ptr(u[1]), _dev_ptr(u[0]), _dev_ptr(outs[0].view),
                                    _dev_ptr(outs[1].view), g64.phi, _stream_handle()))
        cur = (outs[0].view.clone(), outs[1].view.clone())
        first = first or cur
        assert all(torch.equal(x, y) for x, y in zip(first, cur)), "trace GEMM is not deterministic"
    assert outs[0].intact() and outs[1].intact()
    tr = Guarded(T64, cuda_device)
    check(L_.mfhe_transpose(g64._h, _dev_ptr(u[0]), _dev_ptr(tr.view), T64 // g64.n2, _stream_handle()))
    rs = Guarded(g.phi * (g.L - 1) * g.n2, cuda_device)
    check(L_.mfhe_rescale_last(g._h, _dev_ptr(a), _dev_ptr(rs.view), _stream_handle()))
    torch.cuda.synchronize()
    assert tr.intact() and rs.intact()


@pytest.mark.parametrize("logn,limbs,batch", [(12, 3, 5), (14, 2, 3), (16, 2, 2), (17, 1, 1)])
def test_canaries_and_determinism_flat_family(cuda_device, logn, limbs, batch):
    import torch
    import matrix_fhe_gpu_b200 as m
    import matrix_fhe_gpu_b200.scheme as S
    N = 1 << logn
    mods = [int(v) for v in O.coeff_modulus_create(N, 60, limbs)]
    plan = m.NttPlan(logn, mods)
    x = dev(O.synth_polys(1, batch, mods, N), cuda_device).reshape(-1)
    buf = Guarded(x.numel(), cuda_device, x)
    plan.forward_(buf.view)
    f0 = buf.view.clone()
    for _ in range(11):
        buf.view.copy_(x); plan.forward_(buf.view)
        assert torch.equal(buf.view, f0), "forward NTT is not deterministic"
    plan.inverse_(buf.view)
    assert torch.equal(buf.view, x) and buf.intact()
    # fused product: a <- INTT(NTT(a) . NTT(b)), b left in NTT form
    b = Guarded(x.numel(), cuda_device, x.flip(0) % mods[0])
    a = Guarded(x.numel(), cuda_device, x)
    plan.mul_intt_(a.view, b.view)
    assert a.intact() and b.intact()
    # base conversion, plain and scatter forms
    if limbs >= 2:
        q, p = mods[:limbs - 1], mods[limbs - 1:]
        bc = S.BconvPlan(q, p)
        xin = x.view(batch, limbs, N)[:, : limbs - 1].contiguous().view(-1)
        out = Guarded(batch * len(p) * N, cuda_device)
        from matrix_fhe_gpu_b200 import check, _dev_ptr, _stream_handle
        check(S._L().mfhe_bconv(bc._h, _dev_ptr(xin), _dev_ptr(out.view), N, batch, _stream_handle()))
        o0 = out.view.clone()
        sc = Guarded(batch * len(p) * N, cuda_device)
        bc.partial_scatter(xin, 0, len(q), [sc.view.data_ptr() + 8 * j * N for j in range(len(p))], [len(p) * N] * len(p), N, batch)
        torch.cuda.synchronize()
        assert out.intact() and sc.intact() and torch.equal(sc.view, o0)


def test_canaries_and_determinism_round2_kernels(cuda_device):
    """The kernels added late in round 2: tcgen05 encrypted matrix product (persistent CTAs, TMEM reuse across problems, an
    mbarrier per column block), its fused decryption (shared-memory tiles read down their columns), the FP64-pipe NTT
    kernels, and the mod-down on the static scratch.  Outputs between canaries, 12 runs on the same input, same words."""
    import torch
    import matrix_fhe_gpu_b200 as m
    import matrix_fhe_gpu_b200.scheme as S
    from matrix_fhe_gpu_b200 import check, _dev_ptr, _stream_handle
    L_ = S._L()
    g = S.Context(64, 5, O.primes_congruent_one(2 * 64 * 15, 45, 1) + O.primes_congruent_one(2 * 64 * 15, 35, 2), 2.0 ** 30)
    T = g.T
    cts = [Guarded(2 * T, cuda_device, torch.cat([g.sample_uniform(), g.sample_gaussian()])) for _ in range(4)]
    prod = Guarded(8 * T, cuda_device)
    o_re, o_im = Guarded(T, cuda_device), Guarded(T, cuda_device)
    sk = g.generate_secret_key()
    first = None
    for _ in range(12):
        prod.view.fill_(-1); o_re.view.fill_(-1); o_im.view.fill_(-1)
        check(L_.mfhe_ct_matmul(g._h, *[_dev_ptr(c.view) for c in cts], _dev_ptr(prod.view), _stream_handle()))
        p = prod.view.clone()
        check(L_.mfhe_ct_matmul_decrypt(g._h, _dev_ptr(prod.view), _dev_ptr(sk), _dev_ptr(o_re.view), _dev_ptr(o_im.view), _stream_handle()))
        cur = (p, o_re.view.clone(), o_im.view.clone())
        first = first or cur
        assert all(torch.equal(x, y) for x, y in zip(first, cur)), "matrix product / decryption is not deterministic"
    assert all(c.intact() for c in cts) and prod.intact() and o_re.intact() and o_im.intact()
    q = torch.from_numpy(np.repeat(np.asarray(g.moduli, dtype=np.int64), g.n2)[None].repeat(g.phi, 0).reshape(-1)).to(cuda_device)
    assert int(first[0].min()) >= 0 and bool((first[0].view(8, T) < q).all())            # canonical residues everywhere
    # FP64-pipe flat NTT (moduli below 2^50), PB groups + remainder
    for logn, limbs, batch in ((12, 3, 5), (16, 2, 3), (17, 1, 1)):
        N = 1 << logn
        mods = [int(v) for v in O.coeff_modulus_create(N, 50, limbs)]
        plan = m.NttPlan(logn, mods)
        assert plan.arith == 1
        x = dev(O.synth_polys(2, batch, mods, N), cuda_device).reshape(-1)
        buf = Guarded(x.numel(), cuda_device, x)
        plan.forward_(buf.view)
        f0 = buf.view.clone()
        for _ in range(11):
            buf.view.copy_(x); plan.forward_(buf.view)
            assert torch.equal(buf.view, f0), "FP64-pipe forward NTT is not deterministic"
        plan.inverse_(buf.view)
        assert torch.equal(buf.view, x) and buf.intact()
    # mod-down on the library's static scratch
    import ctypes as C
    L_.mfhe_moddown_static.argtypes = [C.c_void_p] * 4 + [C.c_size_t, C.c_void_p]
    logn, lq, lp, batch = 12, 3, 2, 4
    N = 1 << logn
    primes = [int(v) for v in O.get_primes(2 * N, 55, lq + lp)]
    plan, conv = m.NttPlan(logn, primes), S.BconvPlan(primes[lq:], primes[:lq])
    y = Guarded(batch * (lq + lp) * N, cuda_device, dev(O.synth_polys(5, batch, primes, N), cuda_device).reshape(-1))
    out = Guarded(batch * lq * N, cuda_device)
    first = None
    for _ in range(12):
        out.view.fill_(-1)
        check(L_.mfhe_moddown_static(plan._h, conv._h, _dev_ptr(y.view), _dev_ptr(out.view), batch, _stream_handle()))
        first = out.view.clone() if first is None else first
        assert torch.equal(out.view, first), "static mod-down is not deterministic"
    assert y.intact() and out.intact()
