"""Flat NTT on the FP64 pipe (moduli below 2^50; csrc/modmath.cuh, ntt_flat.cu<AR = 1>) against the CPU oracle and against
the integer kernels, word for word: the two arithmetics are interchangeable, so every test here runs both and asserts
the same words.  Covers every N the flat path supports, PB groups + remainder, limb windows, chunking, the fused
product, inputs at the edges of the contract (0, q-1, lazy values up to 2q-1), and the full config-2 shape at 50 bits."""
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu


def to_dev(x, dev):
    import torch
    return torch.from_numpy(np.ascontiguousarray(x).view(np.int64)).to(dev)


def to_host(t):
    return t.cpu().numpy().view(np.uint64)


@pytest.fixture(scope="module")
def m():
    import matrix_fhe_gpu_b200 as m
    return m


def test_arith_selection(m, cuda_device, monkeypatch):
    n = 4096
    p50 = m.NttPlan(12, O.coeff_modulus_create(n, 50, 2))
    assert p50.arith == 1                                         # every modulus below 2^50: FP64 pipe by default
    p50.set_arith(0); assert p50.arith == 0
    p50.set_arith(1); assert p50.arith == 1
    with pytest.raises(m.MfheError):
        p50.set_arith(2)
    p51 = m.NttPlan(12, list(O.coeff_modulus_create(n, 50, 1)) + list(O.coeff_modulus_create(n, 51, 1)))
    assert p51.arith == 0                                         # one wider modulus: integer kernels
    with pytest.raises(m.MfheError):
        p51.set_arith(1)
    monkeypatch.setenv("MFHE_NTT_ARITH", "int")
    assert m.NttPlan(12, O.coeff_modulus_create(n, 50, 2)).arith == 0
    monkeypatch.delenv("MFHE_NTT_ARITH")
    assert m.NttPlan(6, [17592186435073]).arith == 1              # n <= 256: the flag drives the fused row kernel of the scheme
    assert m.NttPlan(6, [17592186435073], convention=m.GL_XN_MINUS_I).arith == 1   # all three conventions share the network
    assert m.NttPlan(6, [1152921504606830593 - 0], convention=m.PHANTOM_NEGACYCLIC).arith == 0   # a 60-bit modulus


@pytest.mark.parametrize("logn,bits,limbs,batch", [
    (12, 50, 1, 1), (12, 50, 3, 5), (13, 45, 2, 3), (14, 50, 4, 2), (15, 35, 3, 3), (16, 50, 4, 3), (17, 50, 3, 2),
    (16, 30, 2, 1), (17, 49, 2, 3),
])
def test_forward_inverse_both_arithmetics(m, cuda_device, logn, bits, limbs, batch):
    n = 1 << logn
    mods = O.coeff_modulus_create(n, bits, limbs)
    assert int(max(mods)) < 2 ** 50
    t = O.PhantomTables(logn, mods)
    plan = m.NttPlan(logn, mods)
    x = O.synth_polys(42 + logn, batch, mods, n)
    q = np.asarray(mods, dtype=np.uint64).reshape(1, -1, 1)
    x[0, :, :64] = 0
    x[0, :, 64:128] = np.broadcast_to(q - np.uint64(1), (1, limbs, 64))[0]
    ref = t.forward(x)
    for arith in (1, 0):
        plan.set_arith(arith)
        d = to_dev(x, cuda_device)
        plan.forward_(d)
        assert (to_host(d) == ref).all(), arith
        plan.inverse_(d)
        assert (to_host(d) == x).all(), arith
    # lazy inputs: anything below 2q is accepted by both directions
    plan.set_arith(1)
    lazy = x + q * (O.splitmix64(np.arange(x.size, dtype=np.uint64)).reshape(x.shape) & np.uint64(1))
    lazy[0, :, 128:192] = np.broadcast_to(2 * q - np.uint64(1), (1, limbs, 64))[0]
    xl = lazy % q
    d = to_dev(lazy, cuda_device)
    plan.forward_(d)
    assert (to_host(d) == t.forward(xl)).all()
    d = to_dev(t.forward(xl) + (lazy - xl), cuda_device)
    plan.inverse_(d)
    assert (to_host(d) == xl).all()


def test_constant_vectors_and_worst_case_magnitudes(m, cuda_device):
    logn, n = 16, 1 << 16
    mods = [int(v) for v in O.get_primes(n, 50, 2)]                # the largest 50-bit NTT primes
    t = O.PhantomTables(logn, mods)
    plan = m.NttPlan(logn, mods)
    assert plan.arith == 1
    q = np.asarray(mods, dtype=np.uint64).reshape(1, 2, 1)
    cases = {
        "zeros": np.zeros((1, 2, n), dtype=np.uint64),
        "ones": np.ones((1, 2, n), dtype=np.uint64),
        "qm1": np.broadcast_to(q - np.uint64(1), (1, 2, n)).copy(),
        "2qm1": np.broadcast_to(2 * q - np.uint64(1), (1, 2, n)).copy(),
        "alternating": np.where(np.arange(n) % 2 == 0, q - np.uint64(1), np.uint64(0)).astype(np.uint64).reshape(1, 2, n).copy(),
        "delta": np.zeros((1, 2, n), dtype=np.uint64),
    }
    cases["delta"][0, :, 1] = 1
    for name, x in cases.items():
        xr = x % q
        d = to_dev(x, cuda_device)
        plan.forward_(d)
        f = to_host(d)
        assert (f == t.forward(xr)).all(), name
        plan.inverse_(d)
        assert (to_host(d) == xr).all(), name
        assert int(f.max()) < int(q.max())


def test_windows_chunks_and_fused_product(m, cuda_device):
    logn, n, L = 13, 8192, 6
    mods = O.coeff_modulus_create(n, 50, L)
    t = O.PhantomTables(logn, mods)
    plan = m.NttPlan(logn, mods)
    assert plan.arith == 1
    x = O.synth_polys(7, 5, mods, n)
    ref = t.forward(x)
    for chunk in (0, 1, 3, 6, 12, 7, 1000):
        plan.set_chunk(chunk)
        d = to_dev(x, cuda_device)
        plan.forward_(d)
        assert (to_host(d) == ref).all(), chunk
        plan.inverse_(d)
        assert (to_host(d) == x).all(), chunk
    plan.set_chunk(0)
    xs = np.ascontiguousarray(x[:, 2:5, :])
    d = to_dev(xs, cuda_device)
    plan.forward_limbs_(d, 2, 3)
    assert (to_host(d) == ref[:, 2:5, :]).all()
    plan.inverse_limbs_(d, 2, 3)
    assert (to_host(d) == xs).all()
    # NTT . pointwise . INTT with the product inside the first inverse pass: both arithmetics, same words
    a, b = O.synth_polys(3, 3, mods, n), O.synth_polys(4, 3, mods, n)
    outs = []
    for arith in (0, 1):
        plan.set_arith(arith)
        fa, fb = to_dev(a, cuda_device), to_dev(b, cuda_device)
        plan.mul_intt_(fa, fb)
        outs.append(to_host(fa).copy())
        ua, ub = to_dev(a, cuda_device), to_dev(b, cuda_device)
        plan.forward_(ua); plan.forward_(ub); plan.mul(ua, ub, out=ua); plan.inverse_(ua)
        assert (to_host(ua) == outs[-1]).all()
    assert (outs[0] == outs[1]).all()
    ref0 = np.zeros(n, dtype=np.uint64)
    q0 = int(mods[0])
    O.lib().mfo_schoolbook_mul(a[0, 0].copy(), b[0, 0].copy(), ref0, n, q0, q0 - 1)
    assert (outs[1][0, 0] == ref0).all()


def test_config2_shape_at_50_bits(m, cuda_device):
    """SURVEY config 2's secondary run: N = 2^16, L = 32, batch 64, 50-bit primes.  Oracle on polys 0 and 63, the integer
    kernels on every word, round trip on every word."""
    import torch
    logn, n, L, batch = 16, 1 << 16, 32, 64
    mods = O.coeff_modulus_create(n, 50, L)
    t = O.PhantomTables(logn, mods)
    plan = m.NttPlan(logn, mods)
    assert plan.arith == 1
    x = O.synth_polys(2, batch, mods, n)
    d = to_dev(x, cuda_device)
    plan.forward_(d)
    f = to_host(d)
    for b in (0, 63):
        assert (f[b] == t.forward(x[b:b + 1])[0]).all()
    plan.set_arith(0)
    di = to_dev(x, cuda_device)
    plan.forward_(di)
    assert torch.equal(d, di)
    plan.set_arith(1)
    plan.inverse_(d)
    assert (to_host(d) == x).all()


@pytest.mark.parametrize("n,f,bits,L", [(64, 5, 45, 3), (64, 5, 50, 2), (16, 5, 35, 4), (32, 17, 49, 2)])
def test_row_kernel_both_arithmetics_and_oracle(cuda_device, n, f, bits, L):
    """Fused encrypt / decrypt row kernel (xrow_mul_s): NTT_X -> key product -> INTT_X on the FP64 pipe equals the
    integer kernel and the oracle on every ciphertext word, single and dual (pair) forms."""
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    mods = O.primes_congruent_one(2 * n * 3 * f, bits, L)
    assert max(mods) < 2 ** 50
    g = S.Context(n, f, mods, 2.0 ** 30)
    o = O.SchemeCtx(n, f, mods, 2.0 ** 30)
    assert g.row_arith == 1
    rng = np.random.default_rng(n + bits)
    q = np.repeat(np.asarray(mods, dtype=np.uint64), n * n)[None].repeat(g.phi, 0).reshape(-1)
    m_re = rng.integers(0, 2 ** 62, size=g.T, dtype=np.uint64) % q
    m_im = rng.integers(0, 2 ** 62, size=g.T, dtype=np.uint64) % q
    m_re[:64] = 0; m_im[:64] = (q - 1)[:64]
    sk = g.generate_secret_key()
    outs = {}
    for arith in (1, 0):
        g.set_row_arith(arith)
        cr, ci = g.encrypt_pair(to_dev(m_re, cuda_device), to_dev(m_im, cuda_device), sk)
        dr = g.decrypt_to_eval(cr, sk, S.POLY)
        # pair form (one launch for both ciphertexts, shared and distinct a) through decrypt_and_decode
        dec = g.decrypt_and_decode(cr, ci, sk).clone()
        ci2 = ci.clone(); ci2[g.T + 5] = (ci2[g.T + 5] + 1) % int(mods[0])          # one word of a differs: second pass
        dec2 = g.decrypt_and_decode(cr, ci2, sk).clone()
        outs[arith] = (cr.clone(), ci.clone(), dr.clone(), dec.view(torch.float64), dec2.view(torch.float64))
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a, b)
    a = to_host(g.sample_uniform()); e = to_host(g.sample_gaussian())
    ocr, oci = o.encrypt_pair_with(m_re, m_im, to_host(sk), a, e, e)
    assert (to_host(outs[1][0]) == ocr).all() and (to_host(outs[1][1]) == oci).all()
    assert (to_host(outs[1][2]) == o.decrypt_to_eval(ocr, to_host(sk))).all()
    g.set_row_arith(1)
