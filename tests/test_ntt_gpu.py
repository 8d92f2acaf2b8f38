"""GPU parity tests: the sm_100a NTT kernels (through the C ABI) against the CPU oracle.

Bit-exact everywhere (integer path).  Small shapes compare every word with the oracle; the
BASELINE full shape (N=2^16, L=32, batch 64) uses the oracle on polys 0 and 63 and
size-independent properties (round trip, linearity, NTT(a)*NTT(b) = NTT(a*b) on a subset).
"""
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu

REF_MODULI = [17592186435073, 17182765057, 17184541441, 17186120449, 17186515201, 17186909953,
              17188883713, 17190462721, 17190857473, 17191844353, 17192831233]


def to_dev(x, dev):
    import torch
    return torch.from_numpy(np.ascontiguousarray(x).view(np.int64)).to(dev)


def to_host(t):
    return t.cpu().numpy().view(np.uint64)


@pytest.fixture(scope="module")
def m():
    import matrix_fhe_gpu_b200 as m
    return m


@pytest.mark.parametrize("logn,bits,limbs,batch", [
    (12, 60, 1, 1), (12, 60, 3, 5), (13, 50, 2, 3), (14, 55, 4, 2), (15, 60, 3, 2),
    (16, 60, 4, 3), (16, 50, 2, 2), (17, 55, 3, 2),
])
def test_flat_forward_inverse_bit_exact(m, cuda_device, logn, bits, limbs, batch):
    n = 1 << logn
    mods = O.coeff_modulus_create(n, bits, limbs)
    t = O.PhantomTables(logn, mods)
    plan = m.NttPlan(logn, mods)
    for l in range(limbs):                                  # plan tables == oracle (== Phantom) tables
        assert plan.root(l) == t.psi(l)
        w, ws = plan.table(l)
        assert (w == t.tw[l]).all() and (ws == t.tws[l]).all()
    x = O.synth_polys(42, batch, mods, n)
    d = to_dev(x, cuda_device)
    plan.forward_(d)
    f = to_host(d)
    assert (f == t.forward(x)).all()
    plan.inverse_(d)
    assert (to_host(d) == x).all()
    # inverse alone on oracle output
    d2 = to_dev(t.forward(x), cuda_device)
    plan.inverse_(d2)
    assert (to_host(d2) == x).all()


def test_flat_edge_inputs(m, cuda_device):
    logn, n = 12, 4096
    mods = O.coeff_modulus_create(n, 60, 2)
    t = O.PhantomTables(logn, mods)
    plan = m.NttPlan(logn, mods)
    q = mods.reshape(1, 2, 1)
    cases = {
        "zeros": np.zeros((1, 2, n), dtype=np.uint64),
        "ones": np.ones((1, 2, n), dtype=np.uint64),        # ph/test/ntt_test.cu constant vectors
        "twos": np.full((1, 2, n), 2, dtype=np.uint64),
        "qm1": np.broadcast_to(q - np.uint64(1), (1, 2, n)).copy(),
        "delta": np.zeros((1, 2, n), dtype=np.uint64),
    }
    cases["delta"][0, :, 1] = 1
    for name, x in cases.items():
        d = to_dev(x, cuda_device)
        plan.forward_(d)
        assert (to_host(d) == t.forward(x)).all(), name
        plan.inverse_(d)
        assert (to_host(d) == x).all(), name
    # empty batch is a no-op
    import torch
    e = torch.empty((0, 2, n), dtype=torch.int64, device=cuda_device)
    plan.forward_(e)
    # wrong shape is rejected
    with pytest.raises(m.MfheError):
        plan.forward_(torch.zeros(n + 1, dtype=torch.int64, device=cuda_device))
    with pytest.raises(m.MfheError):
        plan.forward_(torch.zeros(2 * n, dtype=torch.int64))          # CPU tensor: no fallback


def test_polymul_n4096_vs_schoolbook(m, cuda_device):
    # BASELINE config 1a on the GPU path: NTT, pointwise product, INTT == schoolbook mod X^N+1
    q, logn, n = 1152921504606830593, 12, 4096
    plan = m.NttPlan(logn, [q])
    a = (np.arange(1, n + 1, dtype=np.uint64) % np.uint64(q)).reshape(1, 1, n)
    b = (np.arange(3, n + 3, dtype=np.uint64) % np.uint64(q)).reshape(1, 1, n)
    for (a, b) in ((a, b), (O.synth_polys(1, 1, [q], n), O.synth_polys(2, 1, [q], n))):
        da, db = to_dev(a, cuda_device), to_dev(b, cuda_device)
        plan.forward_(da); plan.forward_(db)
        plan.mul(da, db, out=da)
        plan.inverse_(da)
        ref = np.zeros(n, dtype=np.uint64)
        O.lib().mfo_schoolbook_mul(a.reshape(-1).copy(), b.reshape(-1).copy(), ref, n, q, q - 1)
        assert (to_host(da).reshape(-1) == ref).all()


def test_mul_intt_fused(m, cuda_device):
    """mfhe_ntt_mul_intt (product applied inside the first inverse pass) == schoolbook at N = 4096 (BASELINE config 1a),
    == the unfused NTT / pointwise / INTT sequence word for word at N = 2^16 with an odd batch (PB groups + remainder),
    operand reuse through b_is_ntt, and the small-n composition."""
    q, logn, n = 1152921504606830593, 12, 4096
    plan = m.NttPlan(logn, [q])
    a, b = O.synth_polys(1, 1, [q], n), O.synth_polys(2, 1, [q], n)
    da, db = to_dev(a, cuda_device), to_dev(b, cuda_device)
    plan.mul_intt_(da, db)
    ref = np.zeros(n, dtype=np.uint64)
    O.lib().mfo_schoolbook_mul(a.reshape(-1).copy(), b.reshape(-1).copy(), ref, n, q, q - 1)
    assert (to_host(da).reshape(-1) == ref).all()
    dbn = to_dev(b, cuda_device); plan.forward_(dbn)
    assert (to_host(db) == to_host(dbn)).all()                      # b is left in NTT form
    for logn, L, batch in ((16, 3, 3), (13, 2, 5)):
        n = 1 << logn
        mods = O.coeff_modulus_create(n, 60 if logn == 16 else 50, L)
        plan = m.NttPlan(logn, mods)
        a, b = O.synth_polys(3, batch, mods, n), O.synth_polys(4, batch, mods, n)
        ua, ub = to_dev(a, cuda_device), to_dev(b, cuda_device)
        plan.forward_(ua); plan.forward_(ub); plan.mul(ua, ub, out=ua); plan.inverse_(ua)
        fa, fb = to_dev(a, cuda_device), to_dev(b, cuda_device)
        plan.mul_intt_(fa, fb)
        assert (to_host(fa) == to_host(ua)).all() and (to_host(fb) == to_host(ub)).all()
        ga = to_dev(a, cuda_device)
        plan.mul_intt_(ga, fb, b_is_ntt=True)                        # reuse the transformed operand
        assert (to_host(ga) == to_host(ua)).all()
    q64 = 17592186435073                                             # n = 64, the reference's first modulus
    plan = m.NttPlan(6, [q64])
    a, b = O.synth_polys(5, 7, [q64], 64), O.synth_polys(6, 7, [q64], 64)
    fa, fb = to_dev(a, cuda_device), to_dev(b, cuda_device)
    plan.mul_intt_(fa, fb)
    got = to_host(fa).reshape(7, 64)
    for i in range(7):
        ref = np.zeros(64, dtype=np.uint64)
        O.lib().mfo_schoolbook_mul(a.reshape(7, 64)[i].copy(), b.reshape(7, 64)[i].copy(), ref, 64, q64, q64 - 1)
        assert (got[i] == ref).all()


def test_limb_window_and_chunking(m, cuda_device):
    logn, n, L = 13, 8192, 6
    mods = O.coeff_modulus_create(n, 55, L)
    t = O.PhantomTables(logn, mods)
    plan = m.NttPlan(logn, mods)
    x = O.synth_polys(7, 4, mods, n)
    ref = t.forward(x)
    for chunk in (0, 1, 2, 3, 6, 12, 7, 1000):
        plan.set_chunk(chunk)
        d = to_dev(x, cuda_device)
        plan.forward_(d)
        assert (to_host(d) == ref).all(), chunk
        plan.inverse_(d)
        assert (to_host(d) == x).all(), chunk
    plan.set_chunk(0)
    # window: limbs 2..4 of every poly
    xs = np.ascontiguousarray(x[:, 2:5, :])
    d = to_dev(xs, cuda_device)
    plan.forward_limbs_(d, 2, 3)
    assert (to_host(d) == ref[:, 2:5, :]).all()
    plan.inverse_limbs_(d, 2, 3)
    assert (to_host(d) == xs).all()


def test_host_buffer_entry_points(m, cuda_device):
    logn, n = 12, 4096
    mods = O.coeff_modulus_create(n, 60, 3)
    t = O.PhantomTables(logn, mods)
    plan = m.NttPlan(logn, mods)
    x = O.synth_polys(11, 2, mods, n)
    h = x.copy()
    plan.forward_host_(h)
    assert (h == t.forward(x)).all()
    plan.inverse_host_(h)
    assert (h == x).all()


@pytest.mark.parametrize("conv", [0, 1, 2])
@pytest.mark.parametrize("logn", [4, 5, 6, 7, 8])
def test_small_transforms_all_conventions(m, cuda_device, logn, conv):
    n = 1 << logn
    if logn <= 6:
        mods = np.array(REF_MODULI, dtype=np.uint64)
    else:
        # primes = 1 mod 4n so that all three conventions exist
        mods = O.get_primes(2 * n, 50, 3)
    L = len(mods)
    plan = m.NttPlan(logn, mods, conv)
    batch = 37
    x = O.synth_polys(9, batch, mods, n)
    lib = O.lib()
    if conv == 0:
        ref = O.PhantomTables(logn, mods).forward(x)
    else:
        ref = x.copy()
        f = lib.mfo_gl_forward if conv == 1 else lib.mfo_cyclic_forward
        for b in range(batch):
            for l in range(L):
                f(ref[b, l], n, int(mods[l]))
    assert plan.arith == 1                               # moduli below 2^50: the FP64-pipe networks are the default
    for arith in (1, 0):                                 # both arithmetics against the oracle, every word
        plan.set_arith(arith)
        d = to_dev(x, cuda_device)
        plan.forward_(d)
        assert (to_host(d) == ref).all(), arith
        plan.inverse_(d)
        assert (to_host(d) == x).all(), arith
    # extreme residues through the FP64-pipe networks
    plan.set_arith(1)
    q = np.asarray(mods, dtype=np.uint64).reshape(1, L, 1)
    for fill in (np.zeros_like(x), np.broadcast_to(q - np.uint64(1), x.shape).copy()):
        want = fill.copy()
        if conv == 0:
            want = O.PhantomTables(logn, mods).forward(fill)
        else:
            for b in range(batch):
                for l in range(L):
                    f(want[b, l], n, int(mods[l]))
        d = to_dev(fill, cuda_device)
        plan.forward_(d)
        assert (to_host(d) == want).all()
        plan.inverse_(d)
        assert (to_host(d) == fill).all()


def test_reference_driver_patterns_n64(m, cuda_device):
    # test/test_custom_ntt_roundtrip.cu: h_in = (b+l+x+1) mod q round trips at batch 1 and
    # phi*n = 32768 for the phantom path (:63-112) and (b+l+x+7) for gl (:115-166); then the
    # X^64 - i product of a_j=j+1, b_j=j+3 on limb 0 vs schoolbook (:256-319).
    n, L = 64, 11
    mods = np.array(REF_MODULI, dtype=np.uint64)
    for conv, off in ((0, 1), (1, 7)):
        plan = m.NttPlan(6, mods, conv)
        for batch in (1, 32768):
            b = np.arange(batch, dtype=np.uint64).reshape(-1, 1, 1)
            l = np.arange(L, dtype=np.uint64).reshape(1, -1, 1)
            xx = np.arange(n, dtype=np.uint64).reshape(1, 1, -1)
            x = ((b + l + xx + np.uint64(off)) % mods.reshape(1, -1, 1)).astype(np.uint64)
            d = to_dev(x, cuda_device)
            plan.forward_(d)
            if batch == 1:
                assert not (to_host(d) == x).all()
            plan.inverse_(d)
            assert (to_host(d) == x).all()                      # max_err == 0
    q = REF_MODULI[0]
    plan = m.NttPlan(6, [q], 1)
    a = np.arange(1, n + 1, dtype=np.uint64).reshape(1, 1, n)
    b = np.arange(3, n + 3, dtype=np.uint64).reshape(1, 1, n)
    da, db = to_dev(a, cuda_device), to_dev(b, cuda_device)
    plan.forward_(da); plan.forward_(db)
    plan.mul(da, db, out=da)
    plan.inverse_(da)
    lib = O.lib()
    ref = np.zeros(n, dtype=np.uint64)
    lib.mfo_schoolbook_mul(a.reshape(-1).copy(), b.reshape(-1).copy(), ref, n, q,
                           lib.mfo_powmod(lib.mfo_gl_get_psi(q, n), n, q))
    assert (to_host(da).reshape(-1) == ref).all()                # mul max_err == 0


def test_pointwise_ops(m, cuda_device):
    logn, n = 12, 4096
    mods = O.coeff_modulus_create(n, 60, 3)
    plan = m.NttPlan(logn, mods)
    a, b = O.synth_polys(1, 2, mods, n), O.synth_polys(2, 2, mods, n)
    q = mods.reshape(1, -1, 1)
    da, db = to_dev(a, cuda_device), to_dev(b, cuda_device)
    import torch
    out = torch.empty_like(da)
    plan.add(da, db, out=out)
    assert (to_host(out) == (a + b) % q).all()
    plan.sub(da, db, out=out)
    assert (to_host(out) == (a + (q - b)) % q).all()
    plan.mul(da, db, out=out)
    ref = np.array([(int(x) * int(y)) % int(qq) for x, y, qq in
                    zip(a.reshape(-1), b.reshape(-1), np.broadcast_to(q, a.shape).reshape(-1))], dtype=np.uint64)
    assert (to_host(out).reshape(-1) == ref).all()


def test_baseline_config2_full_shape(m, cuda_device):
    """N=2^16, L=32 (60-bit primes of CoeffModulus::Create), batch 64 = 1 GiB."""
    import torch
    logn, n, L, batch = 16, 65536, 32, 64
    mods = O.coeff_modulus_create(n, 60, L)
    plan = m.NttPlan(logn, mods)
    t = O.PhantomTables(logn, mods)
    # inputs generated on the host in slices (splitmix64(seed + flat idx) mod q), seed 42
    x = O.synth_polys(42, batch, mods, n)
    d = to_dev(x, cuda_device)
    plan.forward_(d)
    f0 = to_host(d[0]); f63 = to_host(d[63])
    assert (f0 == t.forward(x[0:1])[0]).all()                  # all 32 limbs of poly 0
    assert (f63 == t.forward(x[63:64])[0]).all()               # and of poly 63
    # canonical outputs
    dq = torch.from_numpy(mods.view(np.int64)).to(cuda_device).view(1, L, 1)
    assert bool(((d >= 0) & (d < dq)).all())                   # q < 2^60 so int64 compare is valid
    # linearity on the device: NTT(x) + NTT(y) == NTT(x + y)
    y = O.synth_polys(43, 4, mods, n)
    dy = to_dev(y, cuda_device)
    dxy = to_dev((x[:4] + y) % mods.reshape(1, L, 1), cuda_device)
    plan.forward_(dy); plan.forward_(dxy)
    s = torch.empty_like(dy)
    plan.add(d[:4].contiguous(), dy, out=s)
    assert bool((s == dxy).all())
    # chunked (L2-resident) launches give the same words
    plan.set_chunk(64)
    d3 = to_dev(x, cuda_device)
    plan.forward_(d3)
    assert bool((d3 == d).all())
    plan.inverse_(d3)
    assert (to_host(d3[5]) == x[5]).all()
    plan.set_chunk(0)
    plan.inverse_(d)
    assert bool((d == to_dev(x, cuda_device)).all())           # round trip of the whole GiB
