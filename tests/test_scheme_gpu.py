"""GPU parity tests of the matrix pipeline (family B): W-CRT, W-DFT, XY embedding, CRT, encode / decode,
samplers, keygen, encrypt / decrypt, ct ops, trace step, GL perm, base conversion -- all through the C ABI,
against the CPU oracle (oracle/scheme_oracle.c) on the same inputs.

Integer stages: bit-exact.  fp64 stages: max-abs error <= 2^-20 * max|reference| (BASELINE north_star),
plus the reference drivers' own thresholds (1e-3 / 1e-4) on the end-to-end patterns.
Small shapes compare every word; the native shape (n=64, p=771, L=11) compares the ciphertext words of one
full encrypt with the oracle and otherwise uses sampled columns and round-trip properties.
"""
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
REL = 2.0 ** -20

SHAPES = {
    # name: (n, f, moduli builder, delta)
    "tiny": (16, 5, lambda: O.primes_congruent_one(960, 35, 3), 2.0 ** 30),
    "mid": (16, 17, lambda: O.primes_congruent_one(64 * 51, 45, 1) + O.primes_congruent_one(64 * 51, 35, 3), 2.0 ** 30),
    "wide": (32, 5, lambda: O.primes_congruent_one(128 * 15, 55, 2) + O.primes_congruent_one(128 * 15, 40, 1), 2.0 ** 35),
    # the reference's p = 771 (phi = 512) and its first three moduli (45, 35, 35 bits) at n = 16: the smallest shape the
    # tcgen05 W-CRT takes (both digit counts, ragged last column block: 256 = 5*48 + 16 = 6*40 + 16), so every test
    # below also checks that kernel against the full oracle
    "p771": (16, 257, lambda: [17592186435073, 17182765057, 17184541441], 2.0 ** 35),
}


def dev(x, device):
    import torch
    x = np.ascontiguousarray(x)
    if x.dtype == np.uint64:
        x = x.view(np.int64)
    return torch.from_numpy(x).to(device)


def host(t, dtype=np.uint64):
    a = t.cpu().numpy()
    return a.view(dtype) if a.dtype == np.int64 and dtype == np.uint64 else a


@pytest.fixture(scope="module")
def S():
    import matrix_fhe_gpu_b200.scheme as s
    return s


_ctx_cache = {}


def ctxs(S, name):
    if name not in _ctx_cache:
        n, f, mods, delta = SHAPES[name]
        mods = mods()
        _ctx_cache[name] = (S.Context(n, f, mods, delta), O.SchemeCtx(n, f, mods, delta, gauss_jordan=(name == "tiny")))
    return _ctx_cache[name]


def rand_polyset(o, seed):
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 2 ** 63, size=(o.phi, o.L, o.n2), dtype=np.uint64)
    return (x % o.moduli.reshape(1, -1, 1)).reshape(-1)


@pytest.mark.parametrize("name", list(SHAPES))
def test_tables_match_oracle(S, cuda_device, name):
    g, o = ctxs(S, name)
    assert g.copy_device_moduli() == [int(q) for q in o.moduli]
    for l in range(o.L):
        assert g.eta(l) == o.eta(l)
        assert (g.wcrt_table(l) == o.wntt()[l]).all()
        assert (g.wcrt_table(l, inverse=True) == o.wntt(True)[l]).all()      # tiny: the reference's Gauss-Jordan
    t, r = g.crt_tables(), o.crt()
    assert g.big_words == o.big_words
    for k in ("Q", "Qhalf", "M", "inv"):
        assert (t[k] == r[k]).all()


@pytest.mark.parametrize("name", list(SHAPES))
def test_wcrt_matrix_bit_exact_all_layouts(S, cuda_device, name):
    g, o = ctxs(S, name)
    x = rand_polyset(o, 1)
    ref_eval_poly = o.wntt_forward_matrix(x)
    d = dev(x, cuda_device)
    y = g.wntt_forward_matrix(d)                               # MATRIX -> POLY, the reference's call
    assert (host(y) == ref_eval_poly).all()
    assert (host(g.wntt_inverse_matrix(y)) == x).all()         # POLY -> MATRIX
    assert (host(g.wntt_inverse_matrix(dev(ref_eval_poly, cuda_device))) == o.wntt_inverse_matrix(ref_eval_poly)).all()
    ym = g.wntt_forward_matrix(d, S.MATRIX, S.MATRIX)
    assert (host(ym) == o.poly_to_matrix(ref_eval_poly)).all()
    xp = dev(o.matrix_to_poly(x), cuda_device)
    assert (host(g.wntt_forward_matrix(xp, S.POLY, S.POLY)) == ref_eval_poly).all()
    assert (host(g.wntt_inverse_matrix(ym, S.MATRIX, S.POLY)) == o.matrix_to_poly(x)).all()
    # edge inputs: zeros, all q-1
    z = dev(np.zeros(o.T, dtype=np.uint64), cuda_device)
    assert not host(g.wntt_forward_matrix(z)).any()
    top = np.broadcast_to((o.moduli - np.uint64(1)).reshape(1, -1, 1), (o.phi, o.L, o.n2)).reshape(-1).copy()
    assert (host(g.wntt_forward_matrix(dev(top, cuda_device))) == o.wntt_forward_matrix(top)).all()


@pytest.mark.parametrize("name", ["tiny", "mid"])
def test_wcrt_unit_vector_known_answer(S, cuda_device, name):
    """test/test_custom_ntt_roundtrip.cu:169-254: unit vector at W-coeff r0 = 7 evaluates to eta^(exp[w] * 7)."""
    g, o = ctxs(S, name)
    r0 = 7
    x = np.zeros((o.phi, o.L, o.n2), dtype=np.uint64)
    x[r0] = 1
    y = host(g.wntt_forward_matrix(dev(x.reshape(-1), cuda_device), S.MATRIX, S.MATRIX)).reshape(o.phi, o.L, o.n2)
    wexp = o.wexp()
    for l in range(o.L):
        q = int(o.moduli[l])
        for w in range(o.phi):
            assert (y[w, l] == pow(o.eta(l), int(wexp[w]) * r0, q)).all()


@pytest.mark.parametrize("name", list(SHAPES))
def test_wcrt_vector_centered_layout(S, cuda_device, name):
    g, o = ctxs(S, name)
    rng = np.random.default_rng(2)
    v = (rng.integers(0, 2 ** 63, size=(o.phi, o.L, o.n), dtype=np.uint64) % o.moduli.reshape(1, -1, 1)).reshape(-1)
    assert (host(g.wntt_forward_vector(dev(v, cuda_device))) == o.wntt_forward_vector(v)).all()
    # test/test_wcrt_roundtrip.cu:34-41 pattern plus large magnitudes and INT64_MIN
    w, y, x = np.meshgrid(np.arange(o.phi), np.arange(o.n), np.arange(o.n), indexing="ij")
    c = (((w + x + y) % 17) - 8).astype(np.int64).reshape(-1)
    big = rng.integers(-2 ** 62, 2 ** 62, size=c.size, dtype=np.int64)
    big[0] = np.iinfo(np.int64).min
    for data in (c, big):
        f_ref = o.wntt_forward_centered(data)
        assert (host(g.wntt_forward_centered(dev(data, cuda_device)), np.int64) == f_ref).all()
        assert (host(g.wntt_inverse_centered(dev(data, cuda_device)), np.int64) == o.wntt_inverse_centered(data)).all()
    x = rand_polyset(o, 3)
    assert (host(g.layout_convert(dev(x, cuda_device), S.MATRIX, S.POLY)) == o.matrix_to_poly(x)).all()
    assert (host(g.layout_convert(dev(x, cuda_device), S.POLY, S.MATRIX)) == o.poly_to_matrix(x)).all()


def close(a, b, scale=None):
    a, b = np.asarray(a), np.asarray(b)
    scale = max(np.abs(b).max(), 1e-300) if scale is None else scale
    return np.abs(a - b).max() <= REL * scale


@pytest.mark.parametrize("name", list(SHAPES))
def test_fp64_transforms(S, cuda_device, name):
    g, o = ctxs(S, name)
    rng = np.random.default_rng(4)
    z = (rng.standard_normal(o.phi * o.n2) + 1j * rng.standard_normal(o.phi * o.n2)) * 1000.0
    d = dev(z, cuda_device)
    assert close(host(g.wdft_forward_complex(d)), o.wdft_forward_complex(z))
    assert close(host(g.wdft_inverse_complex(d)), o.w_idft(z))
    assert close(host(g.wdft_inverse_complex(g.wdft_forward_complex(d))), z)
    re = rng.integers(-2 ** 40, 2 ** 40, size=z.size, dtype=np.int64); im = rng.integers(-2 ** 40, 2 ** 40, size=z.size, dtype=np.int64)
    gr, gi = g.wdft_forward_centered_pair(dev(re, cuda_device), dev(im, cuda_device))
    orr, oi = o.wdft_forward_pair(re, im)
    assert close(host(gr), orr) and close(host(gi), oi)
    gr, gi = g.wdft_inverse_pair(dev(z.real.copy(), cuda_device), dev(z.imag.copy(), cuda_device))
    orr, oi = o.wdft_inverse_pair(z.real.copy(), z.imag.copy())
    assert close(host(gr), orr) and close(host(gi), oi)
    for lanes in (1, o.phi):
        m = z[: lanes * o.n2]
        assert close(host(g.idft2(dev(m, cuda_device), lanes)), o.xy_idft2(m, lanes))
        assert close(host(g.dft2(dev(m, cuda_device), lanes)), o.xy_dft2(m, lanes))
        assert close(host(g.dft2(g.idft2(dev(m, cuda_device), lanes), lanes)), m)


@pytest.mark.parametrize("name", list(SHAPES))
def test_quantize_and_crt_bit_exact(S, cuda_device, name):
    g, o = ctxs(S, name)
    rng = np.random.default_rng(5)
    # values whose scaled image is an exact integer + 0.25: no rounding-boundary ambiguity
    ints = rng.integers(-2 ** 50, 2 ** 50, size=(2, o.phi * o.n2)).astype(np.float64)
    z = (ints[0] + 0.25) / o.delta + 1j * (ints[1] - 0.25) / o.delta
    gr, gi = g.quantize(dev(z, cuda_device), o.phi)
    orr, oi = o.quantize(z, o.phi)
    assert (host(gr) == orr).all() and (host(gi) == oi).all()
    # CRT compose of arbitrary residues: (magnitude words, sign) and the fp64 image are exact
    res = (rng.integers(0, 2 ** 63, size=(o.L, o.n2), dtype=np.uint64) % o.moduli.reshape(-1, 1)).reshape(-1)
    small = np.array([int(v) % int(q) for q in o.moduli for v in rng.integers(-2 ** 60, 2 ** 60, size=o.n2)], dtype=np.uint64)
    for data in (res, small):
        for limbs in (o.L, 1):
            mag, neg = g.crt_compose_centerlift_big(dev(data, cuda_device), o.n2, limbs)
            omag, oneg = o.crt_compose_big(data, o.n2, limbs)
            assert (host(mag) == omag).all() and (host(neg, np.uint8) == oneg).all()
            assert (host(g.crt_compose_centerlift(dev(data, cuda_device), o.n2, limbs), np.int64) == o.crt_compose_i64(data, o.n2, limbs)).all()
    # lanes > 1 (the reference loops over lanes on the host)
    x = rand_polyset(o, 6)
    mag, neg = g.crt_compose_centerlift_big(dev(x, cuda_device), o.n2, o.L, lanes=o.phi)
    mag, neg = host(mag).reshape(o.phi, -1), host(neg, np.uint8).reshape(o.phi, -1)
    for ell in (0, o.phi - 1):
        omag, oneg = o.crt_compose_big(x.reshape(o.phi, -1)[ell])
        assert (mag[ell] == omag).all() and (neg[ell] == oneg).all()
    c = g.big_pair_to_complex(dev(host(g.crt_compose_centerlift_big(dev(x, cuda_device), o.n2, o.L, lanes=o.phi)[0]), cuda_device),
                              dev(neg.reshape(-1), cuda_device), dev(mag.reshape(-1), cuda_device), dev(neg.reshape(-1), cuda_device))
    oc = o.big_pair_to_complex(mag.reshape(-1), neg.reshape(-1), mag.reshape(-1), neg.reshape(-1))
    assert (host(c) == oc).all()
    re, im = x.reshape(o.phi, -1)[0], x.reshape(o.phi, -1)[1]
    assert (host(g.dequantize_exact(dev(re, cuda_device), dev(im, cuda_device))) == o.dequantize_exact(re, im)).all()
    # the two-limb fast path of crt_to_complex (crt.cu: crt_small) and its fall-back, mixed in one launch: small
    # values of both signs, the edges of (-q0 q1 / 2, q0 q1 / 2), and arbitrary residues -- the fp64 image must
    # equal the full composition's bit for bit
    if o.L >= 2:
        q01 = int(o.moduli[0]) * int(o.moduli[1])
        vals = [0, 1, -1, 2 ** 63, -2 ** 63, 2 ** 64 - 1, 2 ** 64, -(2 ** 64) - 5, q01 // 2, -(q01 // 2), q01 // 2 + 1,
                -(q01 // 2) - 1, q01 - 1, q01, int(o.moduli[0]), -int(o.moduli[1])]
        vals += [int(v) for v in rng.integers(-2 ** 62, 2 ** 62, size=o.n2 // 2)]
        Q = 1
        for q in o.moduli:
            Q *= int(q)
        vals = [v for v in vals if abs(v) < Q // 2][: o.n2]
        mixed = res.reshape(o.L, o.n2).copy()
        for i, v in enumerate(vals):
            for k in range(o.L):
                mixed[k, i] = v % int(o.moduli[k])
        re = mixed.reshape(-1)
        im = np.ascontiguousarray(mixed[:, ::-1]).reshape(-1)
        got = host(g.dequantize_exact(dev(re, cuda_device), dev(im, cuda_device)))
        assert (got == o.dequantize_exact(re, im)).all()
        for i, v in enumerate(vals[:16]):
            assert got[i].real == float(np.float64(v) / o.delta) or abs(v) >= 2 ** 53


def message(o, kind, seed=7):
    """Driver patterns: test_encode_decode_wcrt.cu:40-45, test_encode_encrypt_decrypt_decode_wcrt.cu:45-51,
    main.cu:62-69, plus uniform [-1,1]^2."""
    ell, i = np.meshgrid(np.arange(o.phi, dtype=np.float64), np.arange(o.n2, dtype=np.float64), indexing="ij")
    if kind == "big":
        v = ell * 10000.0 + i; return (v - 1j * v).reshape(-1)
    if kind == "e2e":
        v = ell + 0.001 * i; return (v - 1j * v).reshape(-1)
    if kind == "main":
        return ((ell + 1e-5 * i) + 1j * (ell - 1e-5 * i)).reshape(-1)
    rng = np.random.default_rng(seed)
    return rng.uniform(-1, 1, o.phi * o.n2) + 1j * rng.uniform(-1, 1, o.phi * o.n2)


@pytest.mark.parametrize("name", list(SHAPES))
def test_encode_decode(S, cuda_device, name):
    g, o = ctxs(S, name)
    for kind in ("uniform", "e2e"):
        msg = message(o, kind)
        scale = np.abs(msg).max()
        gre, gim = g.encode_to_wntt_eval(dev(msg, cuda_device))
        ore, oim = o.encode(msg)
        # H6: llround of an fp64 transform may differ by one unit between summation orders.  Compare the
        # W-coefficient integers (exact inverse W-CRT of both results, limb 0, centred): |diff| <= 1.
        q0 = int(o.moduli[0])
        for gg, oo in ((gre, ore), (gim, oim)):
            gc = host(g.wntt_inverse_matrix(gg, S.MATRIX, S.MATRIX)).reshape(o.phi, o.L, o.n2)[:, 0].astype(np.int64)
            oc = o.wntt_inverse_matrix(o.matrix_to_poly(oo)).reshape(o.phi, o.L, o.n2)[:, 0].astype(np.int64)
            d = (gc - oc) % q0
            d = np.minimum(d, q0 - d)
            assert d.max() <= 1 and (d != 0).mean() < 1e-3
        # decode of the GPU encoding == message; decode of the oracle's encoding agrees with the oracle's decode
        out = host(g.decode_eval_pair(gre, gim, S.MATRIX))
        assert close(out, msg, scale) and np.abs(out - msg).max() < 1e-3
        ep_r, ep_i = o.matrix_to_poly(ore), o.matrix_to_poly(oim)
        assert close(host(g.decode_eval_pair(dev(ep_r, cuda_device), dev(ep_i, cuda_device), S.POLY)), o.decode_eval_pair(ep_r, ep_i), scale)
    lane = message(o, "uniform")[: o.n2]
    lr, li = g.encode_lane(dev(lane, cuda_device))
    olr, oli = o.encode_lane(lane)
    qt = np.repeat(o.moduli, o.n2).astype(np.int64)
    dl = (host(lr).astype(np.int64) - olr.astype(np.int64)) % qt
    assert (np.minimum(dl, qt - dl) <= 1).all() and (dl != 0).mean() < 1e-3
    assert close(host(g.decode_lane_from_rns_eval(dev(olr, cuda_device), dev(oli, cuda_device))), o.decode_lane_from_rns_eval(olr, oli))


@pytest.mark.parametrize("name", list(SHAPES))
def test_samplers_and_keygen(S, cuda_device, name):
    g, o = ctxs(S, name)
    assert (host(g.sample_uniform()) == o.sample_uniform()).all()
    assert (host(g.sample_ternary()) == o.sample_ternary()).all()
    ge, oe = host(g.sample_gaussian()), o.sample_gaussian()
    # H5: fp64 log/sqrt/cos + llround may differ by one on rare rounding boundaries between libm and libdevice
    bad = ge != oe
    assert bad.mean() <= 1e-5
    if bad.any():
        q = np.tile(np.repeat(o.moduli, o.n2), o.phi)[bad]
        d = (ge[bad].astype(object) - oe[bad].astype(object)) % q.astype(object)
        assert all(int(v) in (1, int(m) - 1) for v, m in zip(d, q))
    assert (host(g.generate_secret_key()) == o.keygen()).all()


@pytest.mark.parametrize("name", list(SHAPES))
def test_encrypt_decrypt_bit_exact(S, cuda_device, name):
    g, o = ctxs(S, name)
    sk = o.keygen()
    m_re, m_im, a, e_re, e_im = (rand_polyset(o, s) for s in (10, 11, 12, 13, 14))
    d = lambda x: dev(x, cuda_device)
    cr, ci = g.encrypt_pair_with(d(m_re), d(m_im), d(sk), d(a), d(e_re), d(e_im))
    ocr, oci = o.encrypt_pair_with(m_re, m_im, sk, a, e_re, e_im)
    assert (host(cr) == ocr).all() and (host(ci) == oci).all()
    # sampler-fused encrypt_pair / encrypt: a is the exact uniform stream; e is whatever the GPU's gaussian
    # kernel draws (checked against libm separately), fed to the oracle's integer core
    ge = host(g.sample_gaussian())
    cr2, ci2 = g.encrypt_pair(d(m_re), d(m_im), d(sk))
    ocr2, oci2 = o.encrypt_pair_with(m_re, m_im, sk, o.sample_uniform(), ge, ge)
    assert (host(cr2) == ocr2).all() and (host(ci2) == oci2).all()
    assert (host(g.encrypt(d(m_re), d(sk))) == ocr2).all()
    # decrypt: both layouts, and m = b + a*s recovers m + e
    dp = host(g.decrypt_to_eval(cr, d(sk), S.POLY))
    assert (dp == o.decrypt_to_eval(ocr, sk)).all()
    assert (host(g.decrypt_to_eval(cr, d(sk), S.MATRIX)) == o.poly_to_matrix(dp)).all()
    e_eval = o.poly_to_matrix(o.wntt_forward_matrix(e_re)).reshape(o.phi, o.L, o.n2)
    want = (m_re.reshape(o.phi, o.L, o.n2).astype(object) + e_eval.astype(object)) % o.moduli.reshape(1, -1, 1).astype(object)
    assert (o.poly_to_matrix(dp).reshape(o.phi, o.L, o.n2) == want.astype(np.uint64)).all()
    # ciphertext ops, both modulus-indexing modes
    assert (host(g.add_ciphertexts(cr, ci, True)) == o.ct_add(ocr, oci)).all()
    for gg, oo in zip(g.multiply_ciphertexts_raw(cr, ci, True), o.ct_mul_raw(ocr, oci)):
        assert (host(gg) == oo).all()
    qm = np.tile(np.repeat(o.moduli, o.n2), 2 * o.phi).astype(object)
    assert (host(g.add_ciphertexts(cr, ci, False)) == ((ocr.astype(object) + oci.astype(object)) % qm).astype(np.uint64)).all()
    d0 = host(g.multiply_ciphertexts_raw(cr, ci, False)[0])
    assert (d0 == ((ocr[: o.T].astype(object) * oci[: o.T].astype(object)) % qm[: o.T]).astype(np.uint64)).all()


@pytest.mark.parametrize("name", list(SHAPES))
def test_full_pipeline_roundtrip(S, cuda_device, name):
    g, o = ctxs(S, name)
    sk = g.generate_secret_key()
    for kind, thr in (("e2e", 1e-3), ("main", 1e-4), ("uniform", 1e-4)):
        msg = message(o, kind)
        re, im = g.encode_to_wntt_eval(dev(msg, cuda_device))
        cr, ci = g.encrypt_pair(re, im, sk)
        out = host(g.decrypt_and_decode(cr, ci, sk))
        assert np.abs(out - msg).max() < thr          # the drivers' own pass rule
    # the GPU's decrypt_and_decode on the oracle's ciphertexts agrees with the oracle's
    ore, oim = o.encode(msg)
    ocr, oci = o.encrypt_pair(ore, oim, host(sk))
    ref = o.decrypt_and_decode(ocr, oci, host(sk))
    got = host(g.decrypt_and_decode(dev(ocr, cuda_device), dev(oci, cuda_device), sk))
    assert close(got, ref, max(np.abs(msg).max(), 1.0))
    # zero key, a = 0: test/test_encode_decode_wcrt.cu
    ct_r, ct_i = g.allocate_ciphertext(), g.allocate_ciphertext()
    big = message(o, "big")
    re, im = g.encode_to_wntt_eval(dev(big, cuda_device))
    ct_r[: o.T] = re; ct_i[: o.T] = im
    import torch
    out = host(g.decrypt_and_decode(ct_r, ct_i, torch.zeros_like(sk)))
    assert np.abs(out - big).max() < 1e-3 * max(1.0, np.abs(big).max() / 5.2e6)


@pytest.mark.parametrize("name", list(SHAPES))
def test_encode_encrypt_fused_bit_exact(S, cuda_device, name):
    """mfhe_encode_encrypt_pair == encode_to_wntt_eval followed by encrypt_pair, word for word (the W-CRT is linear,
    so message and error may share a transform), and against the oracle's integer core fed with the GPU's encoding."""
    g, o = ctxs(S, name)
    sk = g.generate_secret_key()
    for kind in ("e2e", "uniform"):
        m = dev(message(o, kind), cuda_device)
        re, im = g.encode_to_wntt_eval(m)
        cr, ci = g.encrypt_pair(re, im, sk)
        fr, fi = g.encode_encrypt_pair(m, sk)
        assert (host(fr) == host(cr)).all() and (host(fi) == host(ci)).all()
    ge = host(g.sample_gaussian())
    ocr, oci = o.encrypt_pair_with(host(re), host(im), host(sk), o.sample_uniform(), ge, ge)
    assert (host(fr) == ocr).all() and (host(fi) == oci).all()
    out = host(g.decrypt_and_decode(fr, fi, sk))
    assert np.abs(out - message(o, "uniform")).max() < 1e-4


@pytest.mark.parametrize("name", list(SHAPES))
def test_decrypt_pair_shared_and_distinct_a(S, cuda_device, name):
    """decrypt_and_decode runs both ciphertexts through one row-kernel launch and shares t = INTT(NTT(a) s) between
    them when their a parts agree (encrypt_pair stores the same a twice, HE.cu:1538-1545).  Equal a, entirely different
    a and a single differing word must all give exactly what two separate decrypt_to_eval calls give."""
    import torch
    g, o = ctxs(S, name)
    sk = g.generate_secret_key()
    m = dev(message(o, "uniform"), cuda_device)
    re, im = g.encode_to_wntt_eval(m)
    cr, ci = g.encrypt_pair(re, im, sk)
    e2 = g.sample_gaussian()
    _, ci2 = g.encrypt_pair_with(re, im, sk, dev(rand_polyset(o, 31), cuda_device), e2, e2)
    ci3 = ci.clone()
    ci3[o.T + 5] = 1 if int(ci3[o.T + 5]) != 1 else 2
    for x, y in ((cr, ci), (cr, ci2), (cr, ci3), (ci2, cr)):
        want = g.decode_eval_pair(g.decrypt_to_eval(x, sk, S.MATRIX), g.decrypt_to_eval(y, sk, S.MATRIX), S.MATRIX)
        got = g.decrypt_and_decode(x, y, sk)
        assert torch.equal(torch.view_as_real(got), torch.view_as_real(want))
    assert np.abs(host(g.decrypt_and_decode(cr, ci, sk)) - message(o, "uniform")).max() < 1e-4


@pytest.mark.parametrize("name", list(SHAPES))
def test_trace_and_glperm(S, cuda_device, name):
    g, o = ctxs(S, name)
    rng = np.random.default_rng(20)
    for batch in (1, 3):
        mats = [(rng.integers(0, 2 ** 63, size=(batch, o.L, o.n2), dtype=np.uint64) % o.moduli.reshape(1, -1, 1)).reshape(-1) for _ in range(4)]
        d = [dev(m, cuda_device) for m in mats]
        bpr, bpi = g.map_B_to_Bprime(d[2], d[3], batch)
        obr, obi = o.trace_map(mats[2], mats[3], batch)
        assert (host(bpr) == obr).all() and (host(bpi) == obi).all()
        cr, ci = g.trace_gemm(d[0], d[1], bpr, bpi, batch)
        ocr, oci = o.trace_gemm(mats[0], mats[1], obr, obi, batch)
        assert (host(cr) == ocr).all() and (host(ci) == oci).all()
        inv = [pow(int(o.delta) % int(q), int(q) - 2, int(q)) for q in o.moduli[:3]] + [0, 0]
        rr, ri = o.trace_rescale(ocr, oci, batch, inv[0], inv[1], inv[2])
        g.rescale_by_delta(cr, ci, inv[0], inv[1], inv[2], batch)
        assert (host(cr) == rr).all() and (host(ci) == ri).all()
    n = o.n
    p, ip = S.gl_perm_tables(n)
    op, oip = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
    O.lib().mfo_gl_perm(n, op, oip)
    assert (p == op).all() and (ip == oip).all()
    x = rng.integers(0, 2 ** 63, size=(5, o.L, n), dtype=np.uint64)
    y = host(S.apply_gl_perm(dev(x.reshape(-1), cuda_device), o.L, 5, n)).reshape(5, o.L, n)
    want = np.zeros_like(x); want[..., p] = x
    assert (y == want).all()
    back = host(S.apply_gl_perm(dev(y.reshape(-1), cuda_device), o.L, 5, n, inverse=True)).reshape(5, o.L, n)
    want2 = np.zeros_like(x); want2[..., ip] = y
    assert (back == want2).all()


@pytest.mark.parametrize("N,l_in,l_out,bits", [(1024, 8, 3, 58), (1024, 48, 4, 55), (1024, 9, 3, 58), (512, 2, 2, 60)])
def test_bconv_extreme_residues_one_word_barrett(S, cuda_device, N, l_in, l_out, bits):
    """The kernel reduces the accumulated 128-bit sum with a one-word Barrett whenever in_count * max q <= 2^61
    (8 x 58-bit is the boundary case, 48 x 55-bit the unsharded config-5 shape; 9 x 58-bit and 2 x 60-bit... the
    latter still qualifies, the former takes the 128-bit form).  Inputs at the extremes (all q_i - 1, all 0, random)
    make the sums as large as they can get; every output must equal the oracle's."""
    primes = [int(v) for v in O.get_primes(N, bits, l_in + l_out)]
    q, p = primes[:l_in], primes[l_in:]
    qa = np.array(q, dtype=np.uint64).reshape(-1, 1)
    rng = np.random.default_rng(31)
    x = np.stack([np.broadcast_to(qa - np.uint64(1), (l_in, N)), np.zeros((l_in, N), dtype=np.uint64),
                  rng.integers(0, 2 ** 63, size=(l_in, N), dtype=np.uint64) % qa]).astype(np.uint64)
    plan = S.BconvPlan(q, p)
    got = host(plan.convert(dev(np.ascontiguousarray(x).reshape(-1), cuda_device), N, 3)).reshape(3, l_out, N)
    for b in range(3):
        assert (got[b] == O.bconv(x[b], q, p)).all()


@pytest.mark.parametrize("N,l_in,l_out,bits,batch", [(4096, 6, 3, 55, 2), (8192, 12, 4, 60, 1), (1024, 3, 5, 36, 3)])
def test_bconv_bit_exact(S, cuda_device, N, l_in, l_out, bits, batch):
    primes = [int(v) for v in O.get_primes(N, bits, l_in + l_out)]
    q, p = primes[:l_in], primes[l_in:]
    rng = np.random.default_rng(30)
    x = rng.integers(0, 2 ** 63, size=(batch, l_in, N), dtype=np.uint64) % np.array(q, dtype=np.uint64).reshape(1, -1, 1)
    plan = S.BconvPlan(q, p)
    got = host(plan.convert(dev(x.reshape(-1), cuda_device), N, batch)).reshape(batch, l_out, N)
    for b in range(batch):
        assert (got[b] == O.bconv(x[b], q, p)).all()
    # limb-sharded form: partial sums over windows of the input base add up to the same residues
    import torch
    cuts = [0, l_in // 2, l_in]
    parts = []
    for s, e in zip(cuts[:-1], cuts[1:]):
        xl = np.ascontiguousarray(x[:, s:e]).reshape(-1)
        parts.append(plan.partial(dev(xl, cuda_device), s, e - s, N, batch))
    red = host(plan.reduce(torch.cat(parts), len(parts), 0, l_out, N, batch)).reshape(batch, l_out, N)
    assert (red == got).all()
