"""Pins the matrix-pipeline oracle (oracle/wcrt_oracle.c, scheme_oracle.c) against the reference's known answers.

Sources of truth:
  * constants derived from the reference's own algorithms in SURVEY.md 8(c): eta per limb (h_find_eta,
    HE.cu:119-133), k_wntt_exp == (a*257 + b*3) mod 771 (HE.cu:72-105), Q = prod q has 385 bits / 7 words
    (encoder.cu:23, HE.cu:27);
  * the literal Gauss-Jordan inversion of the reference (HE.cu:135-235) against the closed form the oracle
    normally uses, on one native limb and on a small ring;
  * the reference drivers' checks: W-CRT unit vector -> eta^(exp[w]*7) (test_custom_ntt_roundtrip.cu:169-254),
    int64 W-CRT pattern ((w+x+y) mod 17) - 8 (test_wcrt_roundtrip.cu:34-72), encode->decode patterns with the
    drivers' thresholds (test_encode_decode_wcrt.cu:40-45,115; test_encode_encrypt_decrypt_decode_wcrt.cu:45-51,109;
    main.cu:62-69,150);
  * definitional evaluation (Python integers) of every integer map.
"""
import numpy as np
import pytest

import oracle as O

import json
import os

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "scheme_constants.json")))   # make_golden_scheme.py


@pytest.fixture(scope="module")
def small():
    mods = O.primes_congruent_one(960, 35, 3)
    return O.SchemeCtx(16, 5, mods, 2.0 ** 30, gauss_jordan=True), O.SchemeCtx(16, 5, mods, 2.0 ** 30)


@pytest.fixture(scope="module")
def mid():
    mods = O.primes_congruent_one(64 * 51, 45, 1) + O.primes_congruent_one(64 * 51, 35, 2)
    return O.SchemeCtx(16, 17, mods, 2.0 ** 30)


def test_survey_constants_native():
    L = O._scheme_lib()
    assert [L.mfo_find_eta(q, 257) for q in O.REF_MODULI[:3]] == [10467697596351, 6077882864, 12895515671]
    Q = 1
    for q in O.REF_MODULI:
        Q *= q
    assert Q.bit_length() == 385 and (Q.bit_length() + 63) // 64 == 7


def test_wexp_is_the_reference_table():
    c = O.SchemeCtx(moduli=O.REF_MODULI[:1])
    e = [int(v) for v in c.wexp()]
    assert e == GOLD["k_wntt_exp"]                      # the literal table of HE.cu:72-105
    assert O.REF_MODULI == GOLD["RNS_MODULI"] and O.REF_DELTA == GOLD["SCALING_FACTOR"]
    assert (c.n, c.phi, c.p) == (GOLD["MATRIX_N"], GOLD["BATCH_SIZE"], GOLD["BATCH_PRIME_P"])
    assert e == [(a * 257 + b * 3) % 771 for a in (1, 2) for b in range(1, 257)]
    assert len(set(e)) == 512 and all(np.gcd(v, 771) == 1 for v in e)


def test_gauss_jordan_equals_closed_form_native_limb():
    """One native limb: the reference's O(phi^3) Gauss-Jordan inverse == the closed form (few seconds)."""
    q = O.REF_MODULI[1]
    a = O.SchemeCtx(moduli=[q], gauss_jordan=True)
    b = O.SchemeCtx(moduli=[q])
    assert (a.wntt(True) == b.wntt(True)).all() and (a.wntt() == b.wntt()).all()
    V, Vi = a.wntt()[0], a.wntt(True)[0]                   # Vi[w][r] = V^-1[r][w]
    # spot-check V^-1 V = I on a few rows with Python integers
    for r in (0, 1, 255, 511):
        row = [sum(int(Vi[w][r]) * int(V[w][c]) for w in range(512)) % q for c in (0, 1, r, 300)]
        assert row == [1 if c == r else 0 for c in (0, 1, r, 300)]
    assert np.abs(a.cx_table("wdft_inv") - b.cx_table("wdft_inv")).max() < 1e-12


def test_small_ring_tables(small):
    gj, cf = small
    assert (gj.wntt(True) == cf.wntt(True)).all()
    assert np.abs(gj.cx_table("wdft_inv") - cf.cx_table("wdft_inv")).max() < 1e-13
    V = cf.cx_table("V"); Vi = cf.cx_table("Vinv")
    assert np.abs(V @ Vi - np.eye(cf.n)).max() < 1e-12
    # V[j][k] = zeta^(5^j k), zeta = exp(2 pi i / 4n)  (encoder.cu:425-444)
    j, k = 3, 5
    assert abs(V[j, k] - np.exp(2j * np.pi * (pow(5, j, 4 * cf.n) * k) / (4 * cf.n))) < 1e-12
    crt = cf.crt()
    Q = 1
    for q in cf.moduli:
        Q *= int(q)
    assert sum(int(w) << (64 * i) for i, w in enumerate(crt["Q"])) == Q
    for l, q in enumerate(cf.moduli):
        M = sum(int(w) << (64 * i) for i, w in enumerate(crt["M"][l]))
        assert M == Q // int(q) and (M * int(crt["inv"][l])) % int(q) == 1


def test_wcrt_definitional_and_unit_vector(small, mid):
    for c in (small[1], mid):
        rng = np.random.default_rng(0)
        x = (rng.integers(0, 2 ** 63, size=(c.phi, c.L, c.n2), dtype=np.uint64) % c.moduli.reshape(1, -1, 1))
        y = c.wntt_forward_matrix(x).reshape(c.phi, c.n, c.L, c.n)        # [(w, y)][l][x]
        wexp = c.wexp()
        for (w, l, yy, xx) in ((0, 0, 0, 0), (c.phi - 1, c.L - 1, c.n - 1, 3), (5, 1, 2, 7)):
            q, eta = int(c.moduli[l]), c.eta(l)
            want = sum(int(x[r, l, yy * c.n + xx]) * pow(eta, int(wexp[w]) * r, q) for r in range(c.phi)) % q
            assert int(y[w, yy, l, xx]) == want
        assert (c.wntt_inverse_matrix(y) == x.reshape(-1)).all()
        # unit vector at r0 = 7 (test_custom_ntt_roundtrip.cu:169-254)
        u = np.zeros((c.phi, c.L, c.n2), dtype=np.uint64); u[7] = 1
        ev = c.poly_to_matrix(c.wntt_forward_matrix(u)).reshape(c.phi, c.L, c.n2)
        for w in range(c.phi):
            assert (ev[w, 0] == pow(c.eta(0), int(wexp[w]) * 7, int(c.moduli[0]))).all()
        v = x[:, :, : c.n].copy()
        got = c.wntt_forward_vector(v).reshape(c.phi, c.L, c.n)
        full = np.zeros_like(x); full[:, :, : c.n] = v
        assert (got == c.poly_to_matrix(c.wntt_forward_matrix(full)).reshape(c.phi, c.L, c.n2)[:, :, : c.n]).all()


def test_wcrt_centered_semantics(small):
    c = small[1]
    w, y, x = np.meshgrid(np.arange(c.phi), np.arange(c.n), np.arange(c.n), indexing="ij")
    v = (((w + x + y) % 17) - 8).astype(np.int64).reshape(-1)            # test_wcrt_roundtrip.cu:40
    inv = c.wntt_inverse_centered(v)
    # limb-0-only inverse of small evals is a centred residue mod q0
    q0 = int(c.moduli[0])
    assert np.abs(inv).max() <= q0 // 2
    fwd = c.wntt_forward_centered(v)
    # forward of a constant-in-w polynomial: evaluations of c(W) = sum_r v_r W^r differ per limb, so the CRT
    # lift is a ~Q-sized integer and saturates (HE.cu:902-915) unless the polynomial is constant in W
    const = np.tile(np.arange(-8, 8, dtype=np.int64), c.phi * c.n2 // 16)
    z = np.zeros((c.phi, c.n2), dtype=np.int64); z[0] = const[: c.n2]     # degree-0 polynomial in W
    assert (c.wntt_forward_centered(z.reshape(-1)).reshape(c.phi, c.n2) == z[0]).all()
    assert ((fwd == np.iinfo(np.int64).max) | (fwd == np.iinfo(np.int64).min) | (np.abs(fwd) < 2 ** 62)).all()


def test_crt_compose_matches_python(small):
    c = small[1]
    rng = np.random.default_rng(1)
    vals = [int(v) for v in rng.integers(-2 ** 62, 2 ** 62, size=c.n2)]
    res = np.array([[v % int(q) for v in vals] for q in c.moduli], dtype=np.uint64)
    mag, neg = c.crt_compose_big(res)
    mag = mag.reshape(c.n2, c.big_words)
    for i, v in enumerate(vals):
        assert sum(int(w) << (64 * k) for k, w in enumerate(mag[i])) == abs(v) and bool(neg[i]) == (v < 0)
    assert (c.crt_compose_i64(res) == np.array(vals, dtype=np.int64)).all()
    dq = c.dequantize_exact(res, res)
    assert (dq.real == np.array(vals, dtype=np.float64) / c.delta).all()


def _msg(c, kind):
    ell, i = np.meshgrid(np.arange(c.phi, dtype=np.float64), np.arange(c.n2, dtype=np.float64), indexing="ij")
    if kind == "big":
        v = ell * 10000.0 + i; return (v - 1j * v).reshape(-1)
    if kind == "e2e":
        v = ell + 0.001 * i; return (v - 1j * v).reshape(-1)
    return ((ell + 1e-5 * i) + 1j * (ell - 1e-5 * i)).reshape(-1)


def test_encode_decode_driver_patterns(small, mid):
    for c in (small[1], mid):
        for kind, thr in (("big", 1e-3), ("e2e", 1e-3), ("main", 1e-4)):
            m = _msg(c, kind)
            re, im = c.encode(m)
            out = c.decode_eval_pair(c.matrix_to_poly(re), c.matrix_to_poly(im))
            assert np.abs(out - m).max() < thr


def test_encrypt_decrypt_pipeline(small, mid):
    for c in (small[1], mid):
        sk = c.keygen()
        t = c.sample_ternary().reshape(c.phi, c.L, c.n)
        assert set(np.unique(t[:, 0])) <= {0, 1, int(c.moduli[0]) - 1}
        for kind, thr in (("e2e", 1e-3), ("main", 1e-4)):
            m = _msg(c, kind)
            re, im = c.encode(m)
            cr, ci = c.encrypt_pair(re, im, sk)
            assert (cr[c.T:] == ci[c.T:]).all()                          # shared a (HE.cu:1533-1534)
            assert np.abs(c.decrypt_and_decode(cr, ci, sk) - m).max() < thr
        # decrypt recovers m + e exactly; a is the W-CRT of the uniform stream
        dp = c.poly_to_matrix(c.decrypt_to_eval(cr, sk))
        e_eval = c.poly_to_matrix(c.wntt_forward_matrix(c.sample_gaussian()))
        q = np.tile(np.repeat(c.moduli, c.n2), c.phi)
        assert (dp == (re + e_eval) % q).all()
        assert (cr[c.T:] == c.poly_to_matrix(c.wntt_forward_matrix(c.sample_uniform()))).all()
        assert (c.encrypt(re, sk) == cr).all()
        # gaussian stream: small, centred, same integer in every limb
        g = c.sample_gaussian().reshape(c.phi, c.L, c.n2)
        cent = np.where(g[:, 0] > c.moduli[0] // np.uint64(2), g[:, 0].astype(np.int64) - np.int64(c.moduli[0]), g[:, 0].astype(np.int64))
        assert np.abs(cent).max() < 30 and 2.5 < cent.std() < 3.9
        for l in range(1, c.L):
            cl = np.where(g[:, l] > c.moduli[l] // np.uint64(2), g[:, l].astype(np.int64) - np.int64(c.moduli[l]), g[:, l].astype(np.int64))
            assert (cl == cent).all()


def test_xy_ntt_polymul_is_negacyclic(small):
    c = small[1]
    rng = np.random.default_rng(3)
    n, q = c.n, int(c.moduli[0])
    a = rng.integers(0, q, size=n, dtype=np.uint64); b = rng.integers(0, q, size=n, dtype=np.uint64)
    pad = lambda v: np.concatenate([v.reshape(1, n), np.zeros((c.L - 1, n), dtype=np.uint64)]).reshape(-1)
    fa, fb = c.xy_ntt_forward(pad(a), 1)[:n], c.xy_ntt_forward(pad(b), 1)[:n]
    prod = np.array([int(x) * int(y) % q for x, y in zip(fa, fb)], dtype=np.uint64)
    got = c.xy_ntt_backward(pad(prod), 1)[:n]
    want = np.zeros(n, dtype=np.uint64)
    O.lib().mfo_schoolbook_mul(a, b, want, n, q, q - 1)
    assert (got == want).all()


def test_trace_ops_definitional(small):
    c = small[1]
    rng = np.random.default_rng(4)
    n, L = c.n, c.L
    mk = lambda: (rng.integers(0, 2 ** 63, size=(L, n, n), dtype=np.uint64) % c.moduli.reshape(-1, 1, 1))
    Ar, Ai, Br, Bi = mk(), mk(), mk(), mk()
    Bpr, Bpi = c.trace_map(Br.reshape(-1), Bi.reshape(-1), 1)
    Bpr, Bpi = Bpr.reshape(L, n, n), Bpi.reshape(L, n, n)
    for l in range(L):
        q = int(c.moduli[l])
        assert (Bpr[l, 0] == Br[l, 0]).all() and (Bpi[l, 0] == (q - Bi[l, 0].astype(object)) % q).all()
        assert (Bpr[l, n - 1] == (q - Bi[l, 1].astype(object)) % q).all() and (Bpi[l, n - 1] == (q - Br[l, 1].astype(object)) % q).all()
    Cr, Ci = c.trace_gemm(Ar.reshape(-1), Ai.reshape(-1), Bpr.reshape(-1), Bpi.reshape(-1), 1)
    l, row, col = 1, 2, 5
    q = int(c.moduli[l])
    sr = sum(int(Ar[l, row, t]) * int(Bpr[l, col, t]) - int(Ai[l, row, t]) * int(Bpi[l, col, t]) for t in range(n))
    si = sum(int(Ar[l, row, t]) * int(Bpi[l, col, t]) + int(Ai[l, row, t]) * int(Bpr[l, col, t]) for t in range(n))
    assert int(Cr.reshape(L, n, n)[l, row, col]) == (sr * n) % q and int(Ci.reshape(L, n, n)[l, row, col]) == (si * n) % q


def test_bconv_definitional():
    N, l_in, l_out = 64, 5, 3
    pr = [int(v) for v in O.get_primes(4096, 50, l_in + l_out)]
    q, p = pr[:l_in], pr[l_in:]
    rng = np.random.default_rng(5)
    x = rng.integers(0, 2 ** 49, size=(l_in, N), dtype=np.uint64) % np.array(q, dtype=np.uint64).reshape(-1, 1)
    out = O.bconv(x, q, p)
    Q = 1
    for v in q:
        Q *= v
    for d in (0, 17, 63):
        y = [(int(x[i, d]) * pow(Q // q[i], -1, q[i])) % q[i] for i in range(l_in)]
        for j in range(l_out):
            assert int(out[j, d]) == sum(y[i] * ((Q // q[i]) % p[j]) for i in range(l_in)) % p[j]


# ---- encrypted matrix product: the oracle's restatement against plaintext linear algebra ----
def _matmul_ctx():
    # n = 16, p = 15 (phi = 8), five ~45-bit limbs: Q ~ 2^225 holds delta^2 * n * phi * |c|^2 with room to spare
    mods = O.primes_congruent_one(960, 45, 5)
    return O.SchemeCtx(16, 5, mods, 2.0 ** 30)


def test_conj_lane_is_an_involution_matching_the_exponents():
    for o in (_matmul_ctx(), O.SchemeCtx(moduli=O.REF_MODULI[:1])):
        cl, e = o.conj_lane(), [int(v) for v in o.wexp()]
        assert (cl[cl] == np.arange(o.phi)).all()
        assert all((e[i] + e[int(cl[i])]) % o.p == 0 for i in range(o.phi))


def test_trace_map_x_is_the_reference_map_on_transposed_matrices():
    """mfo_trace_map_x twists the fast index; on transposed matrices it is exactly trace.cu:30-62 (which twists the slow one)."""
    o = _matmul_ctx()
    rng = np.random.default_rng(2)
    batch = 3
    q = np.asarray(o.moduli, dtype=np.uint64).reshape(1, -1, 1, 1)
    Br = rng.integers(0, 2 ** 62, size=(batch, o.L, o.n, o.n), dtype=np.uint64) % q
    Bi = rng.integers(0, 2 ** 62, size=(batch, o.L, o.n, o.n), dtype=np.uint64) % q
    xr, xi = o.trace_map_x(Br.reshape(-1), Bi.reshape(-1), batch)
    tr = lambda a: np.ascontiguousarray(a.reshape(batch, o.L, o.n, o.n).transpose(0, 1, 3, 2)).reshape(-1)
    rr, ri = o.trace_map(tr(Br), tr(Bi), batch)
    assert (tr(rr) == xr).all() and (tr(ri) == xi).all()
    # and trace_gemm_ty == trace_gemm on transposed operands
    Ar = rng.integers(0, 2 ** 62, size=(batch, o.L, o.n, o.n), dtype=np.uint64) % q
    Ai = rng.integers(0, 2 ** 62, size=(batch, o.L, o.n, o.n), dtype=np.uint64) % q
    dr, di = o.trace_gemm_ty(Ar.reshape(-1), Ai.reshape(-1), Br.reshape(-1), Bi.reshape(-1), batch)
    cr, ci = o.trace_gemm(tr(Ar), tr(Ai), tr(Br), tr(Bi), batch)
    assert (dr == cr).all() and (di == ci).all()


def test_encrypted_matmul_decrypts_to_the_plaintext_product():
    """encode -> (V: conjugate-lane permutation, X twist) -> encrypt -> 4 component GEMMs -> 4-term decrypt -> decode
    == V^H U per lane (the factor n of the trace GEMM is part of the identity: sum_a xi_a^(r-r') = n [r = r']).
    Tolerance: the product of two encodings carries delta^2; each factor's quantisation error is <= 1/2 and the gaussian
    errors ~ 3.2 per W-coefficient, so the relative error is ~ (phi n)^(1/2) * 4 / (delta * |m|) ~ 2^-24 at delta = 2^30:
    asserted against the north-star bound 2^-20 * max|entry|."""
    o = _matmul_ctx()
    rng = np.random.default_rng(7)
    U = (rng.standard_normal((o.phi, o.n, o.n)) + 1j * rng.standard_normal((o.phi, o.n, o.n))) * 3
    V = (rng.standard_normal((o.phi, o.n, o.n)) + 1j * rng.standard_normal((o.phi, o.n, o.n))) * 3
    cl = o.conj_lane()
    ur, ui = o.encode(U.reshape(-1))
    vr, vi = o.encode(V[cl].reshape(-1))
    vpr, vpi = o.trace_map_x(vr, vi, o.phi)
    sk = o.keygen()
    cur, cui = o.encrypt_pair(ur, ui, sk)
    cvr, cvi = o.encrypt_pair(vpr, vpi, sk)
    prod = o.ct_matmul(cur, cui, cvr, cvi)
    pr, pi = o.ct_matmul_decrypt(prod, sk)
    out = o.decode_eval_pair(o.matrix_to_poly(pr), o.matrix_to_poly(pi)).reshape(o.phi, o.n, o.n) / o.delta
    want = np.einsum("lab,lac->lbc", V.conj(), U)          # V^H U per lane
    err = np.abs(out - want).max()
    print(f"encrypted matmul (oracle, n=16 phi=8): max_err {err:.3e} on entries up to {np.abs(want).max():.1f}")
    assert err <= 2.0 ** -20 * np.abs(want).max()
    # the plaintext side of the same identity, no encryption: n * U^T V' decoded
    dr, di = o.trace_gemm_ty(ur, ui, vpr, vpi, o.phi)
    tr = lambda a: np.ascontiguousarray(a.reshape(o.phi, o.L, o.n, o.n).transpose(0, 1, 3, 2)).reshape(-1)
    plain = o.decode_eval_pair(o.matrix_to_poly(tr(dr)), o.matrix_to_poly(tr(di))).reshape(o.phi, o.n, o.n) / o.delta
    assert np.abs(plain - want).max() <= 2.0 ** -24 * np.abs(want).max()


# ---- RNS rescale of the matrix scheme ----
def test_rescale_last_is_exact_division_on_the_w_coefficients():
    """X' = (X - [X]_{q_last}) / q_last for every integer W-coefficient, checked with Python integers through the
    CRT of the remaining limbs; X spans the whole range (-Q/2, Q/2)."""
    mods = O.primes_congruent_one(960, 45, 1) + O.primes_congruent_one(960, 35, 3)
    o = O.SchemeCtx(16, 5, mods, 2.0 ** 30)
    o1 = O.SchemeCtx(16, 5, mods[:-1], 2.0 ** 30)
    Q = 1
    for q in mods:
        Q *= q
    ql, Q1 = mods[-1], Q // mods[-1]
    rng = np.random.default_rng(9)
    X = [int(rng.integers(0, 2 ** 62)) * int(rng.integers(0, 2 ** 62)) * int(rng.integers(0, 2 ** 40)) % Q - Q // 2
         for _ in range(o.phi * o.n2)]
    X[0], X[1], X[2] = 0, ql, -ql // 2                             # edge cases
    coeff = np.zeros((o.phi, o.L, o.n2), dtype=np.uint64)         # W-coeff, matrix-major [r][l][pos]
    for r in range(o.phi):
        for pos in range(o.n2):
            for l, q in enumerate(mods):
                coeff[r, l, pos] = X[r * o.n2 + pos] % q
    ev = o.poly_to_matrix(o.wntt_forward_matrix(coeff.reshape(-1)))
    out = o.rescale_last(ev)
    back = o1.wntt_inverse_matrix(o1.matrix_to_poly(out)).reshape(o.phi, o.L - 1, o.n2)
    for r in range(0, o.phi):
        for pos in range(0, o.n2, 7):
            x = X[r * o.n2 + pos]
            c = x % ql
            c = c - ql if c > ql // 2 else c
            want = (x - c) // ql
            assert (x - c) % ql == 0
            got = 0
            for l, q in enumerate(mods[:-1]):
                M = Q1 // q
                got += M * (int(back[r, l, pos]) * pow(M % q, -1, q) % q)
            got %= Q1
            got = got - Q1 if got > Q1 // 2 else got
            assert got == want
