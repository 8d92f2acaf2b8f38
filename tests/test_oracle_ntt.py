"""Pins the CPU oracle's NTT part (oracle/ntt_oracle.c) against the reference's known answers.

Sources of truth, in order:
  * tests/golden/phantom_tables.json -- produced by the REAL Phantom host code
    (ph/src/host/ntt.cu, numth.cu, modulus.cu) through oracle/_ref, see make_golden.py;
  * constants derived from the reference's algorithms in SURVEY.md 8(c);
  * the reference drivers' own checks: round trips and the X^64 - i schoolbook product of
    a_j = j+1, b_j = j+3 (test/test_custom_ntt_roundtrip.cu:63-166, 256-319) and Phantom's
    constant-vector round trips (ph/test/ntt_test.cu).
"""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle as O

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "phantom_tables.json")))
REF_MODULI = [17592186435073, 17182765057, 17184541441, 17186120449, 17186515201, 17186909953,
              17188883713, 17190462721, 17190857473, 17191844353, 17192831233]


def test_survey_constants():
    L = O.lib()
    assert [L.mfo_gl_get_psi(q, 64) for q in REF_MODULI[:3]] == [9038641396914, 213176194, 708412574]
    assert [L.mfo_min_primitive_root(128, q) for q in REF_MODULI[:3]] == [719028594519, 96701262, 6333595]
    assert all((q - 1) % 197376 == 0 and L.mfo_is_prime(q) for q in REF_MODULI)


def test_prime_lists_match_phantom():
    assert [int(x) for x in O.get_primes(4096, 60, 1)] == GOLD["get_primes_4096_60_1"] == [1152921504606830593]
    assert [int(x) for x in O.coeff_modulus_create(65536, 60, 32)] == GOLD["create_65536_60x32"]
    assert [int(x) for x in O.coeff_modulus_create(65536, 50, 32)] == GOLD["create_65536_50x32"]
    assert [int(x) for x in O.get_primes(131072, 55, 64)] == GOLD["get_primes_131072_55_64"]
    c = GOLD["create_65536_60x32"]
    assert c[-1] == 1152921504606584833 and c[0] == 1152921504538951681  # k-th found is returned last


@pytest.mark.parametrize("rec", GOLD["tables"], ids=lambda r: f"logn{r['logn']}-q{r['q']}")
def test_tables_match_phantom(rec):
    t = O.PhantomTables(rec["logn"], [rec["q"]])
    assert t.psi(0) == rec["root"]
    assert int(t.n_inv[0]) == rec["n_inv"] and int(t.n_inv_shoup[0]) == rec["n_inv_shoup"]
    for name, arr in (("tw", t.tw), ("tw_shoup", t.tws), ("itw", t.itw), ("itw_shoup", t.itws)):
        assert hashlib.sha256(arr[0].tobytes()).hexdigest() == rec[name + "_sha256"], name
        assert [int(x) for x in arr[0][:8]] == rec[name + "_head"]


@pytest.mark.skipif(O.ref() is None, reason="oracle/_ref not built (reference tree absent)")
def test_tables_against_live_reference():
    import ctypes as C
    R = O.ref()
    for logn, q in ((6, REF_MODULI[0]), (13, int(O.get_primes(8192, 50, 1)[0]))):
        n = 1 << logn
        t = O.PhantomTables(logn, [q])
        a = [np.zeros(n, dtype=np.uint64) for _ in range(4)]
        root, ni, nis = C.c_uint64(), C.c_uint64(), C.c_uint64()
        assert R.ref_ntt_tables(logn, q, C.byref(root), *a, C.byref(ni), C.byref(nis)) == 0
        assert root.value == t.psi(0)
        for x, y in zip(a, (t.tw, t.tws, t.itw, t.itws)):
            assert (x == y[0]).all()


@pytest.mark.parametrize("logn,q", [(6, REF_MODULI[0]), (6, REF_MODULI[1]), (8, 1152921504606830593)])
def test_forward_is_definitional(logn, q):
    n = 1 << logn
    t = O.PhantomTables(logn, [q])
    x = O.synth_polys(1, 1, [q], n)
    out = np.zeros(n, dtype=np.uint64)
    O.lib().mfo_negacyclic_forward_def(x.reshape(-1).copy(), out, logn, q, t.psi(0))
    assert (t.forward(x).reshape(-1) == out).all()


@pytest.mark.parametrize("logn,bits,limbs", [(8, 50, 1), (11, 50, 10), (12, 50, 1), (14, 50, 10), (12, 60, 1)])
def test_phantom_constant_vector_roundtrip(logn, bits, limbs):
    # ph/test/ntt_test.cu: all-ones / all-twos through fwd -> inv must come back exactly
    n = 1 << logn
    mods = O.coeff_modulus_create(n, bits, limbs)
    t = O.PhantomTables(logn, mods)
    for c in (1, 2):
        x = np.full((1, limbs, n), c, dtype=np.uint64)
        f = t.forward(x)
        assert (f < mods.reshape(1, -1, 1)).all()
        assert (t.inverse(f) == x).all()


def test_negacyclic_polymul_vs_schoolbook_n4096():
    # BASELINE config 1a: N=2^12, q = 2^60 - 2^14 + 1, a_j=j+1, b_j=j+3 and seeded inputs
    q, logn = 1152921504606830593, 12
    n = 1 << logn
    t = O.PhantomTables(logn, [q])
    L = O.lib()
    cases = [(np.arange(1, n + 1, dtype=np.uint64) % np.uint64(q), np.arange(3, n + 3, dtype=np.uint64) % np.uint64(q))]
    cases.append((O.synth_polys(1, 1, [q], n).reshape(-1), O.synth_polys(2, 1, [q], n).reshape(-1)))
    for a, b in cases:
        fa, fb = t.forward(a.reshape(1, 1, n)).reshape(-1), t.forward(b.reshape(1, 1, n)).reshape(-1)
        prod = np.array([L.mfo_mulmod(int(x), int(y), q) for x, y in zip(fa, fb)], dtype=np.uint64)
        c = t.inverse(prod.reshape(1, 1, n)).reshape(-1)
        ref = np.zeros(n, dtype=np.uint64)
        L.mfo_schoolbook_mul(np.ascontiguousarray(a), np.ascontiguousarray(b), ref, n, q, q - 1)
        assert (c == ref).all()


@pytest.mark.parametrize("limb", range(11))
def test_gl_roundtrip_and_polymul_n64(limb):
    # test/test_custom_ntt_roundtrip.cu:115-166 (gl round trip) and :256-319 (X^n - i product)
    q, n = REF_MODULI[limb], 64
    L = O.lib()
    x = np.array([(3 + limb + j + 1) % q for j in range(n)], dtype=np.uint64)   # (b+l+x+1) mod q pattern
    y = x.copy()
    L.mfo_gl_forward(y, n, q)
    L.mfo_gl_backward(y, n, q)
    assert (x == y).all()
    a = np.arange(1, n + 1, dtype=np.uint64)
    b = np.arange(3, n + 3, dtype=np.uint64)
    fa, fb = a.copy(), b.copy()
    L.mfo_gl_forward(fa, n, q)
    L.mfo_gl_forward(fb, n, q)
    prod = np.array([L.mfo_mulmod(int(u), int(v), q) for u, v in zip(fa, fb)], dtype=np.uint64)
    L.mfo_gl_backward(prod, n, q)
    psi = L.mfo_gl_get_psi(q, n)
    ref = np.zeros(n, dtype=np.uint64)
    L.mfo_schoolbook_mul(a, b, ref, n, q, L.mfo_powmod(psi, n, q))
    assert (prod == ref).all()
    # definition: A[k] = sum_j a_j psi^j w^(jk) = a(psi^(4k+1))
    for k in (0, 1, 17, 63):
        pt = L.mfo_powmod(psi, 4 * k + 1, q)
        acc, p = 0, 1
        for j in range(n):
            acc = (acc + int(a[j]) * p) % q
            p = p * pt % q
        assert int(fa[k]) == acc


def test_gl_perm_table():
    perm = np.zeros(64, dtype=np.uint32)
    inv = np.zeros(64, dtype=np.uint32)
    O.lib().mfo_gl_perm(64, perm, inv)
    assert sorted(perm) == list(range(64)) and perm[0] == 0
    e = 1
    for j in range(64):
        assert perm[j] == int(format((e - 1) // 4, "06b")[::-1], 2)
        assert inv[perm[j]] == j
        e = e * 5 % 256
