"""BASELINE config 4's shape: the native ring (n = 64, p = 771, phi = 512) with L = 24 moduli (the
reference's rule continued -- its 45-bit prime, its ten 35-bit primes, then the next 13 primes = 1 mod 197376).
The reference cannot run this shape (config.h fixes L = 11), so parity is against the CPU oracle:

  * ciphertext words of sampled rows (w, limb, y) of a full encrypt_pair_with, every word of the row, against the
    oracle's tables and small-NTT restatement evaluated with exact integers (the dense W-CRT couples all 512 lanes of a
    column, the X-NTT couples the 64 x of a row, so a row is the smallest independent unit);
  * the 14-word CRT composition (Q ~ 2^850) forced with MFHE_CRT_FULL=1 against the two-limb fast path, word for word;
  * decode within the drivers' tolerance (< 1e-3) and the north-star bound 2^-20 * max|m|;
  * the same shape at n = 16 bit-exact against the oracle on EVERY ciphertext word (the oracle finishes that in seconds).
"""
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu

L24_EXTRA = [17194410241, 17194804993, 17195989249, 17196581377, 17196976129, 17197963009, 17199936769, 17201318401,
             17201515777, 17203884289, 17206844929, 17207437057, 17209608193]
MODS24 = O.REF_MODULI + L24_EXTRA


def dev(x, device):
    import torch
    x = np.ascontiguousarray(x)
    return torch.from_numpy(x.view(np.int64) if x.dtype == np.uint64 else x).to(device)


def host(t):
    a = t.cpu().numpy()
    return a.view(np.uint64) if a.dtype == np.int64 else a


def msg(phi, n2):
    ell, i = np.meshgrid(np.arange(phi, dtype=np.float64), np.arange(n2, dtype=np.float64), indexing="ij")
    v = ell + 0.001 * i
    return (v - 1j * v).reshape(-1)


def test_moduli_follow_the_reference_rule():
    assert len(MODS24) == 24 and len(set(MODS24)) == 24
    assert all((q - 1) % 197376 == 0 and O.lib().mfo_is_prime(q) for q in MODS24)
    assert all(a < b for a, b in zip(MODS24[1:], MODS24[2:]))          # consecutive primes of the rule, ascending


def test_L24_n16_every_ciphertext_word_vs_oracle(cuda_device):
    import matrix_fhe_gpu_b200.scheme as S
    g = S.Context(16, 257, MODS24, 2.0 ** 35)
    o = O.SchemeCtx(16, 257, MODS24, 2.0 ** 35)
    assert (g.L, g.big_words) == (24, 14) and o.big_words == 14
    sk = g.generate_secret_key()
    assert (host(sk) == o.keygen()).all()
    rng = np.random.default_rng(24)
    q = np.asarray(MODS24, dtype=np.uint64).reshape(1, -1, 1)
    m_re = (rng.integers(0, 2 ** 63, size=(g.phi, g.L, g.n2), dtype=np.uint64) % q).reshape(-1)
    m_im = (rng.integers(0, 2 ** 63, size=(g.phi, g.L, g.n2), dtype=np.uint64) % q).reshape(-1)
    a = host(g.sample_uniform()); e = host(g.sample_gaussian())
    assert (a == o.sample_uniform()).all()
    cr, ci = g.encrypt_pair(dev(m_re, cuda_device), dev(m_im, cuda_device), sk)
    ocr, oci = o.encrypt_pair_with(m_re, m_im, host(sk), a, e, e)
    assert (host(cr) == ocr).all() and (host(ci) == oci).all()
    assert (host(g.decrypt_to_eval(cr, sk, S.POLY)) == o.decrypt_to_eval(ocr, host(sk))).all()
    # full pipeline at this width, decode through the 14-word composition in the oracle and the fast path on the GPU
    m = msg(g.phi, g.n2)
    re, im = g.encode_to_wntt_eval(dev(m, cuda_device))
    cr, ci = g.encrypt_pair(re, im, sk)
    out = host(g.decrypt_and_decode(cr, ci, sk))
    want = o.decrypt_and_decode(host(cr), host(ci), host(sk))
    assert np.abs(out - m).max() < 1e-3
    assert np.abs(out - want).max() <= 2.0 ** -30 * np.abs(m).max()     # same integers in, fp64 round-off only


@pytest.fixture(scope="module")
def G64(cuda_device):
    import matrix_fhe_gpu_b200.scheme as S
    return S, S.Context(64, 257, MODS24, 2.0 ** 35), O.SchemeCtx(64, 257, MODS24, 2.0 ** 35)


def test_L24_native_sampled_rows_vs_oracle(G64, cuda_device):
    S, g, o = G64
    assert (g.n, g.phi, g.L, g.big_words) == (64, 512, 24, 14)
    for l in (0, 11, 23):
        assert (g.wcrt_table(l) == o.wntt()[l]).all()
    sk = g.generate_secret_key()
    a = g.sample_uniform(); e = g.sample_gaussian()
    m_re = (a * 3 + 1) % dev(np.repeat(np.asarray(MODS24, dtype=np.uint64), g.n2)[None].repeat(g.phi, 0).reshape(-1), cuda_device)
    m_im = (a * 5 + 2) % dev(np.repeat(np.asarray(MODS24, dtype=np.uint64), g.n2)[None].repeat(g.phi, 0).reshape(-1), cuda_device)
    cr, ci = g.encrypt_pair(m_re, m_im, sk)
    A = host(a).reshape(g.phi, g.L, g.n, g.n); E = host(e).reshape(g.phi, g.L, g.n, g.n)
    MR = host(m_re).reshape(g.phi, g.L, g.n, g.n); MI = host(m_im).reshape(g.phi, g.L, g.n, g.n)
    CR = host(cr).reshape(2, g.phi, g.L, g.n, g.n); CI = host(ci).reshape(2, g.phi, g.L, g.n, g.n)
    SK = host(sk).reshape(g.phi, g.L, g.n)
    V = o.wntt()
    rng = np.random.default_rng(4)
    for _ in range(6):
        w, l, y = int(rng.integers(g.phi)), int(rng.integers(g.L)), int(rng.integers(g.n))
        qq = MODS24[l]
        row = lambda X: [sum(int(V[l, w, r]) * int(X[r, l, y, x]) for r in range(g.phi)) % qq for x in range(g.n)]
        a_ev, e_ev = row(A), row(E)
        assert [int(v) for v in CR[1, w, l, y]] == a_ev and [int(v) for v in CI[1, w, l, y]] == a_ev      # a, stored in both
        t = O.PhantomTables(6, [qq])
        an = t.forward(np.asarray(a_ev, dtype=np.uint64).reshape(1, 1, g.n))[0, 0]
        prod = np.asarray([int(x_) * int(s_) % qq for x_, s_ in zip(an, SK[w, l])], dtype=np.uint64)
        tt = t.inverse(prod.reshape(1, 1, g.n))[0, 0]
        for C_, M_ in ((CR, MR), (CI, MI)):
            want = [(int(M_[w, l, y, x]) - int(tt[x]) + e_ev[x]) % qq for x in range(g.n)]
            assert [int(v) for v in C_[0, w, l, y]] == want


def test_L24_native_full_crt_equals_fast_path_and_decodes(G64, cuda_device, monkeypatch):
    import torch
    S, g, o = G64
    m = msg(g.phi, g.n2)
    sk = g.generate_secret_key()
    re, im = g.encode_to_wntt_eval(dev(m, cuda_device))
    cr, ci = g.encrypt_pair(re, im, sk)
    fast = g.decrypt_and_decode(cr, ci, sk).clone()
    monkeypatch.setenv("MFHE_CRT_FULL", "1")               # every coefficient through the 14-word composition (H4)
    g2 = S.Context(64, 257, MODS24, 2.0 ** 35)
    full = g2.decrypt_and_decode(cr, ci, sk).clone()
    monkeypatch.delenv("MFHE_CRT_FULL")
    assert torch.equal(fast.view(torch.float64), full.view(torch.float64))
    err = np.abs(host(fast) - m).max()
    assert err < 1e-3 and err <= 2.0 ** -20 * np.abs(m).max()
    fr, fi = g.encode_encrypt_pair(dev(m, cuda_device), sk)
    assert torch.equal(fr, cr) and torch.equal(fi, ci)
    # CRT words themselves: compose_big at 14 words on a lane of decrypted evals against the oracle
    ev = g.decrypt_to_eval(cr, sk, S.MATRIX)
    co = g.wntt_inverse_matrix(ev, S.MATRIX, S.MATRIX)
    lane = host(co).reshape(g.phi, g.L * g.n2)[7]
    mag, neg = g.crt_compose_centerlift_big(dev(lane, cuda_device))
    omag, oneg = o.crt_compose_big(lane)
    assert (host(mag) == omag).all() and (host(neg).view(np.uint8) == oneg).all()
