"""The sound sampler option: Philox4x32-10 against the Random123 known-answer vectors, the
cumulative table of the discrete gaussian against its definition, statistics of the three streams (CPU, oracle), and on
the GPU bit-exactness against the oracle plus the encryption pipeline under it."""
import math

import numpy as np
import pytest

import oracle as O

MODS = O.REF_MODULI[:3]
SEED = (0x0123456789ABCDEF, 0xFEDCBA9876543210)


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32-10
    assert O.philox4x32_10([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert O.philox4x32_10([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert O.philox4x32_10([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    # an independent restatement of the round function in Python integers
    def ref(c, k):
        c, k = list(c), list(k)
        for _ in range(10):
            p0, p1 = 0xD2511F53 * c[0], 0xCD9E8D57 * c[2]
            c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & 0xffffffff, (p0 >> 32) ^ c[3] ^ k[1], p0 & 0xffffffff]
            k = [(k[0] + 0x9E3779B9) & 0xffffffff, (k[1] + 0xBB67AE85) & 0xffffffff]
        return c
    rng = np.random.default_rng(0)
    for _ in range(50):
        c = [int(v) for v in rng.integers(0, 2 ** 32, 4)]; k = [int(v) for v in rng.integers(0, 2 ** 32, 2)]
        assert O.philox4x32_10(c, k) == ref(c, k)


def test_cdt_is_the_discrete_gaussian():
    t = [int(v) for v in O.sound_cdt()]
    assert len(t) == 83 and all(a <= b for a, b in zip(t, t[1:])) and t[-1] == 2 ** 64 - 1
    assert all(a < b for a, b in zip(t[41 - 20:41 + 20], t[41 - 19:41 + 21]))      # tails below 2^-64 round to equal entries
    w = [math.exp(-z * z / (2 * 3.2 * 3.2)) for z in range(-41, 42)]
    s = sum(w)
    acc = 0.0
    for i, z in enumerate(range(-41, 42)):
        acc += w[i] / s
        assert abs(t[i] / 2.0 ** 64 - acc) < 1e-15 or i == 82
    assert abs(t[41] - 2 ** 63) / 2.0 ** 64 < 0.07          # P(Z <= 0) = 1/2 + P(0)/2
    # symmetric: P(Z <= -k-1) == 1 - P(Z <= k)
    for k in range(0, 41):
        assert abs((t[41 - k - 1] + t[41 + k]) / 2.0 ** 64 - 1.0) < 1e-15


def test_streams_statistics_and_independence():
    n, phi = 16, 64
    u = O.sound_sample(0, MODS, n, phi, SEED, 0).reshape(phi, len(MODS), n * n)
    for l, q in enumerate(MODS):
        x = u[:, l].reshape(-1).astype(np.float64) / q
        assert (u[:, l] < q).all()
        assert abs(x.mean() - 0.5) < 4 * (1 / 12 / x.size) ** 0.5
        hist = np.histogram(x, bins=64, range=(0, 1))[0]
        chi = ((hist - x.size / 64) ** 2 / (x.size / 64)).sum()
        assert chi < 64 + 5 * (2 * 64) ** 0.5                      # 63 dof
    g = O.sound_sample(1, MODS, n, phi, SEED, 1).reshape(phi, len(MODS), n * n)
    z = np.where(g[:, 0] > MODS[0] // 2, g[:, 0].astype(np.int64) - MODS[0], g[:, 0].astype(np.int64))
    for l, q in enumerate(MODS):                                   # the same integer in every limb
        zl = np.where(g[:, l] > q // 2, g[:, l].astype(np.int64) - q, g[:, l].astype(np.int64))
        assert (zl == z).all()
    zz = z.reshape(-1).astype(np.float64)
    assert abs(zz.mean()) < 5 * 3.2 / zz.size ** 0.5
    assert abs(zz.var() - 3.2 ** 2) < 5 * (2 * 3.2 ** 4 / zz.size) ** 0.5
    assert np.abs(zz).max() <= 41
    t = O.sound_sample(2, MODS, n, phi, SEED, 2).reshape(phi, len(MODS), n)
    c = [(t[:, 0] == 0).mean(), (t[:, 0] == 1).mean(), (t[:, 0] == MODS[0] - 1).mean()]
    assert abs(sum(c) - 1) < 1e-12 and all(abs(v - 1 / 3) < 5 * (2 / 9 / t[:, 0].size) ** 0.5 for v in c)
    # different draw numbers, seeds and streams are unrelated
    assert (O.sound_sample(0, MODS, n, phi, SEED, 3) != u.reshape(-1)).mean() > 0.999
    assert (O.sound_sample(0, MODS, n, phi, (SEED[0] ^ 1, SEED[1]), 0) != u.reshape(-1)).mean() > 0.999
    assert (O.sound_sample(0, MODS, n, phi, (SEED[0], SEED[1] ^ (1 << 40)), 0) != u.reshape(-1)).mean() > 0.999


def _host(t):
    a = t.cpu().numpy()
    return a.view(np.uint64) if a.dtype == np.int64 else a


@pytest.mark.gpu
def test_sound_sampler_gpu_bit_exact_and_pipeline(cuda_device):
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    mods = O.REF_MODULI[:3]
    g = S.Context(16, 257, mods, 2.0 ** 35)
    assert ([int(v) for v in S.sampler_cdt()] == [int(v) for v in O.sound_cdt()])
    g.set_sampler(1, SEED)
    a = _host(g.sample_uniform()); e = _host(g.sample_gaussian()); t = _host(g.sample_ternary())      # draws 0, 1, 2
    assert (a == O.sound_sample(0, mods, g.n, g.phi, SEED, 0)).all()
    assert (e == O.sound_sample(1, mods, g.n, g.phi, SEED, 1)).all()
    assert (t == O.sound_sample(2, mods, g.n, g.phi, SEED, 2)).all()
    # keygen = W-CRT + X-NTT of draw 3; encrypt_pair = a: draw 4, e_re: 5, e_im: 6, bit-exact against the oracle's integer core
    o = O.SchemeCtx(16, 257, mods, 2.0 ** 35)
    sk = g.generate_secret_key()
    s_coeff = O.sound_sample(2, mods, g.n, g.phi, SEED, 3)
    want_sk = o.xy_ntt_forward(o.wntt_forward_vector(s_coeff), g.phi)
    assert (_host(sk) == want_sk).all()
    ell, i = np.meshgrid(np.arange(g.phi, dtype=np.float64), np.arange(g.n2, dtype=np.float64), indexing="ij")
    msg = ((ell + 0.001 * i) * (1 - 1j)).reshape(-1)
    m = torch.from_numpy(msg).to(cuda_device)
    re, im = g.encode_to_wntt_eval(m)
    cr, ci = g.encrypt_pair(re, im, sk)
    ocr, oci = o.encrypt_pair_with(_host(re), _host(im), _host(sk), O.sound_sample(0, mods, g.n, g.phi, SEED, 4),
                                   O.sound_sample(1, mods, g.n, g.phi, SEED, 5), O.sound_sample(1, mods, g.n, g.phi, SEED, 6))
    assert (_host(cr) == ocr).all() and (_host(ci) == oci).all()
    assert np.abs(_host(g.decrypt_and_decode(cr, ci, sk)) - msg).max() < 1e-3
    # fresh randomness per call: a second encryption of the same message shares no ciphertext half
    cr2, ci2 = g.encrypt_pair(re, im, sk)
    assert (cr2 != cr).float().mean() > 0.99 and (ci2 != ci).float().mean() > 0.99
    assert np.abs(_host(g.decrypt_and_decode(cr2, ci2, sk)) - msg).max() < 1e-3
    # the fused call works under the sound sampler too (encode, then in-place encryption)
    fr, fi = g.encode_encrypt_pair(m, sk)
    assert np.abs(_host(g.decrypt_and_decode(fr, fi, sk)) - msg).max() < 1e-3
    # back to the parity mode: the reference streams again
    g.set_sampler(0)
    assert (_host(g.sample_uniform()) == o.sample_uniform()).all()
