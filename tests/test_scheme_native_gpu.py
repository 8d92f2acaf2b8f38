"""The matrix pipeline at the reference's native shape (n = 64, p = 771, phi = 512, L = 11, delta = 2^35):
one (re, im) pair = 2 x 176 MiB of message evals, 4 x 176 MiB of ciphertext.

  * the three CMake-built end-to-end drivers of the reference, restated: main.cu (< 1e-4),
    test_encode_encrypt_decrypt_decode_wcrt.cu (< 1e-3), test_encode_decode_wcrt.cu (zero key, < 1e-3);
  * ciphertext words of one full encrypt_pair_with, bit-exact against the CPU oracle (about half a minute of
    host time: three dense W-CRTs of 11.8 G modmuls each);
  * W-CRT columns sampled against Python-integer evaluation; round trips on the full buffers.
"""
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu


def dev(x, device):
    import torch
    x = np.ascontiguousarray(x)
    return torch.from_numpy(x.view(np.int64) if x.dtype == np.uint64 else x).to(device)


def host(t):
    a = t.cpu().numpy()
    return a.view(np.uint64) if a.dtype == np.int64 else a


@pytest.fixture(scope="module")
def G(cuda_device):
    import matrix_fhe_gpu_b200.scheme as S
    return S, S.Context()


@pytest.fixture(scope="module")
def Or():
    return O.SchemeCtx()


def msg(kind, phi=512, n2=4096):
    ell, i = np.meshgrid(np.arange(phi, dtype=np.float64), np.arange(n2, dtype=np.float64), indexing="ij")
    if kind == "big":
        v = ell * 10000.0 + i; return (v - 1j * v).reshape(-1)
    if kind == "e2e":
        v = ell + 0.001 * i; return (v - 1j * v).reshape(-1)
    return ((ell + 1e-5 * i) + 1j * (ell - 1e-5 * i)).reshape(-1)


def test_native_tables(G, Or):
    S, g = G
    assert (g.n, g.phi, g.L, g.big_words) == (64, 512, 11, 7)
    assert [g.eta(l) for l in range(3)] == [10467697596351, 6077882864, 12895515671]
    for l in (0, 1, 10):
        assert (g.wcrt_table(l) == Or.wntt()[l]).all() and (g.wcrt_table(l, True) == Or.wntt(True)[l]).all()
    t, r = g.crt_tables(), Or.crt()
    assert all((t[k] == r[k]).all() for k in t)


def test_native_fp64_transforms_vs_oracle(G, Or, cuda_device):
    """n = 64 XY sandwich and phi = 512 W-DFT (the FP64 tensor-core kernels) against the oracle's plain loops;
    tolerance: 2^-40 relative to the largest magnitude (summation-order round-off only)."""
    S, g = G
    rng = np.random.default_rng(3)
    lanes = 3
    m = (rng.standard_normal(lanes * 4096) + 1j * rng.standard_normal(lanes * 4096)) * 1e3
    for got, want in ((host(g.idft2(dev(m, cuda_device), lanes)), Or.xy_idft2(m, lanes)),
                      (host(g.dft2(dev(m, cuda_device), lanes)), Or.xy_dft2(m, lanes))):
        assert np.abs(got - want).max() <= 2.0 ** -40 * np.abs(want).max()
    z = (rng.standard_normal(512 * 4096) + 1j * rng.standard_normal(512 * 4096)) * 1e3
    got = host(g.wdft_forward_complex(dev(z, cuda_device))).reshape(512, 4096)
    cols = [0, 1, 77, 4095]
    want = Or.wdft_forward_complex(z).reshape(512, 4096)
    assert np.abs(got[:, cols] - want[:, cols]).max() <= 2.0 ** -36 * np.abs(want[:, cols]).max()
    zi = z.reshape(512, 4096)
    gr, gi = g.wdft_inverse_pair(dev(zi.real.copy().reshape(-1), cuda_device), dev(zi.imag.copy().reshape(-1), cuda_device))
    wr, wi = Or.wdft_inverse_pair(zi.real.copy().reshape(-1), zi.imag.copy().reshape(-1))
    scale = max(np.abs(wr).max(), np.abs(wi).max())
    assert max(np.abs(host(gr) - wr).max(), np.abs(host(gi) - wi).max()) <= 2.0 ** -30 * scale


def test_native_encode_encrypt_fused(G, cuda_device):
    """Native shape: the fused encode + encrypt call writes the same 4 x 176 MiB of ciphertext words as the two calls."""
    import torch
    S, g = G
    sk = g.generate_secret_key()
    m = dev(msg("e2e"), cuda_device)
    re, im = g.encode_to_wntt_eval(m)
    cr, ci = g.encrypt_pair(re, im, sk)
    fr, fi = g.encode_encrypt_pair(m, sk)
    assert torch.equal(fr, cr) and torch.equal(fi, ci)
    assert np.abs(host(g.decrypt_and_decode(fr, fi, sk)) - msg("e2e")).max() < 1e-3


def test_native_trace_gemm_tensor_cores(G, Or, cuda_device):
    """n = 64: the trace GEMM runs as int8 digit GEMMs on the tensor cores (trace_tc.cu, both digit counts of the
    native moduli); bit-exact against the oracle's loops, extreme residues (0, q-1) included."""
    S, g = G
    rng = np.random.default_rng(44)
    batch = 3
    mods = Or.moduli.reshape(1, -1, 1)
    mats = [(rng.integers(0, 2 ** 63, size=(batch, Or.L, Or.n2), dtype=np.uint64) % mods) for _ in range(4)]
    for m_ in mats:
        m_[0, :, :64] = 0
        m_[1, :, :128] = (mods - 1)[0]
    mats = [m_.reshape(-1) for m_ in mats]
    d = [dev(m_, cuda_device) for m_ in mats]
    bpr, bpi = g.map_B_to_Bprime(d[2], d[3], batch)
    obr, obi = Or.trace_map(mats[2], mats[3], batch)
    assert (host(bpr) == obr).all() and (host(bpi) == obi).all()
    cr, ci = g.trace_gemm(d[0], d[1], bpr, bpi, batch)
    ocr, oci = Or.trace_gemm(mats[0], mats[1], obr, obi, batch)
    assert (host(cr) == ocr).all() and (host(ci) == oci).all()


def test_native_reference_drivers(G, cuda_device):
    import torch
    S, g = G
    sk = g.generate_secret_key()
    for kind, thr in (("main", 1e-4), ("e2e", 1e-3)):
        m = msg(kind)
        re, im = g.encode_to_wntt_eval(dev(m, cuda_device))
        cr, ci = g.encrypt_pair(re, im, sk)
        out = host(g.decrypt_and_decode(cr, ci, sk))
        err = np.abs(out - m).max()
        print(f"native {kind}: max_err = {err:.3e}")
        assert err < thr
    m = msg("big")
    re, im = g.encode_to_wntt_eval(dev(m, cuda_device))
    cr, ci = g.allocate_ciphertext(), g.allocate_ciphertext()
    cr[: g.T] = re; ci[: g.T] = im
    err = np.abs(host(g.decrypt_and_decode(cr, ci, torch.zeros_like(sk))) - m).max()
    print(f"native zero-key big: max_err = {err:.3e}")
    assert err < 1e-3 and err <= 2.0 ** -20 * np.abs(m).max()


def test_native_wcrt_sampled_columns_and_roundtrip(G, Or, cuda_device):
    S, g = G
    a = g.sample_uniform()
    ev = g.wntt_forward_matrix(a, S.MATRIX, S.MATRIX)
    assert (host(g.wntt_inverse_matrix(ev, S.MATRIX, S.MATRIX)) == host(a)).all()
    x = host(a).reshape(512, 11, 4096); y = host(ev).reshape(512, 11, 4096)
    assert (host(a) == Or.sample_uniform()).all()
    V = Or.wntt()
    rng = np.random.default_rng(0)
    for _ in range(12):
        l, c, w = int(rng.integers(11)), int(rng.integers(4096)), int(rng.integers(512))
        q = int(Or.moduli[l])
        assert int(y[w, l, c]) == sum(int(V[l, w, r]) * int(x[r, l, c]) for r in range(512)) % q


def test_native_encrypt_bit_exact_vs_oracle(G, Or, cuda_device):
    S, g = G
    sk = g.generate_secret_key()
    assert (host(sk) == Or.keygen()).all()
    rng = np.random.default_rng(1)
    q = Or.moduli.reshape(1, -1, 1)
    m_re = (rng.integers(0, 2 ** 63, size=(512, 11, 4096), dtype=np.uint64) % q).reshape(-1)
    m_im = (rng.integers(0, 2 ** 63, size=(512, 11, 4096), dtype=np.uint64) % q).reshape(-1)
    a = host(g.sample_uniform()); e = host(g.sample_gaussian())
    cr, ci = g.encrypt_pair(dev(m_re, cuda_device), dev(m_im, cuda_device), sk)
    ocr, oci = Or.encrypt_pair_with(m_re, m_im, host(sk), a, e, e)
    assert (host(cr) == ocr).all() and (host(ci) == oci).all()
    assert (host(g.decrypt_to_eval(cr, sk, S.POLY)) == Or.decrypt_to_eval(ocr, host(sk))).all()


def test_native_wcrt_three_paths_identical(G, cuda_device, monkeypatch):
    """The tcgen05 (TMEM) kernel, the mma.sync kernel and the CUDA-core kernel give the same words for every
    layout pair and every fused source; the CUDA-core kernel is the one checked against the oracle on the
    small shapes of test_scheme_gpu.py, the default (tcgen05) one by the tests above."""
    import torch
    S, g = G
    a = g.sample_uniform()
    m = dev(msg("e2e"), cuda_device)
    i64 = torch.arange(g.phi * g.n2, dtype=torch.int64, device=cuda_device).reshape(g.phi, g.n2) % 17 - 8
    ref = None
    for path in ("cuda", "mma", "umma"):
        monkeypatch.setenv("MFHE_WCRT_PATH", path)
        got = []
        a_poly = g.layout_convert(a, S.MATRIX, S.POLY)          # inputs must be canonical residues of THEIR limb
        for lin, lout in ((S.MATRIX, S.POLY), (S.POLY, S.MATRIX), (S.MATRIX, S.MATRIX), (S.POLY, S.POLY)):
            x = a_poly if lin == S.POLY else a
            got.append(g.wntt_forward_matrix(x, lin, lout).clone())
            got.append(g.wntt_inverse_matrix(x, lin, lout).clone())
        re, im = g.encode_to_wntt_eval(m)                 # quantiser sources
        got += [re.clone(), im.clone()]
        got.append(g.wntt_forward_centered(i64.reshape(-1)).clone())      # int64 source + CRT compose
        got.append(g.sample_gaussian().clone())
        torch.cuda.synchronize()
        if ref is None:
            ref = got
        else:
            for k, (x, y) in enumerate(zip(ref, got)):
                assert torch.equal(x, y), f"path {path} differs from cuda-core at output {k}"
    monkeypatch.delenv("MFHE_WCRT_PATH")


def test_device_built_tables_equal_host_closed_form(G, cuda_device, monkeypatch):
    """the W-CRT tables come from wtables_gpu.cu (one CTA per (limb, lane), closed-form inverse);
    MFHE_TABLES_HOST=1 selects the host closed form.  Every table word of every limb and the transforms through all
    four device layouts (GEMM tables, digit planes, UMMA tiles) must agree, and both equal the oracle (test_native_tables)."""
    import torch
    S, g = G
    monkeypatch.setenv("MFHE_TABLES_HOST", "1")
    h = S.Context()
    monkeypatch.delenv("MFHE_TABLES_HOST")
    for l in range(g.L):
        assert (g.wcrt_table(l) == h.wcrt_table(l)).all() and (g.wcrt_table(l, True) == h.wcrt_table(l, True)).all()
    a = g.sample_uniform()
    for path in ("cuda", "mma", "umma"):
        monkeypatch.setenv("MFHE_WCRT_PATH", path)
        assert torch.equal(g.wntt_forward_matrix(a), h.wntt_forward_matrix(a))
        assert torch.equal(g.wntt_inverse_matrix(a, S.MATRIX, S.MATRIX), h.wntt_inverse_matrix(a, S.MATRIX, S.MATRIX))
    monkeypatch.delenv("MFHE_WCRT_PATH")
