"""Encrypted matrix product on the GPU: every new entry point bit-exact against the oracle's
restatement, and the whole driver (encode -> twist -> encrypt -> 4 component GEMMs on the tensor cores -> 4-term
decrypt -> decode) against plain numpy linear algebra at the reference's native shape."""
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


def rand_sets(rng, o, count):
    q = np.repeat(np.asarray(o.moduli, dtype=np.uint64), o.n2)[None].repeat(o.phi, 0).reshape(-1)
    out = [rng.integers(0, 2 ** 62, size=o.T, dtype=np.uint64) % q for _ in range(count)]
    for v in out:                                   # extreme residues
        v[:64] = 0
        v[64:128] = (q - 1)[64:128]
    return out


@pytest.mark.parametrize("n,f,bits,L", [(64, 5, 45, 3), (64, 5, 35, 2), (64, 5, 30, 2), (64, 5, 48, 1), (16, 5, 45, 5), (32, 17, 50, 2)])
def test_matmul_pieces_bit_exact_vs_oracle(cuda_device, n, f, bits, L):
    """n = 64 runs the tensor-core kernels -- mfhe_trace_gemm_ty on trace_tc.cu (mma.sync, transposed operand staging),
    mfhe_ct_matmul on ct_matmul_umma.cu (tcgen05) -- with 6-, 5- and 4-digit moduli (48 / 45, 35, 30 bits); the other shapes
    the CUDA-core kernel; 50-bit moduli (7 digits) exercise the CUDA-core fallback's range."""
    import matrix_fhe_gpu_b200.scheme as S
    mods = O.primes_congruent_one(2 * n * 3 * f, bits, L)
    g = S.Context(n, f, mods, 2.0 ** 30)
    o = O.SchemeCtx(n, f, mods, 2.0 ** 30)
    rng = np.random.default_rng(n + f)
    ur, ui, vr, vi = rand_sets(rng, o, 4)
    d = [dev(x, cuda_device) for x in (ur, ui, vr, vi)]
    # twist along x
    gpr, gpi = g.trace_map_x(d[2], d[3])
    opr, opi = o.trace_map_x(vr, vi, o.phi)
    assert (host(gpr) == opr).all() and (host(gpi) == opi).all()
    # GEMM contracting over y
    gdr, gdi = g.trace_gemm_ty(d[0], d[1], gpr, gpi)
    odr, odi = o.trace_gemm_ty(ur, ui, opr, opi, o.phi)
    assert (host(gdr) == odr).all() and (host(gdi) == odi).all()
    # transpose, per-limb rescale
    t = host(g.transpose(d[0])).reshape(o.phi * o.L, n, n)
    assert (t == ur.reshape(o.phi * o.L, n, n).transpose(0, 2, 1)).all()
    inv = [pow(n, -1, int(q)) for q in mods]
    rr, ri = g.rescale_limbs(gdr.clone(), gdi.clone(), inv)
    qv = np.repeat(np.asarray(mods, dtype=object), o.n2)[None].repeat(o.phi, 0).reshape(-1)
    iv = np.repeat(np.asarray(inv, dtype=object), o.n2)[None].repeat(o.phi, 0).reshape(-1)
    assert (host(rr).astype(object) == (odr.astype(object) * iv) % qv).all()
    assert (host(ri).astype(object) == (odi.astype(object) * iv) % qv).all()
    # ciphertext-level product and its decryption: every word
    sk = g.generate_secret_key()
    assert (host(sk) == o.keygen()).all()
    cu = g.encrypt_pair(d[0], d[1], sk)
    cv = g.encrypt_pair(gpr, gpi, sk)
    prod = g.ct_matmul(cu[0], cu[1], cv[0], cv[1])
    oprod = o.ct_matmul(host(cu[0]), host(cu[1]), host(cv[0]), host(cv[1]))
    assert (host(prod) == oprod).all()
    pr, pi = g.ct_matmul_decrypt(prod, sk)
    wr, wi = o.ct_matmul_decrypt(oprod, host(sk))
    assert (host(pr) == wr).all() and (host(pi) == wi).all()
    # and it decrypts to the plaintext bilinear form n * U^T V' exactly up to the two ciphertexts' errors: compare with
    # the product of the DECRYPTED operands (exact identity in Z_q)
    du = [o.poly_to_matrix(o.decrypt_to_eval(host(c), host(sk))) for c in cu]
    dv = [o.poly_to_matrix(o.decrypt_to_eval(host(c), host(sk))) for c in cv]
    xr, xi = o.trace_gemm_ty(du[0], du[1], dv[0], dv[1], o.phi)
    tr = lambda a: np.ascontiguousarray(a.reshape(o.phi * o.L, n, n).transpose(0, 2, 1)).reshape(-1)
    assert (wr == tr(xr)).all() and (wi == tr(xi)).all()


def test_encrypted_matmul_small_vs_plaintext(cuda_device):
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    mods = O.primes_congruent_one(2 * 64 * 15, 45, 5)
    g = S.Context(64, 5, mods, 2.0 ** 30)
    rng = np.random.default_rng(3)
    U = (rng.standard_normal((g.phi, g.n, g.n)) + 1j * rng.standard_normal((g.phi, g.n, g.n))) * 3
    V = (rng.standard_normal((g.phi, g.n, g.n)) + 1j * rng.standard_normal((g.phi, g.n, g.n))) * 3
    sk = g.generate_secret_key()
    R = host(g.encrypted_matmul(dev(U, cuda_device), dev(V, cuda_device), sk))
    want = np.einsum("lab,lac->lbc", V.conj(), U)
    err = np.abs(R - want).max()
    print(f"encrypted matmul n=64 phi=8: max_err {err:.3e}, entries up to {np.abs(want).max():.1f}")
    assert err <= 2.0 ** -20 * np.abs(want).max()


def test_encrypted_matmul_native_shape(cuda_device):
    """512 lanes of 64 x 64 complex matrices, L = 11, delta = 2^35: R_l = V_l^H U_l for every lane from ciphertexts only.
    Tolerance 2^-20 * max|entry| (north star); the noise analysis in DESIGN.md predicts ~2^-28."""
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    g = S.Context()
    rng = np.random.default_rng(11)
    U = (rng.standard_normal((g.phi, g.n, g.n)) + 1j * rng.standard_normal((g.phi, g.n, g.n)))
    V = (rng.standard_normal((g.phi, g.n, g.n)) + 1j * rng.standard_normal((g.phi, g.n, g.n)))
    U[0] = np.eye(g.n); V[1] = np.eye(g.n)                    # R_0 = V_0^H, R_1 = U_1: easy to eyeball
    sk = g.generate_secret_key()
    R = host(g.encrypted_matmul(dev(U, cuda_device), dev(V, cuda_device), sk))
    want = np.einsum("lab,lac->lbc", V.conj(), U)
    err = np.abs(R - want).max()
    print(f"encrypted matmul native shape: max_err {err:.3e}, entries up to {np.abs(want).max():.1f}")
    assert err <= 2.0 ** -20 * np.abs(want).max()
    assert np.abs(R[0] - V[0].conj().T).max() < 1e-4 and np.abs(R[1] - U[1]).max() < 1e-4


def test_ct_matmul_tcgen05_equals_mma_sync_native_shape(cuda_device, monkeypatch):
    """mfhe_ct_matmul at the native shape (512 lanes, L = 11: ten 5-digit limbs and one 6-digit limb): the tcgen05 kernel
    (ct_matmul_umma.cu: stacked M = 128, digit-stack instructions, TMEM accumulators) against the four mma.sync GEMMs of
    trace_tc.cu, every word of the 8 poly sets; extreme residues (0, q - 1) included; twice (persistent CTAs reuse TMEM)."""
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    g = S.Context()
    q = dev(np.repeat(np.asarray(g.moduli, dtype=np.uint64), g.n2)[None].repeat(g.phi, 0).reshape(-1), cuda_device)
    cts = []
    for seed in range(4):
        a = g.sample_uniform()
        x = torch.cat([(a * (3 + seed) + seed) % q, (a * (7 + seed) + 1) % q])       # [b | a] halves, canonical residues
        x[:64] = 0
        x[64:128] = q[64:128] - 1
        x[g.T + 128:g.T + 192] = q[128:192] - 1
        cts.append(x)
    monkeypatch.setenv("MFHE_MATMUL_PATH", "m")
    ref = g.ct_matmul(*cts).clone()
    monkeypatch.delenv("MFHE_MATMUL_PATH")
    for _ in range(2):
        got = g.ct_matmul(*cts)
        assert torch.equal(got, ref)


@pytest.mark.parametrize("L", [11, 3])
def test_ct_matmul_decrypt_fused_equals_pass_sequence(cuda_device, monkeypatch, L):
    """mfhe_ct_matmul_decrypt at n = 64: the fused kernel (R1, R2 in shared memory, the outer convolution down the columns
    of the tile; scheme.cu) against the sequence of three row-kernel passes and four transposes, every word, on both
    arithmetics of the row networks (FP64 pipe / integer)."""
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    g = S.Context(moduli=S.REF_MODULI[:L])
    q = dev(np.repeat(np.asarray(g.moduli, dtype=np.uint64), g.n2)[None].repeat(g.phi, 0).reshape(-1), cuda_device)
    a = g.sample_uniform()
    prod = torch.cat([(a * (k + 2) + k) % q for k in range(8)])
    prod[:64] = 0
    prod[4 * g.T:4 * g.T + 64] = q[:64] - 1
    prod[6 * g.T + 64:6 * g.T + 128] = q[64:128] - 1
    sk = g.generate_secret_key()
    monkeypatch.setenv("MFHE_MATMUL_DECRYPT_PATH", "r")
    ref = [t.clone() for t in g.ct_matmul_decrypt(prod.clone(), sk)]
    monkeypatch.delenv("MFHE_MATMUL_DECRYPT_PATH")
    for arith in (1, 0):
        g.set_row_arith(arith)
        for _ in range(2):
            got = g.ct_matmul_decrypt(prod.clone(), sk)
            assert torch.equal(got[0], ref[0]) and torch.equal(got[1], ref[1]), arith
    g.set_row_arith(1)
