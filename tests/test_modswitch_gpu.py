"""SURVEY 8(f) rank 2 on the GPU: the matrix scheme's RNS rescale and the flat family's mod-up / mod-down with an
auxiliary base, bit-exact against the oracle, plus their meaning (division by the dropped modulus / by P)."""
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


@pytest.mark.parametrize("n,f,L", [(16, 5, 4), (16, 257, 3), (32, 17, 2)])
def test_rescale_last_bit_exact_vs_oracle(cuda_device, n, f, L):
    """(16, 257, 3) runs the tcgen05 W-CRT for both passes (phi = 512), the others the CUDA-core kernels."""
    import matrix_fhe_gpu_b200.scheme as S
    mods = O.primes_congruent_one(2 * n * 3 * f, 45, 1) + O.primes_congruent_one(2 * n * 3 * f, 35, L - 1)
    g = S.Context(n, f, mods, 2.0 ** 30)
    o = O.SchemeCtx(n, f, mods, 2.0 ** 30)
    rng = np.random.default_rng(L)
    q = np.repeat(np.asarray(mods, dtype=np.uint64), o.n2)[None].repeat(o.phi, 0).reshape(-1)
    x = rng.integers(0, 2 ** 62, size=o.T, dtype=np.uint64) % q
    x[:32] = 0; x[32:64] = (q - 1)[32:64]
    assert (host(g.rescale_last(dev(x, cuda_device))) == o.rescale_last(x)).all()


def test_matmul_then_rescale_decodes_at_the_original_scale(cuda_device):
    """A product carries delta^2; rescaling by the last modulus q_last ~ delta brings it back to delta^2 / q_last ~ delta,
    one limb shorter -- the CKKS level structure the reference's moduli are laid out for (delta = 2^35 next to ten 35-bit
    primes, config.h:25-44).  Decoded in a context on the remaining limbs.
    (a) plaintext bilinear form n U^T V' (exact integers): the rescale adds at most 1/2 per W-coefficient -> 2^-20 bound.
    (b) the 4-component product CIPHERTEXT, every component rescaled, then decrypted: the rounding of the (a_U, a_V')
        component meets the key twice at decryption, and one product with the ternary key in Z[W]/Phi_771 amplifies a
        uniform rounding error by ~300 (512-term convolution x reduction modulo Phi_771), so without a relinearisation
        before the rescale the error is ~2^-11 relative at this scale: asserted at 2^-8, printed."""
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    mods = O.REF_MODULI[:7]                               # 45 + 6 x 35 bits
    g = S.Context(16, 257, mods, 2.0 ** 35)
    g1 = S.Context(16, 257, mods[:-1], 2.0 ** 35)
    rng = np.random.default_rng(8)
    U = (rng.standard_normal((g.phi, g.n, g.n)) + 1j * rng.standard_normal((g.phi, g.n, g.n)))
    V = (rng.standard_normal((g.phi, g.n, g.n)) + 1j * rng.standard_normal((g.phi, g.n, g.n)))
    want = np.einsum("lab,lac->lbc", V.conj(), U)
    sk = g.generate_secret_key()
    cl = torch.from_numpy(g.conj_lane()).to(cuda_device)
    ur, ui = g.encode_to_wntt_eval(dev(U, cuda_device).reshape(-1))
    vr, vi = g.encode_to_wntt_eval(dev(V, cuda_device).reshape(g.phi, -1)[cl].reshape(-1))
    vpr, vpi = g.trace_map_x(vr, vi)
    scale = g.delta / mods[-1]                            # decode divides by delta once more
    # (a) plaintext
    dr, di = g.trace_gemm_ty(ur, ui, vpr, vpi)
    pr, pi = g.transpose(g.rescale_last(dr).view(-1)), g.transpose(g.rescale_last(di).view(-1))
    out = host(g1.decode_eval_pair(pr, pi, S.MATRIX)).reshape(g.phi, g.n, g.n) / scale
    err = np.abs(out - want).max()
    print(f"plaintext product -> rescale -> decode: max_err {err:.3e}, entries up to {np.abs(want).max():.1f}")
    assert err <= 2.0 ** -20 * np.abs(want).max()
    # (b) ciphertext
    cu = g.encrypt_pair(ur, ui, sk); cv = g.encrypt_pair(vpr, vpi, sk)
    prod = g.ct_matmul(cu[0], cu[1], cv[0], cv[1])
    T1 = g1.T
    prod1 = torch.empty(8 * T1, dtype=torch.int64, device=cuda_device)
    for k in range(8):
        prod1[k * T1:(k + 1) * T1] = g.rescale_last(prod[k * g.T:(k + 1) * g.T])
    sk1 = sk.view(g.phi, g.L, g.n)[:, : g.L - 1].contiguous().view(-1)       # the same key on the remaining limbs
    pr, pi = g1.ct_matmul_decrypt(prod1, sk1)
    out = host(g1.decode_eval_pair(pr, pi, S.MATRIX)).reshape(g.phi, g.n, g.n) / scale
    err = np.abs(out - want).max()
    print(f"ciphertext product -> rescale x 8 -> decrypt -> decode: max_err {err:.3e} (relative 2^{np.log2(err / np.abs(want).max()):.1f})")
    assert err <= 2.0 ** -8 * np.abs(want).max()


@pytest.mark.parametrize("logn,lq,lp,bits,batch", [(12, 4, 2, 50, 3), (13, 6, 3, 55, 1), (12, 3, 3, 60, 2)])
def test_modup_moddown_bit_exact_and_meaning(cuda_device, logn, lq, lp, bits, batch):
    import matrix_fhe_gpu_b200 as m
    import matrix_fhe_gpu_b200.scheme as S
    N = 1 << logn
    primes = [int(v) for v in O.get_primes(2 * N, bits, lq + lp)]
    q, p = primes[:lq], primes[lq:]
    plan = m.NttPlan(logn, q + p)
    tq, tqp = O.PhantomTables(logn, q), O.PhantomTables(logn, q + p)
    x = O.synth_polys(3, batch, q, N)                                  # coefficient form over Q
    # mod-up: [x | bConv(x)] then NTT of every limb
    up = host(S.modup(plan, S.BconvPlan(q, p), dev(x, cuda_device), batch)).reshape(batch, lq + lp, N)
    ext = np.concatenate([x, np.stack([O.bconv(x[b], q, p) for b in range(batch)])], axis=1)
    assert (up == tqp.forward(ext)).all()
    # mod-down of an arbitrary NTT-form element over Q u P
    y = O.synth_polys(5, batch, q + p, N)
    dn = host(S.moddown(plan, S.BconvPlan(p, q), dev(y, cuda_device), batch)).reshape(batch, lq, N)
    coeff = tqp.inverse(y)
    t = tq.forward(np.stack([O.bconv(coeff[b, lq:], p, q) for b in range(batch)]))
    P = 1
    for v in p:
        P *= v
    want = np.empty_like(t)
    for i, qi in enumerate(q):
        pinv = pow(P % qi, -1, qi)
        want[:, i] = ((y[:, i].astype(object) - t[:, i].astype(object)) * pinv % qi).astype(np.uint64)
    assert (dn == want).all()
    # meaning: mod-down(mod-up(x) * P) == x, and mod-down of mod-up(x) alone is x / P rounded (|error| < lp) -- checked
    # on the integers through the CRT of poly 0's first coefficients
    xp = up.astype(object)
    for i, mod in enumerate(q + p):
        xp[:, i] = xp[:, i] * (P % mod) % mod
    back = host(S.moddown(plan, S.BconvPlan(p, q), dev(xp.astype(np.uint64), cuda_device), batch)).reshape(batch, lq, N)
    assert (tq.inverse(back) == x).all()
