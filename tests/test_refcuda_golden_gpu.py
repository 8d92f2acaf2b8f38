"""Second parity pin: our CUDA path against OUTPUTS OF THE REFERENCE'S OWN CUDA CODE.

tests/golden/refcuda_golden.jsonl was produced on a B200 by baseline/_ref/ref_dump (baseline/ref_dump.cu), which
calls the UNMODIFIED reference -- src/core/*.cu built with -rdc for sm_100a by baseline/build_ref.py -- and Phantom's
kernels on closed-form inputs and records, per output array, SHA-256 + leading words + 64 sampled words
(tests/golden/make_golden_refcuda.sh is the generating command).  Here the same inputs go through libmfhe_b200.so
and every integer output must hash to the same digest (bit-exact, whole array); fp64 outputs are compared on the
sampled values with the tolerance written in the test.
"""
import hashlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD_PATH = os.path.join(os.path.dirname(__file__), "golden", "refcuda_golden.jsonl")
MATRIX, POLY = 0, 1


def _load():
    if not os.path.exists(GOLD_PATH):
        return {}
    out = {}
    with open(GOLD_PATH) as f:
        for line in f:
            line = line.strip()
            if line:
                r = json.loads(line)
                out[r["name"]] = r
    return out


GOLD = _load()


def splitmix64(x):
    with np.errstate(over="ignore"):
        z = x.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def synth(seed, total, layout, n, moduli):
    """baseline/ref_dump.cu synth(): splitmix64(seed + idx) mod q_limb(idx)."""
    L = len(moduli)
    idx = np.arange(total, dtype=np.uint64)
    limb = ((idx // np.uint64(n * n)) if layout == MATRIX else (idx // np.uint64(n))) % np.uint64(L)
    q = np.asarray(moduli, dtype=np.uint64)[limb.astype(np.int64)]
    return splitmix64(idx + np.uint64(seed)) % q


def dev(x, device="cuda"):
    import torch
    x = np.ascontiguousarray(x)
    return torch.from_numpy(x.view(np.int64) if x.dtype == np.uint64 else x).to(device)


def host(t):
    a = t.detach().cpu().numpy()
    return a.view(np.uint64) if a.dtype == np.int64 else a


def check(name, t):
    """Whole-array digest equality with the reference's output; on a mismatch say how the sampled words differ."""
    assert name in GOLD, f"fixture has no record {name}"
    g = GOLD[name]
    a = np.ascontiguousarray(host(t)).reshape(-1)
    assert a.size == g["count"], f"{name}: {a.size} words, reference wrote {g['count']}"
    digest = hashlib.sha256(a.tobytes()).hexdigest()
    if digest != g["sha256"]:
        au = a.view(np.uint64)
        bad = [(i, int(au[i]), v) for i, v in g["samples"] if int(au[i]) != v]
        head_bad = [(i, int(au[i]), v) for i, v in enumerate(g["head"]) if int(au[i]) != v]
        raise AssertionError(f"{name}: digest differs from the reference CUDA output; {len(bad)}/64 sampled words differ "
                             f"(first: {bad[:3]}), head mismatches: {head_bad[:3]}")


needs_gold = pytest.mark.skipif(not GOLD, reason="tests/golden/refcuda_golden.jsonl missing")


@pytest.fixture(scope="module")
def G(cuda_device):
    import matrix_fhe_gpu_b200.scheme as S
    g = S.Context()
    assert [int(m) for m in g.moduli] == GOLD["moduli"]["values"]
    return S, g


@needs_gold
def test_fixture_is_from_the_reference_on_b200():
    assert "B200" in GOLD["meta"]["device"] and GOLD["meta"]["sm"] == 100


@needs_gold
def test_samplers_and_keygen_vs_reference_cuda(G):
    """uniform_random_kernel / gaussian_noise_kernel / ternary_secret_kernel (HE.cu:564-627, 690-713) and
    generate_secret_key (HE.cu:1272-1307): every word equal, the fp64 Box-Muller gaussian included."""
    S, g = G
    check("uniform_random", g.sample_uniform())
    check("gaussian_noise", g.sample_gaussian())
    check("ternary_secret", g.sample_ternary())
    check("secret_key", g.generate_secret_key())


@needs_gold
def test_wcrt_vs_reference_cuda(G):
    """wntt_forward_matrix / wntt_inverse_matrix (HE.cu:716-781) at the native shape, whole 176 MiB arrays."""
    S, g = G
    a = g.sample_uniform()
    ev = g.wntt_forward_matrix(a, S.MATRIX, S.POLY)
    check("wcrt_forward_uniform", ev)
    back = g.wntt_inverse_matrix(ev, S.POLY, S.MATRIX)
    check("wcrt_inverse_of_forward", back)
    assert GOLD["wcrt_inverse_of_forward"]["sha256"] == GOLD["uniform_random"]["sha256"]
    x = synth(5005, g.T, POLY, g.n, g.moduli)
    check("wcrt_inverse_synth5005", g.wntt_inverse_matrix(dev(x), S.POLY, S.MATRIX))


@needs_gold
def test_x_transforms_vs_reference_cuda(G):
    """xy_ntt_{forward,backward}_{phantom,gl} and apply_gl_perm (ntt_core.cu:433-481) on 1024 polys x 11 limbs."""
    import matrix_fhe_gpu_b200 as m
    S, g = G
    cnt = 1024
    x = synth(6006, cnt * g.L * g.n, POLY, g.n, g.moduli)
    ph = m.NttPlan(6, g.moduli, m.PHANTOM_NEGACYCLIC)
    gl = m.NttPlan(6, g.moduli, m.GL_XN_MINUS_I)
    d = dev(x)
    ph.forward_(d); check("xy_forward_phantom", d)
    ph.inverse_(d); check("xy_backward_phantom", d)
    assert (host(d) == x).all()
    gl.forward_(d); check("xy_forward_gl", d)
    gl.inverse_(d); check("xy_backward_gl", d)
    check("gl_perm", S.apply_gl_perm(d, g.L, cnt, g.n, False))
    check("gl_perm_inverse", S.apply_gl_perm(d, g.L, cnt, g.n, True))


@needs_gold
def test_encrypt_decrypt_ct_ops_vs_reference_cuda(G):
    """encrypt_pair / encrypt (HE.cu:1370-1552), decrypt_to_eval_matrix (:1603-1617), add_ciphertexts and
    multiply_ciphertexts_raw (:1710-1740): all ciphertext words of the native shape (4 x 176 MiB per pair)."""
    S, g = G
    sk = g.generate_secret_key()
    m_re = dev(synth(1001, g.T, MATRIX, g.n, g.moduli))
    m_im = dev(synth(2002, g.T, MATRIX, g.n, g.moduli))
    cr, ci = g.encrypt_pair(m_re, m_im, sk)
    check("encrypt_pair_ct_re", cr)
    check("encrypt_pair_ct_im", ci)
    check("decrypt_to_eval_matrix_re", g.decrypt_to_eval(cr, sk, S.MATRIX))
    check("encrypt_single_ct", g.encrypt(m_re, sk))
    check("add_ciphertexts", g.add_ciphertexts(cr, ci, reference_indexing=True))
    d0, d1, d2 = g.multiply_ciphertexts_raw(cr, ci, reference_indexing=True)
    check("mul_raw_d0", d0); check("mul_raw_d1", d1); check("mul_raw_d2", d2)


@needs_gold
def test_centered_wcrt_reference_is_broken_in_every_build(G):
    """wntt_forward_centered / wntt_inverse_centered (HE.cu:1029-1114), the pattern of test_wcrt_roundtrip.cu:34-72.

    What the reference's OWN code produces on a B200, recorded in the fixture: in the stock build the kernel reads
    zero-filled copies of d_crt_* (SURVEY fact 8); in the -rdc build the constants are one instance, but the reference's
    upload of d_crt_M fails (cudaMemcpyToSymbol -> cudaErrorInvalidValue, never checked; the fixture's probe_constants
    record shows d_crt_M[0] == 0 next to correct d_crt_Q / d_crt_inv), so every composed value is 0: the forward output is
    all zeros, the inverse of zeros is zeros, and the driver prints max_err=8, worst_idx=0 -- its own pass rule
    (max_err == 0) fails on the unmodified reference.  There is therefore no reference output to match for this row;
    we pin (a) those facts about the reference, and (b) our kernel to the oracle's restatement of the kernel WITH correct
    tables (tests/test_scheme_gpu.py::test_wcrt_vector_centered_layout, test_dropin_gpu.py)."""
    S, g = G
    zeros = hashlib.sha256(bytes(8 * g.phi * g.n2)).hexdigest()
    assert GOLD["wcrt_forward_centered"]["sha256"] == zeros and GOLD["wcrt_inverse_centered"]["sha256"] == zeros
    assert GOLD["probe_constants"]["d_crt_M0"] == 0 and GOLD["probe_constants"]["d_crt_Q0"] != 0
    assert (GOLD["wcrt_roundtrip_driver"]["max_err"], GOLD["wcrt_roundtrip_driver"]["worst_idx"]) == (8, 0)
    # ours, with correct tables: the composed value of 11 limbs of eval residues is ~385 bits wide, so the int64
    # conversion saturates (he_big_to_i64_checked, HE.cu:904-915) -- equal to the oracle's restatement
    import oracle as O
    w, y, x = np.meshgrid(np.arange(g.phi), np.arange(g.n), np.arange(g.n), indexing="ij")
    v = ((w + x + y) % 17 - 8).astype(np.int64).reshape(g.phi, g.n2)
    ev = host(g.wntt_forward_centered(dev(v.reshape(-1)))).view(np.int64).reshape(g.phi, g.n2)
    V = O.SchemeCtx().wntt()                                    # oracle tables (pinned to the reference's constants)
    mods = [int(q) for q in g.moduli]
    Q = 1
    for q in mods:
        Q *= q
    rng = np.random.default_rng(5)
    for _ in range(6):                                          # exact integers: kernel definition with correct tables
        wo, pos = int(rng.integers(g.phi)), int(rng.integers(g.n2))
        val = 0
        for l, q in enumerate(mods):
            acc = sum((int(v[r, pos]) % q) * int(V[l, wo, r]) for r in range(g.phi)) % q
            M = Q // q
            val += M * (acc * pow(M % q, -1, q) % q)
        val %= Q
        neg = val > Q // 2
        mag = Q - val if neg else val
        want = (np.iinfo(np.int64).min if neg else np.iinfo(np.int64).max) if mag > np.iinfo(np.int64).max else (-mag if neg else mag)
        assert int(ev[wo, pos]) == want
    sat = np.isin(ev, [np.iinfo(np.int64).max, np.iinfo(np.int64).min]).mean()
    print(f"centred forward: {sat:.3f} of the outputs saturate int64 (the composed value does not fit)")


@needs_gold
def test_trace_step_vs_reference_cuda(G):
    """map_B_to_Bprime_batched / trace_gemm_batched / rescale_by_delta_batched (batched_trace.cu:37-197) on 16 lanes and
    the single-lane trace.cu:30-161 forms."""
    S, g = G
    batch = 16
    tot = batch * g.L * g.n2
    m = [dev(synth(3001 + k, tot, MATRIX, g.n, g.moduli)) for k in range(4)]
    bpr, bpi = g.map_B_to_Bprime(m[2], m[3], batch)
    check("trace_map_batched_re", bpr); check("trace_map_batched_im", bpi)
    cr, ci = g.trace_gemm(m[0], m[1], bpr, bpi, batch)
    check("trace_gemm_batched_re", cr); check("trace_gemm_batched_im", ci)
    cr, ci = g.rescale_by_delta(cr, ci, 123456789, 987654321, 555555555, batch)
    check("trace_rescale_batched_re", cr); check("trace_rescale_batched_im", ci)
    one = g.L * g.n2
    bpr1, bpi1 = g.map_B_to_Bprime(m[2][:one].clone(), m[3][:one].clone(), 1)
    c1r, c1i = g.trace_gemm(m[0][:one].clone(), m[1][:one].clone(), bpr1, bpi1, 1)
    check("trace_gemm_single_re", c1r); check("trace_gemm_single_im", c1i)


@needs_gold
def test_fp64_pipeline_vs_reference_cuda(G):
    """encode -> encrypt_pair -> decrypt_and_decode and the zero-key encode -> decode loop on the e2e driver's message:
    our decoded slots against the REFERENCE's decoded slots at the 64 sampled positions.  Tolerance: both pipelines
    carry the same gaussian error (the sampler words are equal, test above), so the outputs differ only by fp64
    round-off in encode (+-1 ulp of the 2^35 quantiser) and decode: |ours - reference| <= 2^-20 * max|m| (the north-star
    bound); our own max error against the message must also pass the drivers' 1e-3 like the reference's did."""
    S, g = G
    ell, i = np.meshgrid(np.arange(g.phi, dtype=np.float64), np.arange(g.n2, dtype=np.float64), indexing="ij")
    v = ell + 0.001 * i
    msg = (v - 1j * v).reshape(-1)
    sk = g.generate_secret_key()
    re, im = g.encode_to_wntt_eval(dev(msg))
    cr, ci = g.encrypt_pair(re, im, sk)
    out = host(g.decrypt_and_decode(cr, ci, sk))
    import torch
    zr, zi = g.allocate_ciphertext(), g.allocate_ciphertext()
    zr[: g.T] = re; zi[: g.T] = im
    out0 = host(g.decrypt_and_decode(zr, zi, torch.zeros_like(sk)))
    tol = 2.0 ** -20 * np.abs(msg).max()
    # the reference's decode only works in its stock build (under -rdc its d_crt_M upload fails and decode returns zeros:
    # the rdc records carry max_err 7.3e2), so the fp64 records compared here are the stock_* ones
    assert GOLD["decrypt_and_decode_e2e"]["max_err"] > 100 and GOLD["stock_decrypt_and_decode_e2e"]["max_err"] < 1e-5
    for name, o in (("stock_decrypt_and_decode_e2e", out), ("stock_decode_zero_key_e2e", out0)):
        r = GOLD[name]
        assert r["max_err"] < 1e-3 and np.abs(o - msg).max() < 1e-3
        worst = max(abs(o[idx] - complex(a, b)) for idx, a, b in r["samples"])
        print(f"{name}: reference max_err {r['max_err']:.3e}, ours {np.abs(o - msg).max():.3e}, |ours - reference| <= {worst:.3e}")
        assert worst <= tol
    # encode output words: the reference's fp64 summation order differs from ours, so words may differ where
    # llround(z * 2^35) falls on the other side of .5; report the agreement, require the decode bound above
    for name, t in (("stock_encode_e2e_re", re), ("stock_encode_e2e_im", im)):
        a = host(t)
        same = sum(int(a[i_]) == v_ for i_, v_ in GOLD[name]["samples"])
        print(f"{name}: {same}/64 sampled words equal to the reference's")


# ---------------------------------------------------------------- Phantom kernels (configs 2 and 5) ----
@needs_gold
def test_flat_ntt_vs_phantom_cuda(cuda_device):
    """nwt_2d_radix8_forward_inplace / backward_inplace (ph/src/ntt/fntt_2d.cu:620-653, intt_2d.cu:724-757) at config 2
    (N = 2^16, 32 x 60-bit), polys 0 and 63 of the bench's input: forward words equal, round trip equal to the input."""
    import matrix_fhe_gpu_b200 as m
    mods = GOLD["phantom_cfg2_moduli"]["values"]
    N, L = 1 << 16, 32
    plan = m.NttPlan(16, mods)
    q = np.asarray(mods, dtype=np.uint64).reshape(L, 1)
    for b in (0, 63):
        idx = np.arange(L * N, dtype=np.uint64).reshape(L, N) + np.uint64(42 + b * L * N)
        x = splitmix64(idx) % q
        d = dev(x.reshape(1, L, N))
        plan.forward_(d)
        check(f"phantom_cfg2_forward_poly{b}", d)
        plan.inverse_(d)
        check(f"phantom_cfg2_roundtrip_poly{b}", d)
        assert (host(d).reshape(L, N) == x).all()


@needs_gold
def test_bconv_step_vs_phantom_cuda(cuda_device):
    """bConv_BEHZ (ph/src/rns_bconv.cu:212-229) and the whole config-5 step INTT(48) -> bConv 48->16 -> NTT(16) at
    N = 2^17 on poly 0."""
    import matrix_fhe_gpu_b200 as m
    import matrix_fhe_gpu_b200.scheme as S
    mods = GOLD["phantom_cfg5_moduli"]["values"]
    assert GOLD["phantom_cfg5_primes_support_2N"]["value"] is True
    N, LI, LO = 1 << 17, 48, 16
    qi, pj = mods[:LI], mods[LI:]
    q = np.asarray(qi, dtype=np.uint64).reshape(LI, 1)
    idx = np.arange(LI * N, dtype=np.uint64).reshape(LI, N) + np.uint64(42)
    x = splitmix64(idx) % q
    bc = S.BconvPlan(qi, pj)
    d = dev(x.reshape(1, LI, N))
    check("phantom_cfg5_bconv_poly0", bc.convert(d, N, 1))
    plan = m.NttPlan(17, mods)
    plan.inverse_limbs_(d, 0, LI)
    check("phantom_cfg5_intt_poly0", d)
    o = bc.convert(d, N, 1)
    check("phantom_cfg5_intt_bconv_poly0", o)
    plan.forward_limbs_(o, LI, LO)
    check("phantom_cfg5_step_poly0", o)
