"""Randomised parity sweep of the kernels added in round 2 against the CPU oracle / the alternative kernels (not a pytest:
run by hand on a GPU box, result recorded under profiles/).   python tools/fuzz_r2.py [seconds] [seed]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import matrix_fhe_gpu_b200 as m
import matrix_fhe_gpu_b200.scheme as S
import oracle as O

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
dev = torch.device("cuda", 0)


def to_dev(x):
    return torch.from_numpy(np.ascontiguousarray(x).view(np.int64)).to(dev)


def host(t):
    return t.cpu().numpy().view(np.uint64)


counts = {"fp64_ntt": 0, "bconv_umma": 0, "small_fp64": 0, "matmul": 0}
t_end = time.time() + budget
while time.time() < t_end:
    kind = rng.integers(4)
    if kind == 0:                                       # flat NTT on the FP64 pipe vs oracle and integer kernels
        logn = int(rng.integers(12, 16)); n = 1 << logn
        bits = int(rng.integers(30, 51)); limbs = int(rng.integers(1, 5)); batch = int(rng.integers(1, 6))
        try:
            mods = O.coeff_modulus_create(n, bits, limbs)
        except ValueError:
            continue
        plan = m.NttPlan(logn, mods)
        assert plan.arith == 1
        x = O.synth_polys(int(rng.integers(1 << 30)), batch, mods, n)
        q = np.asarray(mods, dtype=np.uint64).reshape(1, -1, 1)
        lazy = x + q * (O.splitmix64(np.arange(x.size, dtype=np.uint64)).reshape(x.shape) & np.uint64(1))
        ref = O.PhantomTables(logn, mods).forward(x)
        d = to_dev(lazy); plan.forward_(d)
        assert (host(d) == ref).all(), ("fp64 fwd", logn, bits, limbs, batch)
        plan.inverse_(d)
        assert (host(d) == x).all(), ("fp64 inv", logn, bits, limbs, batch)
        counts["fp64_ntt"] += 1
    elif kind == 1:                                     # tensor-core base conversion vs oracle
        logn = 12; n = 1 << logn
        qb, pb = int(rng.integers(30, 61)), int(rng.integers(30, 61))
        l_in, l_out, batch = int(rng.integers(1, 40)), int(rng.integers(1, 36)), int(rng.integers(1, 3))
        try:
            q = [int(v) for v in O.get_primes(2 * n, qb, l_in)]
            p = [int(v) for v in O.get_primes(2 * n, pb, l_out + (l_in if pb == qb else 0))][-l_out:]
        except ValueError:
            continue
        if set(q) & set(p):
            continue
        try:
            bc = S.BconvPlan(q, p)
        except m.MfheError:
            continue                                    # l_in * q * p >= 2^128
        x = O.synth_polys(int(rng.integers(1 << 30)), batch, q, n)
        os.environ["MFHE_BCONV_PATH"] = "t"
        got = host(bc.convert(to_dev(x), n, batch)).reshape(batch, l_out, n)
        del os.environ["MFHE_BCONV_PATH"]
        want = np.stack([O.bconv(x[b], q, p) for b in range(batch)])
        assert (got == want).all(), ("bconv", qb, pb, l_in, l_out)
        counts["bconv_umma"] += 1
    elif kind == 2:                                     # n-point transforms on the FP64 pipe, all conventions
        logn = int(rng.integers(4, 9)); n = 1 << logn
        conv = int(rng.integers(3)); bits = int(rng.integers(28, 51)); L = int(rng.integers(1, 4))
        try:
            mods = O.get_primes(2 * n, bits, L)
        except ValueError:
            continue
        plan = m.NttPlan(logn, mods, conv)
        x = O.synth_polys(int(rng.integers(1 << 30)), 9, mods, n)
        outs = []
        for arith in (1, 0):
            plan.set_arith(arith)
            d = to_dev(x); plan.forward_(d); outs.append(host(d).copy()); plan.inverse_(d)
            assert (host(d) == x).all(), ("small inv", logn, conv, bits, arith)
        assert (outs[0] == outs[1]).all(), ("small fwd", logn, conv, bits)
        counts["small_fp64"] += 1
    else:                                               # tcgen05 matrix product + fused decrypt vs the alternative kernels
        bits = int(rng.integers(26, 49)); L = int(rng.integers(1, 4)); f = int(rng.choice([5, 17]))
        mods = O.primes_congruent_one(2 * 64 * 3 * f, bits, L)
        g = S.Context(64, f, mods, 2.0 ** 30)
        qv = to_dev(np.repeat(np.asarray(mods, dtype=np.uint64), g.n2)[None].repeat(g.phi, 0).reshape(-1))
        a = g.sample_uniform()
        cts = [torch.cat([(a * (3 + k) + k) % qv, (a * (7 + k) + 1) % qv]) for k in range(4)]
        sk = g.generate_secret_key()
        prod = g.ct_matmul(*cts)
        dec = [t.clone() for t in g.ct_matmul_decrypt(prod.clone(), sk)]
        os.environ["MFHE_MATMUL_PATH"] = "m"; os.environ["MFHE_MATMUL_DECRYPT_PATH"] = "r"
        prod2 = g.ct_matmul(*cts)
        dec2 = g.ct_matmul_decrypt(prod2.clone(), sk)
        del os.environ["MFHE_MATMUL_PATH"], os.environ["MFHE_MATMUL_DECRYPT_PATH"]
        assert torch.equal(prod, prod2) and torch.equal(dec[0], dec2[0]) and torch.equal(dec[1], dec2[1]), ("matmul", bits, L, f)
        counts["matmul"] += 1
torch.cuda.synchronize()
print("fuzz ok:", counts)
