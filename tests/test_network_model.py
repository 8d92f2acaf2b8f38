"""CPU check of the kernels' index arithmetic (tests/network_model.py) against the oracle.

The model mirrors csrc/ntt_flat.cu / ntt_small.cu pass by pass; if it reproduces the oracle,
the twiddle indexing, pass decomposition and lazy-bound schedule of the CUDA code are right.
"""
import numpy as np
import pytest

import network_model as M
import oracle as O

REF_Q0 = 17592186435073


@pytest.mark.parametrize("logn", [12, 13])
def test_flat_network_matches_oracle(logn):
    q = int(O.get_primes(1 << logn, 60, 1)[0])
    n = 1 << logn
    t = O.PhantomTables(logn, [q])
    tw, itw, ninv, w1n = M.tree_tables(logn, q, t.psi(0), 2 * n, n)
    assert tw == [int(x) for x in t.tw[0]]          # product-tree builder == Phantom layout
    net = M.Net(logn, q, tw, itw, ninv, w1n)
    x = O.synth_polys(3, 1, [q], n)
    got = M.flat_forward(net, [int(v) for v in x.reshape(-1)])
    assert got == [int(v) for v in t.forward(x).reshape(-1)]
    assert net.max_bound <= 16
    assert M.flat_inverse(net, got) == [int(v) for v in x.reshape(-1)]


@pytest.mark.parametrize("logn", [14, 15, 16, 17])
def test_lazy_bound_schedule_never_overflows(logn):
    B = M.K_FWD_IN_BOUND
    for s in range(logn):
        assert M.fwd_bound_before(s, logn) == B
        T = min(M.fwd_xtarget(B, s, logn), B)
        B = T + (2 if s == logn - 1 else 4)      # approximate quotient: t < 4q; exact (t < 2q) in the last stage
        assert B <= 16
    assert B <= 4      # the row kernel finishes with csub(2q), csub(q)


def test_swizzle_is_a_bijection_and_conflict_free():
    assert sorted(M.swz(e) for e in range(256)) == list(range(256))
    # 8-byte accesses: a half-warp must touch 16 distinct 8-byte bank pairs (mod 16)
    for j in range(8):
        for half in (0, 16):
            a = {M.swz((j << 5) + lane) % 16 for lane in range(half, half + 16)}
            assert len(a) == 16
            b = {M.swz((((lane >> 2) << 5) | (lane & 3)) + (j << 2)) % 16 for lane in range(half, half + 16)}
            assert len(b) == 16
    # 16-byte accesses: a quarter-warp must touch 8 distinct 16-byte chunks (mod 8)
    for k in range(4):
        for qw in range(0, 32, 8):
            c = {(M.swz((lane << 3) + 2 * k) >> 1) % 8 for lane in range(qw, qw + 8)}
            assert len(c) == 8


@pytest.mark.parametrize("logn", [4, 5, 6, 7, 8])
def test_small_network_all_conventions(logn):
    n = 1 << logn
    q = REF_Q0 if logn <= 6 else int(O.get_primes(n, 50, 1)[0])
    L = O.lib()
    t = O.PhantomTables(logn, [q])
    x = O.synth_polys(5, 1, [q], n)
    xs = [int(v) for v in x.reshape(-1)]
    net = M.Net(logn, q, *M.tree_tables(logn, q, t.psi(0), 2 * n, n))
    got = M.small_forward(net, xs, False)
    assert got == [int(v) for v in t.forward(x).reshape(-1)]
    assert M.small_inverse(net, got, False) == xs
    if (q - 1) % (4 * n) == 0:
        psi4 = L.mfo_gl_get_psi(q, n)
        for c0, f in ((n, L.mfo_gl_forward), (0, L.mfo_cyclic_forward)):
            net = M.Net(logn, q, *M.tree_tables(logn, q, psi4, 4 * n, c0))
            y = x.reshape(-1).copy()
            f(y, n, q)
            got = M.small_forward(net, xs, True)
            assert got == [int(v) for v in y]
            assert M.small_inverse(net, got, True) == xs
