"""FP64-pipe arithmetic of the flat NTT kernels (csrc/modmath.cuh, q < 2^50) checked on the CPU: IEEE fma / add / mul
are bit-identical on the host and on the GPU, so the exactness of the modular product, the fold, and the kernels'
compile-time bound schedule are verified here by running the SAME expressions in the SAME stage order (oracle/ntt_oracle.c)
against the integer oracle (Phantom's butterflies restated).  The GPU parity tests then check the kernels word for word
(tests/test_ntt_fp64_gpu.py)."""
import numpy as np
import pytest

import oracle as O

# the largest NTT-friendly primes below 2^50 and two smaller widths (the matrix scheme's 45- and 35-bit moduli)
PRIMES = [int(v) for v in O.get_primes(1 << 17, 50, 2)] + [17592186435073, 17182765057, 786433]


@pytest.mark.parametrize("q", PRIMES)
def test_modmul_exact_and_bounded(q):
    for bound16, limit in ((16, 0.5 + 1 / 16 + 1e-9), (32, 0.5 + 2 / 16 + 1e-9), (64, 0.5 + 4 / 16 + 1e-9)):
        bad, mx = O.fp64_modmul_check(q, 2_000_000, 7 + bound16, bound16)
        assert bad == 0
        assert mx <= limit, (bound16, mx)          # |t| <= q (1/2 + B/16): the schedule's fp_t


@pytest.mark.parametrize("logn", [4, 6, 8, 12, 14, 16, 17])      # 4..8: the n-point networks of the matrix scheme
def test_networks_equal_integer_oracle_and_keep_their_bounds(logn):
    n = 1 << logn
    mods = [int(v) for v in O.get_primes(2 * n, 50, 2)] + [int(O.get_primes(2 * n, 45, 1)[0]), int(O.get_primes(2 * n, 30, 1)[0])]
    if logn == 6:
        mods += [17592186435073, 17182765057]                  # the reference's 45- and 35-bit moduli
    assert max(mods) < 2 ** 50
    t = O.PhantomTables(logn, mods)
    rng = np.random.default_rng(logn)
    for l, q in enumerate(mods):
        cases = [rng.integers(0, q, size=n, dtype=np.uint64),
                 np.full(n, q - 1, dtype=np.uint64),
                 rng.integers(0, 2 * q, size=n, dtype=np.uint64),           # lazy inputs below 2q
                 np.full(n, 2 * q - 1, dtype=np.uint64)]
        for x in cases:
            xin = (x % np.uint64(q))
            ref = xin.copy()
            O.lib().mfo_negacyclic_forward(ref, logn, q, t.tw[l], t.tws[l])
            got, rc, st = O.fp64_ntt(x, logn, q, t.tw[l])
            assert rc == 0, (q, rc, st)
            assert (got == ref).all()
            back = ref.copy()
            O.lib().mfo_negacyclic_inverse(back, logn, q, t.itw[l], t.itws[l], int(t.n_inv[l]), int(t.n_inv_shoup[l]))
            assert (back == xin).all()
            y = ref + (x - xin)                                              # same laziness on the way back
            got, rc, st = O.fp64_ntt(y, logn, q, t.itw[l], inverse=True, n_inv=int(t.n_inv[l]))
            assert rc == 0, (q, rc, st)
            assert (got == xin).all()


def test_schedule_is_the_one_in_the_kernels():
    """The fold stages the C emulation derives are the ones modmath.cuh derives (same three rules); spelled out for
    N = 2^16 so that a change of either side shows up here."""
    def fp_t(B): return 8 + (B + 15) // 16
    B, folds = 32, []
    for s in range(16):
        f = B + fp_t(B) > 64 and s != 15
        folds.append(f)
        B = (9 if f else B) + fp_t(B)
    assert [i for i, f in enumerate(folds) if f] == [2, 7, 12]
    B, folds = 32, []
    for i in range(16):
        f = 2 * B > 32
        assert 2 * B <= 64
        folds.append(f)
        B = max(9 if f else 2 * B, fp_t(2 * B))
    assert [i for i, f in enumerate(folds) if f][:4] == [0, 2, 4, 6]
