"""Integer model of the base-conversion kernel's reduction (matrix_fhe_gpu_b200/csrc/bconv.cu, bconv_fold<W1=true>):
the accumulated sum v = sum_i y_i c_i (y_i < q_i, c_i < p_j) is reduced with ONE 64x64 high product when the host
check `in_count * max q <= 2^61` holds.  This test replays that arithmetic in Python integers at the extremes of the
admitted range, so the bound argument in the kernel comment (qhat in [floor(v/p) - 2, floor(v/p)], r < 3p < 2^64) is
checked independently of a GPU; the GPU-side parity of the kernel itself is tests/test_scheme_gpu.py::test_bconv_*."""
import random

import pytest

import oracle as O

M64 = (1 << 64) - 1


def fold_one_word(v, p):
    """bconv_fold<true>: make_barrett1 + one-word Barrett on the 128-bit value v, 64-bit wrap-around as on the device."""
    k = p.bit_length()
    mu = (1 << 128) // p
    mu_lo, mu_hi = mu & M64, mu >> 64
    mu1 = ((mu_hi << (k - 2)) | (mu_lo >> (66 - k))) & M64
    assert mu1 == (1 << (k + 62)) // p                       # what modmath.cuh's make_barrett1 claims to build
    lo, hi, sh = v & M64, v >> 64, k - 2
    x1 = ((lo >> sh) | (hi << (64 - sh))) & M64
    assert x1 == v >> sh                                      # no bits lost: v < 2^(k+62)
    qhat = (x1 * mu1) >> 64
    r = (lo - qhat * p) & M64
    assert v // p - 2 <= qhat <= v // p
    for m in (2 * p, p):
        if r >= m:
            r -= m
    return r


@pytest.mark.parametrize("bits_q,bits_p,in_count", [(55, 55, 48), (58, 60, 8), (55, 60, 64), (36, 45, 128), (60, 60, 2),
                                                    (61, 36, 1), (50, 61, 24)])
def test_one_word_barrett_covers_the_admitted_range(bits_q, bits_p, in_count):
    qs = [int(v) for v in O.get_primes(1024, bits_q, 2)]
    ps = [int(v) for v in O.get_primes(1024, bits_p, 2)] if bits_p != bits_q else qs
    qmax = max(qs)
    if bits_q == 61:                                          # get_primes stays below 2^bits: model the admitted maximum
        qmax = (1 << 61) - 1
    assert in_count * qmax <= 1 << 61                         # the host-side condition for the one-word form
    rng = random.Random(7)
    for p in ps:
        worst = in_count * (qmax - 1) * (p - 1)
        assert worst < 1 << (p.bit_length() + 61)
        for v in [0, 1, p - 1, p, worst, worst - 1, (worst // p) * p, (worst // p) * p - 1] + \
                 [rng.randrange(worst + 1) for _ in range(2000)]:
            assert fold_one_word(v, p) == v % p


def test_one_word_form_is_refused_outside_the_range():
    """9 x 58-bit exceeds 2^61: the kernel must take the 128-bit form there (host check in bconv_run)."""
    q = int(O.get_primes(1024, 58, 1)[0])
    assert 9 * q > 1 << 61 and 8 * q <= 1 << 61
