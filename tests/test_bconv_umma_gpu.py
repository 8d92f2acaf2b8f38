"""Base conversion on the tensor cores (csrc/bconv_umma.cuh: 8-bit digit GEMM, tcgen05 kind::i8, TMEM accumulators) against
the CPU oracle (Phantom's bConv_BEHZ restated) and against the CUDA-core kernels (MFHE_BCONV_PATH=c), word for word:
modulus widths from 30 to 60 bits on either side (4..8 digits), input windows of 1..48 limbs (K padded to 32 / 64),
output bases that are not a multiple of the 16-limb tile, full / partial / scatter forms, extreme residues."""
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu


def dev(x, device):
    import torch
    return torch.from_numpy(np.ascontiguousarray(x).view(np.int64)).to(device)


def host(t):
    return t.cpu().numpy().view(np.uint64)


def primes(n, bits, count, skip=0):
    return [int(v) for v in O.get_primes(n, bits, count + skip)][skip:]


@pytest.mark.parametrize("logn,qbits,l_in,pbits,l_out,batch", [
    (12, 55, 48, 55, 16, 2),        # config 5's bases
    (12, 60, 7, 60, 3, 3),          # 8 digits each side, 128-bit Barrett
    (12, 30, 5, 45, 17, 2),         # 4 x 6 digits, two output tiles (17 limbs)
    (13, 50, 33, 35, 1, 1),         # K = 64 with 31 padded limbs, a single output limb
    (12, 45, 1, 55, 4, 2),          # a single input limb
    (12, 55, 40, 60, 40, 1),        # three output tiles, K = 64
])
def test_full_conversion_vs_oracle_and_cuda_cores(cuda_device, monkeypatch, logn, qbits, l_in, pbits, l_out, batch):
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    N = 1 << logn
    q = primes(2 * N, qbits, l_in)
    p = primes(2 * N, pbits, l_out, skip=l_in if pbits == qbits else 0)
    x = O.synth_polys(9, batch, q, N)
    qa = np.asarray(q, dtype=np.uint64)
    x[0, :, :64] = 0
    x[0, :, 64:128] = (qa - np.uint64(1))[:, None]
    bc = S.BconvPlan(q, p)
    d = dev(x, cuda_device)
    monkeypatch.setenv("MFHE_BCONV_PATH", "t")            # the tensor-core kernel for every window (default: > 24 input limbs)
    got = bc.convert(d, N, batch)
    want = np.stack([O.bconv(x[b], q, p) for b in range(batch)])
    assert (host(got).reshape(batch, l_out, N) == want).all()
    monkeypatch.setenv("MFHE_BCONV_PATH", "c")
    ref = bc.convert(d, N, batch)
    monkeypatch.setenv("MFHE_BCONV_PATH", "t")
    assert torch.equal(got, ref)
    for _ in range(3):                                   # persistent CTAs, TMEM reuse: same words every time
        assert torch.equal(bc.convert(d, N, batch), ref)
    monkeypatch.delenv("MFHE_BCONV_PATH")
    assert torch.equal(bc.convert(d, N, batch), ref)     # whatever the default picks


@pytest.mark.parametrize("in_start,in_count", [(0, 6), (6, 6), (42, 6), (0, 24), (24, 24), (5, 11), (47, 1), (0, 48)])
def test_partial_and_scatter_windows(cuda_device, monkeypatch, in_start, in_count):
    """The limb-sharded forms: partial sums over a window of input limbs (reduced mod p_j), plain and with one destination
    per output limb; both kernels, same words; the windows of config 5 at 8 / 2 / 1 ranks and odd ones."""
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    logn, l_in, l_out, batch = 12, 48, 16, 3
    N = 1 << logn
    pr = primes(2 * N, 55, l_in + l_out)
    q, p = pr[:l_in], pr[l_in:]
    bc = S.BconvPlan(q, p)
    x = O.synth_polys(4, batch, q[in_start:in_start + in_count], N)
    d = dev(x, cuda_device)
    monkeypatch.setenv("MFHE_BCONV_PATH", "t")
    got = bc.partial(d, in_start, in_count, N, batch)
    sc = torch.full((batch * l_out * N + 64,), -1, dtype=torch.int64, device=cuda_device)
    bc.partial_scatter(d, in_start, in_count, [sc.data_ptr() + 8 * j * N for j in range(l_out)], [l_out * N] * l_out, N, batch)
    monkeypatch.setenv("MFHE_BCONV_PATH", "c")
    ref = bc.partial(d, in_start, in_count, N, batch)
    monkeypatch.delenv("MFHE_BCONV_PATH")
    assert torch.equal(got, ref)
    assert torch.equal(sc[:batch * l_out * N], ref) and bool((sc[batch * l_out * N:] == -1).all())
    # and the meaning: the window's contribution  sum_{i in window} y_i * (Q/q_i mod p_j)  mod p_j  with exact integers
    Q = 1
    for v in q:
        Q *= v
    c0 = [int(v) for v in x[0, :, 5]]
    for j in (0, l_out - 1):
        acc = 0
        for k, xi in enumerate(c0):
            qi = q[in_start + k]
            y = xi * pow((Q // qi) % qi, -1, qi) % qi
            acc += y * ((Q // qi) % p[j])
        assert int(host(got).reshape(batch, l_out, N)[0, j, 5]) == acc % p[j]
