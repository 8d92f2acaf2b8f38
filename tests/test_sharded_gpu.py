"""GPU parity of the limb-sharded INTT -> base conversion -> NTT step (BASELINE config 5).

One GPU: the ranks are emulated in ONE process (each rank's window is computed in turn and the
all-to-all is replaced by its single-process stand-in) and compared bit-exactly with the oracle's
unsharded pipeline.  With >= 2 GPUs the same step also runs under torchrun over NCCL
(tools/run_sharded.py --check)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _oracle_step(x, logn, q, p):
    """x [batch][L_in][N] NTT form -> [batch][L_out][N] NTT form in base P."""
    coeff = O.PhantomTables(logn, q).inverse(x)
    conv = np.stack([O.bconv(coeff[b], q, p) for b in range(x.shape[0])])
    return O.PhantomTables(logn, p).forward(conv)


@pytest.mark.parametrize("logn,l_in,l_out,bits,world,batch", [(12, 6, 4, 55, 4, 2), (13, 5, 3, 60, 2, 1), (12, 8, 8, 50, 8, 1)])
def test_limb_sharded_step_emulated(cuda_device, logn, l_in, l_out, bits, world, batch):
    import torch
    from matrix_fhe_gpu_b200 import sharded as S
    N = 1 << logn
    primes = [int(v) for v in O.get_primes(N, bits, l_in + l_out)]
    q, p = primes[:l_in], primes[l_in:]
    x = O.synth_polys(11, batch, q, N)
    want = _oracle_step(x, logn, q, p)
    ranks = [S.LimbShardedBconv(logn, q, p, r, world) for r in range(world)]
    partials = []
    for r, sh in enumerate(ranks):
        s, c = sh.q_windows[r]
        if c == 0:
            partials.append(torch.zeros(batch, l_out, N, dtype=torch.int64, device=cuda_device))
            continue
        xl = torch.from_numpy(np.ascontiguousarray(x[:, s:s + c]).view(np.int64)).to(cuda_device)
        partials.append(sh.local_partial(xl, batch))
    for r, sh in enumerate(ranks):
        s, c = sh.p_windows[r]
        if c == 0:
            continue
        parts = S.emulate_exchange(partials, sh.p_windows, r)
        got = sh.finish(parts, batch).cpu().numpy().view(np.uint64)
        assert (got == want[:, s:s + c]).all()


@pytest.mark.parametrize("logn,l_in,l_out,bits,world,batch", [(12, 6, 4, 55, 4, 2), (12, 5, 3, 60, 2, 3)])
def test_partial_scatter_matches_partial_plus_exchange(cuda_device, logn, l_in, l_out, bits, world, batch):
    """mfhe_bconv_partial_scatter (fused partial + exchange) against mfhe_bconv_partial followed by the exchange:
    the 'peer' receive buffers are plain local tensors here; under torchrun they are symmetric-memory peers."""
    import torch
    from matrix_fhe_gpu_b200 import sharded as S
    N = 1 << logn
    primes = [int(v) for v in O.get_primes(N, bits, l_in + l_out)]
    q, p = primes[:l_in], primes[l_in:]
    x = O.synth_polys(5, batch, q, N)
    ranks = [S.LimbShardedBconv(logn, q, p, r, world) for r in range(world)]
    wins = ranks[0].p_windows
    cmax = max(c for _, c in wins)
    bufs = [torch.full((world * batch * cmax * N,), -1, dtype=torch.int64, device=cuda_device) for _ in range(world)]
    bases = [b.data_ptr() for b in bufs]
    partials = []
    for r, sh in enumerate(ranks):
        s, c = sh.q_windows[r]
        xl = torch.from_numpy(np.ascontiguousarray(x[:, s:s + c]).view(np.int64)).to(cuda_device)
        sh.plan_q.inverse_limbs_(xl, s, c)
        partials.append(sh.bconv.partial(xl, s, c, N, batch).view(batch, l_out, N))
        ptrs, strides = S.scatter_table(bases, wins, r, batch, N)
        sh.bconv.partial_scatter(xl, s, c, ptrs, strides, N, batch)
    torch.cuda.synchronize()
    for r, (s, c) in enumerate(wins):
        want = S.emulate_exchange(partials, wins, r)
        got = bufs[r][: world * batch * c * N].view(world, batch, c, N)
        assert torch.equal(got, want)
        assert bool((bufs[r][world * batch * c * N:] == -1).all())         # nothing written past the window


@pytest.mark.parametrize("logn,l_in,l_out,bits,batch", [(12, 48, 4, 55, 2), (12, 32, 3, 55, 1), (12, 7, 3, 58, 2)])
def test_unsharded_step_windows(cuda_device, monkeypatch, logn, l_in, l_out, bits, batch):
    """world = 1: a wide base is converted by one mfhe_bconv on the tensor cores (bconv_umma.cuh); with MFHE_BCONV_PATH=c
    window by window on the register-resident CUDA-core forms (48 = 2 x 24, 32 = 2 x 16) + mfhe_bconv_reduce, a base
    without such a split (7) through one mfhe_bconv: all equal the oracle."""
    import torch
    from matrix_fhe_gpu_b200 import sharded as S
    N = 1 << logn
    primes = [int(v) for v in O.get_primes(N, bits, l_in + l_out)]
    q, p = primes[:l_in], primes[l_in:]
    x = O.synth_polys(17, batch, q, N)
    sh = S.LimbShardedBconv(logn, q, p, 0, 1)
    got = sh.step_unsharded(torch.from_numpy(x.view(np.int64).copy()).to(cuda_device), batch)
    assert (got.cpu().numpy().view(np.uint64) == _oracle_step(x, logn, q, p)).all()
    monkeypatch.setenv("MFHE_BCONV_PATH", "c")
    got = sh.step_unsharded(torch.from_numpy(x.view(np.int64).copy()).to(cuda_device), batch)
    assert (got.cpu().numpy().view(np.uint64) == _oracle_step(x, logn, q, p)).all()


def test_ipc_comm_single_rank(cuda_device):
    """mfhe_comm / mfhe_bconv_sharded_step with world = 1 (no peers to open): the C-ABI step equals the oracle."""
    import torch
    from matrix_fhe_gpu_b200 import sharded as S
    logn, l_in, l_out, bits, batch = 12, 6, 4, 55, 3
    N = 1 << logn
    primes = [int(v) for v in O.get_primes(N, bits, l_in + l_out)]
    q, p = primes[:l_in], primes[l_in:]
    x = O.synth_polys(13, batch, q, N)
    sh = S.LimbShardedBconv(logn, q, p, 0, 1)
    comm = S.IpcComm(0, 1, l_out, N, batch)
    for _ in range(3):                                   # both receive buffers, repeated epochs
        got = comm.step(sh, torch.from_numpy(x.view(np.int64).copy()).to(cuda_device), batch)
        assert (got.cpu().numpy().view(np.uint64) == _oracle_step(x, logn, q, p)).all()
    comm.barrier(); torch.cuda.synchronize()
    for groups in (2, 3, 16):                            # sliced pipeline (side stream, strided reduce); 16 > batch clamps to 3
        comm.set_groups(groups)
        for _ in range(3):
            got = comm.step(sh, torch.from_numpy(x.view(np.int64).copy()).to(cuda_device), batch)
            assert (got.cpu().numpy().view(np.uint64) == _oracle_step(x, logn, q, p)).all()
    with pytest.raises(RuntimeError):
        comm.set_groups(0)


@pytest.mark.parametrize("exchange", ["nccl", "pipelined", "p2p", "p2p-groups2", "ipc", "ipc-groups"])
def test_limb_sharded_step_nccl(exchange):
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(ROOT, "tools", "run_sharded.py"), "--logn", "13", "--lin", "6", "--lout", "4",
           "--batch", "3" if exchange in ("pipelined", "p2p-groups2", "ipc-groups") else "2", "--check", "--steps", "3",
           "--exchange", exchange.split("-")[0]]
    if exchange == "p2p-groups2":        # the two-stream form behind the config-5 number; a batch that does not divide evenly
        cmd += ["--p2p-groups", "2"]
    if exchange == "ipc-groups":         # mfhe_comm_set_groups: sliced software pipeline of the C-ABI step, uneven slices
        cmd += ["--ipc-groups", "2,3"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "parity ok" in r.stdout
