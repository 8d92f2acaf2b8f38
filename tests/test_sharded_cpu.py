"""World-size-2 gloo tests of the multi-GPU host logic: limb windows, batch slices and the
all-to-all of base-conversion partial sums.  The partial sums themselves are formed here with Python
integers (the kernels need a GPU); the result is checked against the oracle's full base conversion."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle as O
from matrix_fhe_gpu_b200 import sharded as S


def test_limb_windows_cover_and_balance():
    assert S.limb_windows(48, 8) == [(6 * r, 6) for r in range(8)]
    assert S.limb_windows(16, 8) == [(2 * r, 2) for r in range(8)]
    for total in (1, 5, 7, 48):
        for world in (1, 2, 3, 8):
            w = S.limb_windows(total, world)
            assert sum(c for _, c in w) == total
            assert all(w[i][0] + w[i][1] == w[i + 1][0] for i in range(world - 1))
            assert max(c for _, c in w) - min(c for _, c in w) <= 1


def test_batch_slice_partitions():
    for batch in (1, 7, 128):
        for world in (1, 2, 4, 8):
            seen = sorted(i for r in range(world) for i in S.batch_slice(batch, r, world))
            assert seen == list(range(batch))


def _py_partial(x, q, p, window):
    """sum over input limbs in `window` of y_i * (Q/q_i mod p_j) mod p_j with Python integers; x [L_in][N]."""
    qs = [int(v) for v in q]; ps = [int(v) for v in p]
    Q = 1
    for v in qs:
        Q *= v
    N = x.shape[1]
    out = np.zeros((len(ps), N), dtype=np.uint64)
    s, c = window
    for j, pj in enumerate(ps):
        acc = [0] * N
        for i in range(s, s + c):
            qh = Q // qs[i]
            inv = pow(qh % qs[i], -1, qs[i])
            f = qh % pj
            for d in range(N):
                acc[d] = (acc[d] + (int(x[i, d]) * inv % qs[i]) * f) % pj
        out[j] = np.array(acc, dtype=np.uint64)
    return out


def _worker(rank, world, port, q, p, x, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        batch = x.shape[0]
        qwin = S.limb_windows(len(q), world)
        pwin = S.limb_windows(len(p), world)
        partial = np.stack([_py_partial(x[b], q, p, qwin[rank]) for b in range(batch)])
        t = torch.from_numpy(partial.view(np.int64))
        parts = S.exchange_partials(t, pwin, rank, world)
        assert tuple(parts.shape) == (world, batch, pwin[rank][1], x.shape[2])
        arr = parts.numpy().view(np.uint64)
        ps, pc = pwin[rank]
        out = np.zeros((batch, pc, x.shape[2]), dtype=np.uint64)
        for j in range(pc):
            pj = int(p[ps + j])
            acc = np.zeros((batch, x.shape[2]), dtype=object)
            for r in range(world):
                acc = acc + arr[r, :, j, :].astype(object)
            out[:, j, :] = (acc % pj).astype(np.uint64)
        ret[rank] = out
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("l_in,l_out", [(4, 3), (5, 2)])
def test_exchange_partials_gloo_world2(l_in, l_out):
    world, N, batch = 2, 32, 2
    primes = O.get_primes(2 * N, 40, l_in + l_out)
    q, p = primes[:l_in], primes[l_in:]
    x = O.synth_polys(5, batch, q, N)
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), q, p, x, ret), nprocs=world, join=True)
        pwin = S.limb_windows(l_out, world)
        for b in range(batch):
            want = O.bconv(x[b], q, p)
            for r in range(world):
                s, c = pwin[r]
                assert (ret[r][b] == want[s:s + c]).all()


def test_emulate_exchange_matches_pack():
    world, batch, l_out, N = 3, 2, 5, 8
    win = S.limb_windows(l_out, world)
    partials = [torch.arange(batch * l_out * N, dtype=torch.int64).view(batch, l_out, N) + 1000 * r for r in range(world)]
    for r in range(world):
        parts = S.emulate_exchange(partials, win, r)
        s, c = win[r]
        for src in range(world):
            assert torch.equal(parts[src], partials[src][:, s:s + c, :])
        send, splits = S.pack_for_exchange(partials[r], win)
        assert splits == [batch * cc * N for _, cc in win]
        off = sum(splits[:r])
        assert torch.equal(send[off:off + splits[r]].view(batch, c, N), partials[r][:, s:s + c, :])


@pytest.mark.parametrize("l_out,world,batch,n", [(16, 8, 3, 8), (5, 2, 2, 4), (7, 4, 1, 16)])
def test_scatter_table_addresses_receive_layout(l_out, world, batch, n):
    """scatter_table must address slot [src][b][j - start][coeff] of the owner's [world][batch][count][N] buffer:
    writing partials through the table into numpy stand-ins of the peer buffers gives exactly emulate_exchange."""
    import torch
    wins = S.limb_windows(l_out, world)
    cmax = max(c for _, c in wins)
    words = world * batch * cmax * n
    bases = [r * words * 8 + 4096 for r in range(world)]            # fake, disjoint "peer addresses"
    mem = {r: np.full(words, -1, dtype=np.int64) for r in range(world)}
    rng = np.random.default_rng(3)
    partials = [rng.integers(0, 1 << 60, size=(batch, l_out, n), dtype=np.int64) for _ in range(world)]
    for src in range(world):
        ptrs, strides = S.scatter_table(bases, wins, src, batch, n)
        assert len(ptrs) == l_out and len(strides) == l_out
        for j in range(l_out):
            owner = next(r for r, (s, c) in enumerate(wins) if s <= j < s + c)
            off = (ptrs[j] - bases[owner]) // 8
            assert (ptrs[j] - bases[owner]) % 8 == 0 and 0 <= off
            for b in range(batch):
                lo = off + b * strides[j]
                assert lo + n <= words
                mem[owner][lo:lo + n] = partials[src][b, j]
    for r, (s, c) in enumerate(wins):
        want = S.emulate_exchange([torch.from_numpy(p) for p in partials], wins, r).numpy()
        got = mem[r][: world * batch * c * n].reshape(world, batch, c, n)
        assert (got == want).all()


def _p2p_worker(rank, world, port, bufs, l_out, batch, n, steps, ret):
    """The step_p2p protocol with host stand-ins: `bufs[phase][r]` are shared-memory tensors playing the ranks'
    symmetric receive buffers, the stores go through scatter_table's addresses (base 0 = word offsets), dist.barrier()
    plays the device barrier.  There is NO barrier after the read, as on the device: only the double buffering keeps a
    fast rank's stores of step k+1 away from a slow rank's reads of step k."""
    import time
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        wins = S.limb_windows(l_out, world)
        mine = wins[rank][1]
        ok = True
        for step in range(steps):
            everyone = [torch.from_numpy(np.random.default_rng(1000 * step + r).integers(
                0, 1 << 60, size=(batch, l_out, n), dtype=np.int64)) for r in range(world)]
            partial = everyone[rank]
            want = S.emulate_exchange(everyone, wins, rank)                 # no collective here: nothing else syncs the ranks
            phase = step & 1
            ptrs, strides = S.scatter_table([0] * world, wins, rank, batch, n)
            for j in range(l_out):
                owner = next(r for r, (s, c) in enumerate(wins) if s <= j < s + c)
                for b in range(batch):
                    off = ptrs[j] // 8 + b * strides[j]
                    bufs[phase][owner][off:off + n] = partial[b, j]
            dist.barrier()
            if rank == 0:
                time.sleep(0.05)                                            # rank 0 reads late, rank 1 runs ahead
            got = bufs[phase][rank][: world * batch * mine * n].view(world, batch, mine, n).clone()
            ok = ok and bool(torch.equal(got, want))
        ret[rank] = ok
    finally:
        dist.destroy_process_group()


def test_p2p_protocol_world2_shared_memory():
    world, l_out, batch, n, steps = 2, 5, 2, 16, 6
    wins = S.limb_windows(l_out, world)
    cmax = max(c for _, c in wins)
    bufs = [[torch.full((world * batch * cmax * n,), -1, dtype=torch.int64).share_memory_() for _ in range(world)]
            for _ in range(2)]
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_p2p_worker, args=(world, _free_port(), bufs, l_out, batch, n, steps, ret), nprocs=world, join=True)
        assert all(ret[r] for r in range(world))
