#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path: RNS negacyclic NTT+INTT throughput.

Workload (BASELINE.json configs[1]): N = 2^16, L = 32 sixty-bit primes of Phantom's
CoeffModulus::Create, batch 64 polys = 2048 limb-polys = 1 GiB of u64, synthetic
splitmix64 coefficients.  One step = forward NTT + inverse NTT of the whole batch, in place.
metric = limb-polys/s through NTT+INTT (a limb-poly counts once per NTT+INTT pair).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Prints ONE JSON line (rank 0).  `value` is device-resident throughput (CUDA events on the
launch stream, max over ranks); `e2e` is the same metric through the C-ABI host-buffer entry
points (pinned host memory, H2D and D2H inside the timed region).  `--impl reference` times
the CPU restatement of the reference algorithm (oracle/, OpenMP over all host threads) on a
bounded sample of the same workload: the reference has no CPU implementation of its own.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOGN, LIMBS, BATCH, BITS = 16, 32, 64, 60
N = 1 << LOGN
LIMB_POLYS = LIMBS * BATCH
ALGO_BYTES_PER_TRANSFORM = 16 * N                # 8 B read + 8 B write per coefficient
WORKLOAD = f"N=2^{LOGN} L={LIMBS} batch={BATCH} negacyclic NTT+INTT, {BITS}-bit primes (BASELINE configs[1])"
CPU_SAMPLE_POLYS = 2                             # both CPU numbers (cpu_baseline and --impl reference) use this sample


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per transform (the two launches of a forward / inverse transform of
    this workload) from the newest committed `ncu --set full` capture: profiles/ncu_traffic.json names the capture file
    and the commit it was taken at.  Not live -- DRAM counters need ncu, and a number printed under ncu is never a
    bench value -- so it is labelled with its source."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        return {"forward_bytes": None, "inverse_bytes": None, "source": "no capture committed"}
# register-only butterfly rate of the same arithmetic on this GPU (profiles/r1_04_intpipe.log, "CT bfly approx-quotient"):
# the integer (fmaheavy) pipe, not HBM, binds these kernels -- reported next to the HBM fraction
INT_PIPE_BFLY_PER_S_PER_SM = 7.02e9


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


L24_EXTRA = [17194410241, 17194804993, 17195989249, 17196581377, 17196976129, 17197963009, 17199936769, 17201318401,
             17201515777, 17203884289, 17206844929, 17207437057, 17209608193]


def secondary_metrics(dev, world, dist, pairs_per_rank=2, with_cpu=False):
    """BASELINE configs 3 and 4 on the matrix pipeline (native ring n=64, p=771): device-timed, reported
    next to the headline.  Every rank runs its own (re, im) pairs: batch-sharded, no collective."""
    import torch
    import matrix_fhe_gpu_b200.scheme as S
    out = {}

    def timed(fn, reps):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        fn(); torch.cuda.synchronize()
        ev[0].record()
        for _ in range(reps):
            fn()
        ev[1].record(); torch.cuda.synchronize()
        t = torch.tensor([ev[0].elapsed_time(ev[1]) / reps], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for L, tag in ((11, "config3_encode_decode_L11"), (24, "config4_pipeline_L24")):
        mods = (S.REF_MODULI + L24_EXTRA)[:L]
        torch.cuda.synchronize(); t_ctx = time.perf_counter()
        g = S.Context(moduli=mods)
        torch.cuda.synchronize(); t_ctx = (time.perf_counter() - t_ctx) * 1e3
        ell, i = np.meshgrid(np.arange(g.phi, dtype=np.float64), np.arange(g.n2, dtype=np.float64), indexing="ij")
        m = torch.from_numpy((ell + 0.001 * i) * (1 - 1j)).reshape(-1).to(dev)
        sk = g.generate_secret_key()
        re, im = g.encode_to_wntt_eval(m)
        cr, ci = g.encrypt_pair(re, im, sk)
        dec = g.decrypt_and_decode(cr, ci, sk)
        err = float((dec - m).abs().max())
        assert err < 1e-3, err
        if L == 11:
            # the W-CRT alone: a dense modular GEMM on the tcgen05 tensor cores (int8 digit planes, TMEM accumulators)
            a_u = g.sample_uniform()
            ms_w = timed(lambda: g.wntt_forward_matrix(a_u), 10)
            nd = [(int(q).bit_length() + 7) // 8 for q in mods]
            macs8 = sum(d * d for d in nd) * g.phi * g.phi * g.n2
            try:
                bf16 = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
                peak8, src = 2.0 * bf16 * 1e12 / 2.0, "2 x measured bf16 MAC rate (MEASURED_PEAKS.json; int8 is nominally twice bf16)"
            except Exception:
                peak8, src = 2.25e15, "nominal 4.5 Pop/s int8 dense"
            out["wcrt_forward_L11"] = {"ms": ms_w, "modular_macs_per_s": g.phi * g.phi * L * g.n2 / (ms_w * 1e-3),
                                       "int8_macs_per_s": macs8 / (ms_w * 1e-3),
                                       "roofline": {"bound": "tensor", "achieved": 2e-12 * macs8 / (ms_w * 1e-3),
                                                    "peak": 2e-12 * peak8, "unit": "Top/s (int8)",
                                                    "frac": macs8 / (ms_w * 1e-3) / peak8, "peak_source": src},
                                       "kernel": "wcrt_umma_kernel (tcgen05.mma kind::i8), profiles/r1_12_ncu_umma_summary.txt"}
            del a_u
            # encrypted matrix product: four Gaussian-integer GEMMs on the ciphertext components
            # (one tcgen05 kind::i8 kernel, U components stacked to M = 128, K = 2n) + the 4-term decryption
            prod = g.ct_matmul(cr, ci, cr, ci)
            ms_mm = timed(lambda: g.ct_matmul(cr, ci, cr, ci), 3)
            ms_md = timed(lambda: g.ct_matmul_decrypt(prod, sk), 3)
            macs_mm = 4 * 4 * sum(d * d for d in nd) * g.phi * g.n ** 3      # 4 GEMMs x 4 real products x digit pairs
            out["encrypted_matmul_L11"] = {
                "ct_matmul_ms": ms_mm, "ct_matmul_decrypt_ms": ms_md,
                "lane_products_per_s": g.phi / ((ms_mm + ms_md) * 1e-3),
                "what": "512 lanes of 64x64 complex matrices: R_l = V_l^H U_l from two (re, im) ciphertext pairs",
                "roofline": {"bound": "hbm", "algorithmic_bytes": int(16 * g.T * 8),
                             "achieved": 16 * g.T * 8 / (ms_mm * 1e-3) / 1e9, "peak": peaks()[0], "unit": "GB/s",
                             "frac": 16 * g.T * 8 / (ms_mm * 1e-3) / 1e9 / peaks()[0],
                             "int8_macs_per_s": macs_mm / (ms_mm * 1e-3),
                             "tensor_frac": macs_mm / (ms_mm * 1e-3) / 2.25e15,
                             "note": "ct_matmul_umma_kernel (tcgen05.mma kind::i8, TMEM accumulators): reads the 8 operand "
                                     "poly sets once, writes the 8 product poly sets; tensor_frac against the nominal 2.25 P "
                                     "int8 MAC/s the MMA stream itself reaches (profiles/r2_11_wcrt_limiters.txt); "
                                     "MFHE_MATMUL_PATH=m selects the four mma.sync GEMMs (2.5 ms)"}}
            del prod
            zero = torch.zeros_like(sk)
            ct_r, ct_i = g.allocate_ciphertext(), g.allocate_ciphertext()

            def loop():                          # test_encode_decode_wcrt.cu: encode -> (b = m, a = 0, s = 0) -> decode
                r, q = g.encode_to_wntt_eval(m)
                ct_r[: g.T] = r; ct_i[: g.T] = q
                g.decrypt_and_decode(ct_r, ct_i, zero, dec)
            out[tag] = {"ms_per_batch": timed(loop, 3), "encode_ms": timed(lambda: g.encode_to_wntt_eval(m), 3),
                        "decrypt_decode_ms": timed(lambda: g.decrypt_and_decode(ct_r, ct_i, zero, dec), 3),
                        "batch": "512 lanes x 64x64 complex fp64 (2^21 values)", "max_err": err,
                        "pipeline_pair_ms": timed(lambda: (g.encode_encrypt_pair(m, sk, cr, ci), g.decrypt_and_decode(cr, ci, sk, dec)), 3),
                        "pipeline_pair_three_calls_ms": timed(lambda: (g.encrypt_pair(*g.encode_to_wntt_eval(m), sk, cr, ci),
                                                                       g.decrypt_and_decode(cr, ci, sk, dec)), 3),
                        "context_create_ms": t_ctx,
                        "context_create_note": "mfhe_ctx_create: W-CRT tables built on the device (wtables_gpu.cu); the "
                                               "reference's init_he_backend is baselines.reference_cuda.init_he_backend_ms"}
            S_bytes = g.T * 8
            fl = pipeline_floors(g, mods, 4, 6 * S_bytes + 2 * g.phi * g.n2 * 16)     # encode: 32 MiB + 2S, decode: 4S + 32 MiB
            msb = out[tag]["ms_per_batch"]
            out[tag]["roofline"] = dict(fl, bound="tensor" if fl["tensor_floor_ms"] > fl["hbm_floor_ms"] else "hbm",
                                        frac_of_hbm_floor=fl["hbm_floor_ms"] / msb, frac_of_tensor_floor=fl["tensor_floor_ms"] / msb,
                                        frac=max(fl["hbm_floor_ms"], fl["tensor_floor_ms"]) / msb,
                                        note="2 forward + 2 inverse W-CRTs per batch dominate: the loop is tensor-bound, not HBM-bound")
            if with_cpu:
                cb = cpu_pipeline_sample(L, mods)
                out[tag]["cpu_baseline"] = {"value": cb["encode_decode_loop_ms"], "unit": "ms/batch", "cores": cb["cores"],
                                            "kind": cb["kind"], "sample": cb["sample"]}
        else:
            def pipe():
                for _ in range(pairs_per_rank):
                    r, q = g.encode_to_wntt_eval(m)
                    g.encrypt_pair(r, q, sk, cr, ci)
                    g.decrypt_and_decode(cr, ci, sk, dec)

            def pipe_fused():                    # same ciphertext words, message and error share a W-CRT
                for _ in range(pairs_per_rank):
                    g.encode_encrypt_pair(m, sk, cr, ci)
                    g.decrypt_and_decode(cr, ci, sk, dec)
            ms3 = timed(pipe, 2)
            ms = timed(pipe_fused, 2)
            errf = float((dec - m).abs().max())
            assert errf < 1e-3, errf
            out[tag] = {"pairs_per_s": world * pairs_per_rank / (ms * 1e-3), "ciphertexts_per_s": 2 * world * pairs_per_rank / (ms * 1e-3),
                        "ms_per_pair": ms / pairs_per_rank, "ms_per_pair_three_calls": ms3 / pairs_per_rank,
                        "pairs_per_rank_per_step": pairs_per_rank, "max_err": max(err, errf), "context_create_ms": t_ctx,
                        "note": "mfhe_encode_encrypt_pair -> decrypt_and_decode (three_calls: encode -> encrypt_pair -> "
                                "decrypt_and_decode, identical ciphertexts), one (re, im) pair = 2 ciphertexts of 768 MiB"}
            S_bytes = g.T * 8
            fl = pipeline_floors(g, mods, 5, 12 * S_bytes + 4 * g.phi * g.n2 * 16)     # 12 S + 64 MiB, API boundaries materialised
            msp = ms / pairs_per_rank
            out[tag]["roofline"] = dict(fl, bound="tensor" if fl["tensor_floor_ms"] > fl["hbm_floor_ms"] else "hbm",
                                        frac_of_hbm_floor=fl["hbm_floor_ms"] / msp, frac_of_tensor_floor=fl["tensor_floor_ms"] / msp,
                                        frac=max(fl["hbm_floor_ms"], fl["tensor_floor_ms"]) / msp,
                                        note="fused form: 3 forward (re, im, a) + 2 inverse W-CRTs per pair")
            if with_cpu:
                cb = cpu_pipeline_sample(L, mods)
                out[tag]["cpu_baseline"] = {"value": cb["pipeline_pair_ms"], "unit": "ms/pair", "cores": cb["cores"],
                                            "kind": cb["kind"], "sample": cb["sample"],
                                            "stages_ms": {k: cb[k] for k in ("encode_ms", "encrypt_pair_ms", "decrypt_and_decode_ms")}}
        del g
        torch.cuda.empty_cache()
    return out


def config2_50bit_block(dev, steps=20):
    """SURVEY config 2's secondary run: the headline shape on 50-bit primes.  Moduli below 2^50 let the butterflies run on
    the FP64 pipe (csrc/modmath.cuh); the integer kernels are timed on the same plan beside it."""
    import torch
    import matrix_fhe_gpu_b200 as m
    import oracle as O
    mods = O.coeff_modulus_create(N, 50, LIMBS)
    plan = m.NttPlan(LOGN, mods)
    x = O.synth_polys(43, BATCH, mods, N)
    d = torch.from_numpy(x.view(np.int64)).to(dev)
    stream = torch.cuda.current_stream()
    peak, _ = peaks()
    out = {"workload": WORKLOAD.replace("60-bit", "50-bit"), "default_arith": "fp64" if plan.arith == 1 else "int"}

    def timed(chunk):
        plan.set_chunk(chunk)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        tf, ti = [], []
        for i in range(steps + 3):
            ev[0].record(stream); plan.forward_(d, stream)
            ev[1].record(stream); plan.inverse_(d, stream)
            ev[2].record(stream); torch.cuda.synchronize()
            if i >= 3:
                tf.append(ev[0].elapsed_time(ev[1])); ti.append(ev[1].elapsed_time(ev[2]))
        return statistics.median(tf), statistics.median(ti)

    for name, arith in (("fp64", 1), ("int", 0)):
        if arith == 1 and plan.arith != 1:
            try:
                plan.set_arith(1)
            except Exception as e:                       # noqa: BLE001
                out[name] = {"error": repr(e)[:200]}
                continue
        plan.set_arith(arith)
        tf, ti = timed(0)
        blk = {"ms_forward": tf, "ms_inverse": ti, "limb_polys_per_s": LIMB_POLYS / ((tf + ti) * 1e-3),
               "hbm_frac": ALGO_BYTES_PER_TRANSFORM * LIMB_POLYS / (max(tf, ti) * 1e-3) / 1e9 / peak}
        if arith == 1:                                   # with a cheaper butterfly the L2-resident second pass can pay
            chunks = {}
            for c in (128, 256, 512):
                cf, ci = timed(c)
                chunks[str(c)] = LIMB_POLYS / ((cf + ci) * 1e-3)
            blk["chunk_limb_polys_ab"] = chunks
        out[name] = blk
    plan.set_chunk(0)
    assert (d[0].cpu().numpy().view(np.uint64) == x[0]).all(), "50-bit round trip failed"
    out["note"] = ("same results word for word (tests/test_ntt_fp64_gpu.py); Phantom on the same primes: "
                   "baselines.phantom.cfg2_50bit_limb_poly_pairs_per_s")
    return out


def link_ceiling(hbuf, dev):
    """Host <-> device copy rates of this box with the e2e buffers themselves (pinned, 1 GiB): one direction at a time
    and both at once -- the ceiling the e2e number is read against."""
    import torch
    dbuf = torch.empty_like(hbuf, device=dev)
    dbuf2 = torch.empty_like(hbuf, device=dev)
    hout = torch.empty_like(hbuf).pin_memory()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    nbytes = hbuf.numel() * hbuf.element_size()

    def timed(fn, reps=3):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps

    def h2d():
        with torch.cuda.stream(s1):
            dbuf.copy_(hbuf, non_blocking=True)

    def d2h():
        with torch.cuda.stream(s2):
            hout.copy_(dbuf2, non_blocking=True)

    def both():
        h2d(); d2h()

    return {"h2d_gbs": nbytes / timed(h2d) / 1e9, "d2h_gbs": nbytes / timed(d2h) / 1e9,
            "both_directions_each_gbs": nbytes / timed(both) / 1e9,
            "note": "pinned 1 GiB copies, cudaMemcpyAsync on separate streams; e2e moves 2 GiB each way per step"}


def gpu_baselines():
    """The two GPU baselines BASELINE.json's north_star names, timed on THIS box in THIS run: the reference's own CUDA
    kernels and Phantom's, i.e. the UNMODIFIED sources of /root/reference compiled for sm_100a by baseline/build_ref.py
    (test infrastructure; the stock flavour = the reference's own CMake flags) and driven by baseline/ref_time.cu.
    Runs after our own timed region, on rank 0 of a single-GPU run only."""
    exe = os.path.join(ROOT, "baseline", "_ref", "ref_time_stock")
    if not os.path.exists(exe):
        return {"unavailable": "baseline/_ref/ref_time_stock not built (baseline/build_ref.py needs /root/reference)"}
    out = os.path.join(ROOT, "gpurun_out", "ref_time_bench.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    try:
        r = subprocess.run([exe, out, "2"], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, timeout=420)
        if r.returncode != 0:
            return {"unavailable": f"ref_time_stock exited {r.returncode}: {r.stderr[-200:]}"}
        b = json.load(open(out))
    except Exception as e:                                  # noqa: BLE001 -- a baseline must never take the bench down
        return {"unavailable": f"ref_time_stock failed: {e!r}"[:300]}
    ph, rc = b.get("phantom", {}), b.get("reference_cuda", {})
    b["how"] = ("unmodified /root/reference sources (src/core/*.cu + vendored Phantom), nvcc sm_100a, -O3 --use_fast_math "
                "-DPHANTOM_USE_CUDA_PTX (CMakeLists.txt:11,14,95-99 with the arch changed), driven by baseline/ref_time.cu; " + b.pop("timing", ""))
    b["headline_same_workload"] = {
        "phantom_limb_polys_per_s": ph.get("cfg2_60bit_limb_poly_pairs_per_s"),
        "what": "nwt_2d_radix8_forward_inplace + backward_inplace, 64 calls per direction on [64][32][65536], 60-bit primes"}
    b["notes"] = ["Phantom's CKKS encoder returns garbage on sm_100a in this build (ckks_*_max_err): its times are of the "
                  "code as it runs, not of a verified result",
                  "reference_cuda.pipeline_max_err is the reference's own decode error on its e2e driver message"]
    return b


def _tensor_peak():
    try:
        bf16 = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
        return bf16 * 1e12, "2 x measured bf16 MAC rate (MEASURED_PEAKS.json; int8 is nominally twice bf16)"
    except Exception:
        return 2.25e15, "nominal 4.5 Pop/s int8 dense"


def pipeline_floors(g, mods, n_wcrt, algo_bytes):
    """HBM floor (SURVEY 8(d) algorithmic bytes / measured HBM peak) and tensor floor (#W-CRT x int8 digit MACs / int8
    peak) of a matrix-pipeline call sequence, in ms."""
    hbm, _ = peaks()
    nd = [(int(q).bit_length() + 7) // 8 for q in mods]
    macs8 = sum(d * d for d in nd) * g.phi * g.phi * g.n2
    peak8, src = _tensor_peak()
    return {"hbm_floor_ms": algo_bytes / (hbm * 1e9) * 1e3, "tensor_floor_ms": n_wcrt * macs8 / peak8 * 1e3,
            "algorithmic_bytes": int(algo_bytes), "wcrt_calls": n_wcrt, "int8_macs_per_wcrt": int(macs8), "tensor_peak_source": src}


def cpu_pipeline_sample(L, mods):
    """CPU baseline of the matrix pipeline: the oracle's encode -> encrypt_pair -> decrypt_and_decode at n = 16, i.e. on
    1/16 of the n^2 = 4096 independent (y, x) columns of the native ring (same phi = 512 lanes and limbs; the dense W-CRT,
    ~95 % of the CPU time, is linear in the column count), scaled by 16 and labelled as such."""
    import oracle as O
    cores = O.use_all_cores()
    o = O.SchemeCtx(16, 257, mods, 2.0 ** 35)
    ell, i = np.meshgrid(np.arange(o.phi, dtype=np.float64), np.arange(o.n2, dtype=np.float64), indexing="ij")
    msg = ((ell + 0.001 * i) * (1 - 1j)).reshape(-1)
    sk = o.keygen()
    t0 = time.perf_counter(); re, im = o.encode(msg); t1 = time.perf_counter()
    cr, ci = o.encrypt_pair(re, im, sk); t2 = time.perf_counter()
    out = o.decrypt_and_decode(cr, ci, sk); t3 = time.perf_counter()
    assert np.abs(out - msg).max() < 1e-3
    zk = np.zeros_like(sk); zr = np.zeros_like(cr); zi = np.zeros_like(ci); zr[: o.T] = re; zi[: o.T] = im
    t4 = time.perf_counter(); o.decrypt_and_decode(zr, zi, zk); t5 = time.perf_counter()
    scale = (64 * 64) / o.n2
    return {"encode_ms": (t1 - t0) * 1e3 * scale, "encrypt_pair_ms": (t2 - t1) * 1e3 * scale,
            "decrypt_and_decode_ms": (t3 - t2) * 1e3 * scale, "pipeline_pair_ms": (t3 - t0) * 1e3 * scale,
            "encode_decode_loop_ms": ((t1 - t0) + (t5 - t4)) * 1e3 * scale,
            "cores": cores, "kind": "port",
            "sample": f"oracle pipeline at n=16 (256 of the 4096 (y,x) columns, phi=512, L={L}), one run, times scaled x{scale:.0f}"}


def config5_block(dev, world, rank, dist, steps=10):
    """BASELINE config 5: N = 2^17, 48 -> 16 limbs of 55-bit primes, batch 16: INTT over the input base -> base conversion ->
    NTT over the output base.  1 GPU: unsharded.  N >= 2: limbs sharded over the ranks, the partial sums leave the
    base-conversion kernel as peer stores over NVLink (symmetric memory), one device barrier, then reduce + NTT."""
    import torch
    import oracle as O
    from matrix_fhe_gpu_b200 import sharded as SH
    logn, lin, lout, bits, batch = 17, 48, 16, 55, 16
    n = 1 << logn
    primes = [int(v) for v in O.get_primes(n, bits, lin + lout)]
    q, p = primes[:lin], primes[lin:]
    sh = SH.LimbShardedBconv(logn, q, p, rank, world)
    qs, qc = sh.q_windows[rank]; ps, pc = sh.p_windows[rank]
    idx = (np.arange(batch, dtype=np.uint64)[:, None, None] * np.uint64(lin * n)
           + np.arange(qs, qs + qc, dtype=np.uint64)[None, :, None] * np.uint64(n)
           + np.arange(n, dtype=np.uint64)[None, None, :] + np.uint64(42))
    x_loc = O.splitmix64(idx) % np.asarray(q[qs:qs + qc], dtype=np.uint64).reshape(1, -1, 1)
    x0 = torch.from_numpy(x_loc.view(np.int64)).to(dev)
    work = torch.empty_like(x0)
    kind, groups, ipc, symm_ok = "none (one GPU, unsharded)", 1, None, False
    if world > 1:
        try:
            ipc = SH.IpcComm(rank, world, lout, n, batch)
            kind = ("mfhe_bconv_sharded_step (C ABI): bconv kernel stores partial sums into the owners' CUDA-IPC receive "
                    "buffers over NVLink, flag barrier (release/acquire.sys), no NCCL")
        except Exception as e:                              # noqa: BLE001
            kind = f"NCCL all_to_all of partial sums (CUDA IPC unavailable: {e!r})"[:200]
        try:
            sh.enable_p2p(batch)
            symm_ok = True
        except Exception:                                   # noqa: BLE001
            symm_ok = False

    def step():
        work.copy_(x0)
        if world == 1:
            return sh.step_unsharded(work, batch)
        if ipc is not None:
            return ipc.step(sh, work, batch)
        return sh.step(work, batch)

    out = None
    for _ in range(3):
        out = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(steps):
        out = step()
    ev[1].record(); torch.cuda.synchronize()
    t = torch.tensor([ev[0].elapsed_time(ev[1]) / steps], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    # parity: poly 0 of this rank's output window against the CPU oracle (INTT + bConv_BEHZ + NTT restated)
    x_full0 = O.synth_polys(42, 1, q, n)
    coeff = O.PhantomTables(logn, q).inverse(x_full0)
    want = O.PhantomTables(logn, p).forward(O.bconv(coeff[0], q, p)[None])[0, ps:ps + pc]
    ok = bool((out[0].cpu().numpy().view(np.uint64) == want).all())
    flag = torch.tensor([1 if ok else 0], device=dev)
    stages = None
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if symm_ok:
            acc = np.zeros(4)
            for _ in range(5):
                work.copy_(x0)
                dist.barrier()
                _, st = sh.step_p2p_stages(work, batch)
                acc += np.asarray(st)
            tt = torch.tensor(acc / 5, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            stages = [float(v) for v in tt.tolist()]
    sent = batch * (lout - pc) * n * 8 if world > 1 else 0
    blk = {"polys_per_s": batch / (ms * 1e-3), "ms_per_step": ms, "n_gpus": world,
           "config": f"N=2^{logn}, {lin} -> {lout} limbs of {bits}-bit primes, batch {batch} (BASELINE configs[4])",
           "exchange": kind, "stream_groups": groups, "bytes_sent_per_rank": sent,
           "parity": "ok" if bool(flag.item()) else "FAILED",
           "parity_check": "poly 0, every rank's output limbs, vs the CPU oracle (INTT + bConv_BEHZ + NTT restated)",
           "algorithmic_bytes_per_step": batch * n * 8 * (2 * lin + 2 * lout)}
    if world == 1:
        # the base conversion alone (B2: bConv_BEHZ, 48 -> 16 limbs) -- tensor-core kernel (bconv_umma.cuh) next to the CUDA-core one
        def conv_ms(path):
            if path:
                os.environ["MFHE_BCONV_PATH"] = path
            try:
                sh.bconv.convert(x0, n, batch); torch.cuda.synchronize()
                ev[0].record()
                for _ in range(steps):
                    sh.bconv.convert(x0, n, batch)
                ev[1].record(); torch.cuda.synchronize()
                return ev[0].elapsed_time(ev[1]) / steps
            finally:
                os.environ.pop("MFHE_BCONV_PATH", None)
        ms_t, ms_c = conv_ms(None), conv_ms("c")
        byts = batch * n * 8 * (lin + lout)
        blk["base_conversion_48_16"] = {
            "ms": ms_t, "ms_cuda_core_kernel": ms_c, "modular_macs_per_s": batch * n * lin * lout / (ms_t * 1e-3),
            "roofline": {"bound": "hbm", "algorithmic_bytes": byts, "achieved": byts / (ms_t * 1e-3) / 1e9, "peak": peaks()[0],
                         "unit": "GB/s", "frac": byts / (ms_t * 1e-3) / 1e9 / peaks()[0]},
            "kernel": "bconv_umma_kernel (tcgen05.mma kind::i8 digit GEMM, TMEM accumulators): 128 coefficients x 16 output limbs per "
                      "tile, K = input limbs; ms_cuda_core_kernel = bconv_kernel (128-bit lazy multiply-accumulate)"}
    if world > 1 and symm_ok:                               # the torch symmetric-memory form of the same exchange, two stream groups
        def step2():
            work.copy_(x0)
            return sh.step_p2p(work, batch, 2)
        for _ in range(3):
            step2()
        torch.cuda.synchronize(); dist.barrier()
        ev[0].record()
        for _ in range(steps):
            step2()
        ev[1].record(); torch.cuda.synchronize()
        t2 = torch.tensor([ev[0].elapsed_time(ev[1]) / steps], device=dev)
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        blk["alt_symmetric_memory_two_streams_ms_per_step"] = float(t2.item())
    if stages:
        blk["stage_ms"] = {"intt": stages[0], "scatter": stages[1], "barrier": stages[2], "reduce_ntt": stages[3],
                           "note": "one stream group, CUDA events between the stages, max over ranks"}
        blk["nvlink_gbs_per_rank"] = sent / (stages[1] * 1e-3) / 1e9 if stages[1] > 0 else None
        blk["nvlink_note"] = "bytes this rank stores into peers / the scatter kernel's own duration"
    return blk


class ClockSampler:
    """Samples SM clock / throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def _nvml_loop(self):
        import pynvml as nv
        bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        while not self.halt.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                mx = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
                rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.nv_rows.append((float(sm), float(mx), [k for k, b in bits.items() if rs & b]))
            except Exception:
                pass
            self.halt.wait(0.005)

    def start(self):
        # NVML in-process (5 ms period: short runs, e.g. the multi-GPU scaling points, still get samples);
        # nvidia-smi -lms 50 as the fallback
        self.nv_rows, self.halt = [], threading.Event()
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].strip().isdigit() else self.index
            self.h = nv.nvmlDeviceGetHandleByIndex(idx)
            nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
            self.t = threading.Thread(target=self._nvml_loop, daemon=True)
            self.t.start()
            self.proc = "nvml"
            return
        except Exception:
            pass
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        if self.proc == "nvml":
            self.halt.set(); self.t.join(timeout=2)
            sm = [r[0] for r in self.nv_rows]; mx = [r[1] for r in self.nv_rows]
            reasons = sorted({x for r in self.nv_rows for x in r[2]})
            return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                    "reasons": reasons, "samples": len(sm), "source": "nvml, 5 ms period"}
        time.sleep(0.06)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nme, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_sample(sample_limb_polys: int, reps: int):
    """Oracle (CPU restatement of the reference algorithm) on a bounded sample, all host threads."""
    import oracle as O
    O.use_all_cores()                            # torch.distributed.run exports OMP_NUM_THREADS=1: set and state it
    mods = O.coeff_modulus_create(N, BITS, LIMBS)
    t = O.PhantomTables(LOGN, mods)
    polys = max(1, sample_limb_polys // LIMBS)
    x = O.synth_polys(42, polys, mods, N)
    lib = O.lib()
    flat = x.reshape(-1)
    args_f = (flat, polys, LIMBS, LOGN, t.moduli, t.tw.reshape(-1), t.tws.reshape(-1))
    args_i = (flat, polys, LIMBS, LOGN, t.moduli, t.itw.reshape(-1), t.itws.reshape(-1), t.n_inv, t.n_inv_shoup)
    lib.mfo_negacyclic_forward_batch(*args_f); lib.mfo_negacyclic_inverse_batch(*args_i)     # warm
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        lib.mfo_negacyclic_forward_batch(*args_f)
        lib.mfo_negacyclic_inverse_batch(*args_i)
        times.append(time.perf_counter() - t0)
    assert (flat.reshape(x.shape) == O.synth_polys(42, polys, mods, N)).all()
    return polys * LIMBS, times, lib.mfo_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_lp, times, cores = cpu_sample(CPU_SAMPLE_POLYS * LIMBS, max(1, args.steps))
    ms = 1e3 * min(times)
    value = n_lp / min(times)
    sample = (f"{n_lp} limb-polys ({n_lp // LIMBS} of {BATCH} polys x {LIMBS} limbs) per step, NTT+INTT, OpenMP over "
              f"{cores} threads, best of {len(times)} steps (same rule as the native arm's cpu_baseline)")
    line = {
        "impl": "reference", "metric": "ntt_intt_limb_polys_per_s", "value": value, "unit": "limb-polys/s",
        "n_gpus": args.gpus, "steps": len(times), "warmup": 1, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "note": "CPU restatement of the reference algorithm (oracle/); the reference ships no CPU path: its own "
                           "CUDA kernels and Phantom's are timed by the native arm as `baselines`"},
        "cpu_baseline": {"value": value, "unit": "limb-polys/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "limb-polys/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--chunk", type=int, default=int(os.environ.get("MFHE_CHUNK", "-1")),
                    help="limb-polys per kernel pair (L2-resident chunking); -1 = library default")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the matrix-pipeline secondary metrics")
    ap.add_argument("--no-baselines", action="store_true", help="skip timing the reference's and Phantom's CUDA kernels")
    ap.add_argument("--no-config5", action="store_true", help="skip the limb-sharded config-5 block")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import matrix_fhe_gpu_b200 as m
    import oracle as O          # inputs + cpu_baseline + one-poly self-check only

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    mods = O.coeff_modulus_create(N, BITS, LIMBS)
    plan = m.NttPlan(LOGN, mods)
    if args.chunk >= 0:
        plan.set_chunk(args.chunk)
    # synthetic input, each rank its own batch (weak scaling over the batch axis, no collective)
    x = O.synth_polys(42 + rank, BATCH, mods, N)
    d = torch.from_numpy(x.view(np.int64)).to(dev)
    stream = torch.cuda.current_stream()

    def step():
        plan.forward_(d, stream)
        plan.inverse_(d, stream)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    # correctness guard inside the bench: after W round trips poly 0 is still the input
    assert (d[0].cpu().numpy().view(np.uint64) == x[0]).all(), "round trip failed"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    fwd_ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    barrier()
    if rank == 0:
        sampler.start()
    ev[0].record(stream)
    for _ in range(args.steps):
        step()
    ev[1].record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_total = ev[0].elapsed_time(ev[1])
    # per-transform durations (forward = column+row kernels, inverse = row+column kernels)
    tf, ti = [], []
    for _ in range(10):
        fwd_ev[0].record(stream); plan.forward_(d, stream)
        fwd_ev[1].record(stream); plan.inverse_(d, stream)
        fwd_ev[2].record(stream); torch.cuda.synchronize()
        tf.append(fwd_ev[0].elapsed_time(fwd_ev[1])); ti.append(fwd_ev[1].elapsed_time(fwd_ev[2]))
    t = torch.tensor([ms_total], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_step = ms_total / args.steps
    value = world * LIMB_POLYS / (ms_step * 1e-3)

    # ---- end to end through the C ABI with host buffers (pinned), copies inside the timing
    hbuf = torch.from_numpy(x.view(np.int64).copy()).pin_memory()
    hnp = hbuf.numpy().view(np.uint64)
    plan.forward_host_(hnp); plan.inverse_host_(hnp)        # warm (allocates staging)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        plan.forward_host_(hnp)
        plan.inverse_host_(hnp)
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / args.e2e_steps
    te = torch.tensor([e2e_s], device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())
    assert (hnp.reshape(x.shape)[0] == x[0]).all()
    nbytes = x.nbytes
    try:
        barrier()
        link = link_ceiling(hbuf, dev)
        if world > 1:                                       # all ranks copy at once, like the e2e step: slowest rank
            tl = torch.tensor([link["h2d_gbs"], link["d2h_gbs"], link["both_directions_each_gbs"]], device=dev)
            dist.all_reduce(tl, op=dist.ReduceOp.MIN)
            link.update(h2d_gbs=float(tl[0]), d2h_gbs=float(tl[1]), both_directions_each_gbs=float(tl[2]),
                        ranks="minimum over ranks, all ranks copying at the same time")
    except Exception as e:                                  # noqa: BLE001
        link = {"error": repr(e)[:200]}

    secondary = None
    if not args.no_secondary:
        del d
        torch.cuda.empty_cache()
        secondary = secondary_metrics(dev, world, dist if world > 1 else None, with_cpu=(rank == 0 and world == 1 and not args.no_cpu))
        if rank == 0:
            try:
                secondary["config2_50bit_primes"] = config2_50bit_block(dev)
            except Exception as e:                          # noqa: BLE001
                secondary["config2_50bit_primes"] = {"error": repr(e)[:400]}
        if not args.no_config5:
            try:
                secondary["config5_limb_sharded"] = config5_block(dev, world, rank, dist if world > 1 else None)
            except Exception as e:                          # noqa: BLE001 -- report, never lose the headline line
                secondary["config5_limb_sharded"] = {"error": repr(e)[:400]}
    if rank == 0:
        peak, peak_src = peaks()
        t_fwd = statistics.median(tf) * 1e-3
        t_inv = statistics.median(ti) * 1e-3
        algo = ALGO_BYTES_PER_TRANSFORM * LIMB_POLYS
        worst = max(t_fwd, t_inv)
        achieved = algo / worst / 1e9
        line = {
            "metric": "ntt_intt_limb_polys_per_s", "value": value, "unit": "limb-polys/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "bytes_resident": int(nbytes), "l2": "input (1 GiB) larger than L2, no flush needed",
                       "partition": "batch axis, one batch per GPU, no collective",
                       "chunk_limb_polys": args.chunk},
            "clocks": clocks,
            "e2e": {"value": world * LIMB_POLYS / e2e_s, "unit": "limb-polys/s",
                    "h2d_bytes_per_step": 2 * nbytes, "d2h_bytes_per_step": 2 * nbytes,
                    "ms_per_step": e2e_s * 1e3,
                    "path": "mfhe_ntt_forward_host + mfhe_ntt_inverse_host on pinned host memory",
                    "link_gbs_each_way": 2 * nbytes / e2e_s / 1e9, "link_ceiling": link},
            "gpu_launches": 4 * args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak,
                         "traffic": ncu_traffic()["inverse_bytes" if t_inv >= t_fwd else "forward_bytes"],
                         "traffic_note": "dram read+write bytes of the two launches of that transform; " + str(ncu_traffic().get("source")),
                         "algorithmic_bytes_per_launch_pair": algo,
                         "int_pipe": {"butterflies_per_s": LIMB_POLYS * (N // 2) * LOGN / worst,
                                      "peak_butterflies_per_s": INT_PIPE_BFLY_PER_S_PER_SM * 148,
                                      "frac": LIMB_POLYS * (N // 2) * LOGN / worst / (INT_PIPE_BFLY_PER_S_PER_SM * 148),
                                      "note": "binding roofline: fmaheavy pipe 74-80 % busy in ncu; peak = register-only "
                                              "butterfly loop measured on this B200 (profiles/r1_04_intpipe.log)"},
                         "peak_source": peak_src,
                         "kernel": "slower of forward (ntt_cols_kernel+ntt_rows_fwd_kernel) and inverse "
                                   "(ntt_rows_inv_kernel+ntt_cols_kernel) transform; algorithmic bytes = "
                                   "16*N per limb-poly per transform",
                         "ms_forward": t_fwd * 1e3, "ms_inverse": t_inv * 1e3,
                         "pair_frac": (2 * algo / (t_fwd + t_inv) / 1e9) / peak},
        }
        if secondary is not None:
            line["secondary"] = secondary
        if not args.no_cpu:
            n_lp, times, cores = cpu_sample(CPU_SAMPLE_POLYS * LIMBS, 3)
            line["cpu_baseline"] = {"value": n_lp / min(times), "unit": "limb-polys/s", "cores": cores,
                                    "kind": "port",
                                    "sample": f"{n_lp} limb-polys ({CPU_SAMPLE_POLYS} of {BATCH} polys x {LIMBS} limbs), NTT+INTT, "
                                              f"OpenMP over {cores} threads, best of 3"}
        if world == 1 and not args.no_baselines:
            line["baselines"] = gpu_baselines()
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
