"""Device timings of the remaining SURVEY 8(a) rows at the native shape (n = 64, phi = 512 lanes, L = 11): trace step
(A18), ciphertext ops (A17), layout passes (A13), samplers / keygen (A14), n-point X transforms (A3/A4), vector and
centred W-CRT (A7/A8), single-lane encoder (A10), big CRT compose (A11).  CUDA events, median of 10.
Prints name, ms, algorithmic bytes, GB/s and the fraction of the measured HBM peak.   python tools/time_ops.py"""
import json
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import matrix_fhe_gpu_b200 as m
import matrix_fhe_gpu_b200.scheme as S

try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6447.5
g = S.Context()
dev = torch.device("cuda:0")
T, L, n, n2, phi = g.T, g.L, g.n, g.n2, g.phi
mods = torch.tensor([int(q) for q in g.copy_device_moduli()], dtype=torch.int64, device=dev)


def rand_set():
    x = torch.randint(0, 2 ** 34, (phi, L, n2), dtype=torch.int64, device=dev)
    return (x % mods.view(1, L, 1)).reshape(-1).contiguous()


def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for _ in range(reps):
        ev[0].record(); fn(); ev[1].record(); torch.cuda.synchronize()
        ts.append(ev[0].elapsed_time(ev[1]))
    return statistics.median(ts)


A = [rand_set() for _ in range(4)]
sk = g.generate_secret_key()
ct1 = torch.cat([A[0], A[1]]); ct2 = torch.cat([A[2], A[3]])
S8 = T * 8
bpr, bpi = g.map_B_to_Bprime(A[2], A[3], phi)
cr, ci = g.trace_gemm(A[0], A[1], bpr, bpi, phi)
msg = torch.randn(phi * n2, dtype=torch.complex128, device=dev)
lane = msg[:n2].contiguous()
vec = (torch.randint(0, 2 ** 34, (phi, L, n), dtype=torch.int64, device=dev) % mods.view(1, L, 1)).reshape(-1).contiguous()
cen = torch.randint(-8, 9, (phi * n2,), dtype=torch.int64, device=dev)
xs = A[0].clone()
rows = [
    ("trace map_B_to_Bprime x512 (A18)", lambda: g.map_B_to_Bprime(A[2], A[3], phi), 4 * S8),
    ("trace_gemm x512 (A18)", lambda: g.trace_gemm(A[0], A[1], bpr, bpi, phi), 6 * S8),
    ("trace rescale x512 (A18)", lambda: g.rescale_by_delta(cr, ci, 3, 5, 7, phi), 4 * S8),
    ("add_ciphertexts (A17)", lambda: g.add_ciphertexts(ct1, ct2), 6 * S8),
    ("multiply_ciphertexts_raw (A17)", lambda: g.multiply_ciphertexts_raw(ct1, ct2), 7 * S8),
    ("layout matrix->poly (A13)", lambda: g.layout_convert(A[0], S.MATRIX, S.POLY), 2 * S8),
    ("sample_uniform (A14)", lambda: g.sample_uniform(), S8),
    ("sample_gaussian (A14)", lambda: g.sample_gaussian(), S8),
    ("keygen (A14)", lambda: g.generate_secret_key(), 2 * phi * L * n * 8),
    ("decrypt_to_eval 1 ct (A16)", lambda: g.decrypt_to_eval(ct1, sk), 3 * S8),
    ("wntt_forward_vector (A7)", lambda: g.wntt_forward_vector(vec), 2 * phi * L * n * 8),
    ("wntt_forward_centered (A8)", lambda: g.wntt_forward_centered(cen), 2 * phi * n2 * 8),
    ("encode_lane (A10)", lambda: g.encode_lane(lane), n2 * 16 + 2 * L * n2 * 8),
    ("crt_compose_centerlift_big x512 (A11)", lambda: g.crt_compose_centerlift_big(A[0], lanes=phi), S8 + phi * n2 * (g.big_words * 8 + 1)),
]
print(f"native shape: T = {T} u64, S = {S8 / 2**20:.0f} MiB, HBM peak {PEAK} GB/s")
for name, fn, nbytes in rows:
    try:
        ms = t(fn)
        gbs = nbytes / (ms * 1e-3) / 1e9
        print(f"{name:42s} {ms:9.3f} ms  {nbytes / 2**20:8.1f} MiB  {gbs:8.1f} GB/s  {100 * gbs / PEAK:5.1f} % of HBM")
    except Exception as e:                      # keep going: this is a measurement tool
        print(f"{name:42s} failed: {e}")
# n-point X transforms over all phi*L*n rows (A3 GL convention with twist, A4 Phantom order)
for conv, tag in ((m.PHANTOM_NEGACYCLIC, "A4 phantom"), (m.GL_XN_MINUS_I, "A3 gl")):
    try:
        plan = m.NttPlan(6, [int(q) for q in g.copy_device_moduli()], conv)
        x = xs.view(phi * n, L, n)
        ms = t(lambda: (plan.forward_(x), plan.inverse_(x)))
        gbs = 4 * S8 / (ms * 1e-3) / 1e9
        print(f"{'n=64 X NTT+INTT, ' + tag:42s} {ms:9.3f} ms  {4 * S8 / 2**20:8.1f} MiB  {gbs:8.1f} GB/s  {100 * gbs / PEAK:5.1f} % of HBM")
    except Exception as e:
        print(f"n=64 X transforms {tag} failed: {e}")
