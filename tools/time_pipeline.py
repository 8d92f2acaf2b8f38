"""Per-stage device timings of the matrix pipeline at the native shape (CUDA events, median of N).
   python tools/time_pipeline.py [L]      (L = 11 native; 24 = BASELINE config 4 moduli)"""
import statistics
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import matrix_fhe_gpu_b200.scheme as S

L = int(sys.argv[1]) if len(sys.argv) > 1 else 11
mods = None
if L != 11:
    extra = [17194410241, 17194804993, 17195989249, 17196581377, 17196976129, 17197963009, 17199936769, 17201318401,
             17201515777, 17203884289, 17206844929, 17207437057, 17209608193]
    mods = (S.REF_MODULI + extra)[:L]
g = S.Context(moduli=mods)
dev = torch.device("cuda:0")
ell, i = np.meshgrid(np.arange(g.phi, dtype=np.float64), np.arange(g.n2, dtype=np.float64), indexing="ij")
m = torch.from_numpy(((ell + 1e-5 * i) + 1j * (ell - 1e-5 * i)).reshape(-1)).to(dev)
sk = g.generate_secret_key()
re, im = g.encode_to_wntt_eval(m)
cr, ci = g.encrypt_pair(re, im, sk)
out = g.decrypt_and_decode(cr, ci, sk)
print("max_err", float((out - m).abs().max()))


def t(fn, n=10):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for _ in range(n):
        ev[0].record(); fn(); ev[1].record(); torch.cuda.synchronize()
        ts.append(ev[0].elapsed_time(ev[1]))
    return statistics.median(ts)


a = g.sample_uniform()
S_bytes = g.T * 8
rows = [
    ("wcrt_forward (1 poly set)", lambda: g.wntt_forward_matrix(a)),
    ("wcrt_inverse", lambda: g.wntt_inverse_matrix(a)),
    ("encode", lambda: g.encode_to_wntt_eval(m)),
    ("encrypt_pair", lambda: g.encrypt_pair(re, im, sk, cr, ci)),
    ("encode_encrypt_pair (fused)", lambda: g.encode_encrypt_pair(m, sk, cr, ci)),
    ("decrypt_to_eval (1 ct)", lambda: g.decrypt_to_eval(cr, sk)),
    ("decrypt_and_decode", lambda: g.decrypt_and_decode(cr, ci, sk, out)),
    ("keygen", lambda: g.generate_secret_key()),
    ("idft2 x512", lambda: g.idft2(m, g.phi)),
    ("wdft complex", lambda: g.wdft_forward_complex(m)),
]
print(f"L={L} S={S_bytes / 2**20:.0f} MiB")
for name, fn in rows:
    fn(); torch.cuda.synchronize()
    print(f"{name:30s} {t(fn):9.3f} ms")
