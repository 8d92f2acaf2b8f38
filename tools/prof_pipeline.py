"""One encode -> encrypt_pair -> decrypt_and_decode at the native shape, for an ncu launch list.  python tools/prof_pipeline.py [L]"""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import matrix_fhe_gpu_b200.scheme as S
L = int(sys.argv[1]) if len(sys.argv) > 1 else 11
extra = [17194410241, 17194804993, 17195989249, 17196581377, 17196976129, 17197963009, 17199936769, 17201318401,
         17201515777, 17203884289, 17206844929, 17207437057, 17209608193]
g = S.Context(moduli=(S.REF_MODULI + extra)[:L])
dev = torch.device("cuda:0")
ell, i = np.meshgrid(np.arange(g.phi, dtype=np.float64), np.arange(g.n2, dtype=np.float64), indexing="ij")
m = torch.from_numpy(((ell + 1e-3 * i) * (1 - 1j)).reshape(-1)).to(dev)
sk = g.generate_secret_key()
for rep in range(2):
    torch.cuda.synchronize(); print("== encode", flush=True)
    re, im = g.encode_to_wntt_eval(m)
    torch.cuda.synchronize(); print("== encrypt", flush=True)
    cr, ci = g.encrypt_pair(re, im, sk)
    torch.cuda.synchronize(); print("== decrypt_decode", flush=True)
    out = g.decrypt_and_decode(cr, ci, sk)
    torch.cuda.synchronize(); print("== encode_encrypt_pair (fused)", flush=True)
    fr, fi = g.encode_encrypt_pair(m, sk)
    torch.cuda.synchronize()
    assert torch.equal(fr, cr) and torch.equal(fi, ci)
print("max_err", float((out - m).abs().max()))
