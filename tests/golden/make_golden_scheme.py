"""Generates tests/golden/scheme_constants.json from the reference's own source text.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden_scheme.py
Records the literal k_wntt_exp table (src/core/HE.cu:72-105) and the parameter constants of
include/core/config.h, so that the oracle and the CUDA context are pinned to what the reference ships
rather than to a re-derivation.
"""
import hashlib
import json
import os
import re

REF = "/root/reference"
he = open(os.path.join(REF, "src/core/HE.cu")).read()
cfg = open(os.path.join(REF, "include/core/config.h")).read()
tab = [int(x) for x in re.findall(r"\d+", re.search(r"k_wntt_exp\[BATCH_SIZE\] = \{(.*?)\};", he, re.S).group(1))]
mod = [int(x) for x in re.findall(r"(\d+)ULL", re.search(r"RNS_MODULI\[RNS_NUM_LIMBS\] = \{(.*?)\};", cfg, re.S).group(1))]
pmod = [int(x) for x in re.findall(r"(\d+)ULL", re.search(r"P_MODULI\[P_NUM_LIMBS\] = \{(.*?)\};", cfg, re.S).group(1))]
gold = {
    "source": "Shaibk/Matrix-FHE-GPU @ 00889fc: src/core/HE.cu:72-105, include/core/config.h:7-52",
    "k_wntt_exp": tab,
    "k_wntt_exp_sha256": hashlib.sha256(",".join(map(str, tab)).encode()).hexdigest(),
    "RNS_MODULI": mod, "P_MODULI": pmod,
    "MATRIX_N": int(re.search(r"MATRIX_N = (\d+)", cfg).group(1)),
    "BATCH_SIZE": int(re.search(r"BATCH_SIZE = (\d+)", cfg).group(1)),
    "BATCH_PRIME_P": int(re.search(r"BATCH_PRIME_P = (\d+)", cfg).group(1)),
    "SCALING_FACTOR": float(re.search(r"SCALING_FACTOR = ([\d.]+)", cfg).group(1)),
}
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "scheme_constants.json")
json.dump(gold, open(out, "w"), indent=1)
print("wrote", out)
