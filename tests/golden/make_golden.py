"""Generates tests/golden/phantom_tables.json from the REAL Phantom host code.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
It loads oracle/_ref/libphantom_host.so (ph/src/host/*.cu compiled unmodified, see
oracle/Makefile) and records, for a set of (logN, modulus) pairs, the root, n^-1, a few
table entries and a SHA-256 of each full table, plus the prime lists of the BASELINE
configs.  The oracle and the CUDA plan builder are both checked against this file.
"""
import ctypes as C
import hashlib
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import oracle as O  # noqa: E402

R = O.ref()
assert R is not None, "oracle/_ref not built (make -C oracle ref)"


def primes(n, bits, count):
    out = np.zeros(count, dtype=np.uint64)
    assert R.ref_get_primes(n, bits, count, out) == 0
    return [int(x) for x in out]


def create(n, bits, count):
    out = np.zeros(count, dtype=np.uint64)
    assert R.ref_coeff_modulus_create(n, bits, count, out) == 0
    return [int(x) for x in out]


def tables(logn, q):
    n = 1 << logn
    a = [np.zeros(n, dtype=np.uint64) for _ in range(4)]
    root, ni, nis = C.c_uint64(), C.c_uint64(), C.c_uint64()
    assert R.ref_ntt_tables(logn, q, C.byref(root), a[0], a[1], a[2], a[3], C.byref(ni), C.byref(nis)) == 0
    rec = {"logn": logn, "q": q, "root": root.value, "n_inv": ni.value, "n_inv_shoup": nis.value}
    for name, arr in zip(("tw", "tw_shoup", "itw", "itw_shoup"), a):
        rec[name + "_sha256"] = hashlib.sha256(arr.tobytes()).hexdigest()
        rec[name + "_head"] = [int(x) for x in arr[:8]]
    return rec


REF_MODULI = [17592186435073, 17182765057, 17184541441, 17186120449, 17186515201, 17186909953,
              17188883713, 17190462721, 17190857473, 17191844353, 17192831233]
gold = {
    "source": "extern/phantom-fhe @ 1f4a198 host code via oracle/_ref/libphantom_host.so",
    "get_primes_4096_60_1": primes(4096, 60, 1),
    "create_65536_60x32": create(65536, 60, 32),
    "create_65536_50x32": create(65536, 50, 32),
    "get_primes_131072_55_64": primes(131072, 55, 64),
    "tables": [tables(6, q) for q in REF_MODULI]
    + [tables(12, primes(4096, 60, 1)[0])]
    + [tables(16, q) for q in create(65536, 60, 32)[:2]]
    + [tables(17, primes(131072, 55, 1)[0])],
}
with open(os.path.join(os.path.dirname(__file__), "phantom_tables.json"), "w") as f:
    json.dump(gold, f, indent=1)
print("wrote", len(gold["tables"]), "table records")
