#!/bin/bash
# Regenerates tests/golden/refcuda_golden.jsonl: outputs of the UNMODIFIED reference (and Phantom) run on a B200.
#   1. here (no GPU needed):  python baseline/build_ref.py        # nvcc: /root/reference -> baseline/_ref/{ref_dump,ref_dump_stock}
#   2. on the GPU box:        gpurun -- bash tests/golden/make_golden_refcuda.sh run
#   3. here:                  bash tests/golden/make_golden_refcuda.sh merge
# "rdc" records come from the -rdc=true build (one instance of the cross-TU constant tables), "stock_*" records from the
# build with the reference's own CMake flags (its fp64 decode only works there: under -rdc the reference's upload of
# d_crt_M fails with cudaErrorInvalidValue and decode returns zeros -- see DESIGN.md section 2).
set -e
cd "$(dirname "$0")/../.."
case "${1:-run}" in
  run)
    mkdir -p gpurun_out
    baseline/_ref/ref_dump gpurun_out/refcuda_golden.jsonl 2> gpurun_out/ref_dump.log
    baseline/_ref/ref_dump_stock gpurun_out/refcuda_golden_stock.jsonl 2> gpurun_out/ref_dump_stock.log ;;
  merge)
    python - <<'PY'
import json
seen, out = set(), []
for f in ("gpurun_out/refcuda_golden.jsonl", "gpurun_out/refcuda_golden_stock.jsonl"):
    for line in open(f):
        line = line.strip()
        if line and json.loads(line)["name"] not in seen | {"stock_secret_key"}:
            seen.add(json.loads(line)["name"]); out.append(line)
open("tests/golden/refcuda_golden.jsonl", "w").write("\n".join(out) + "\n")
PY
    ;;
esac
