#!/bin/bash
# p2p exchange with the batch in stream groups (tools/run_sharded.py --exchange p2p --p2p-groups G) on NG GPUs
NG=${1:-2}; OUT=${2:-gpurun_out/p2p_groups.txt}
mkdir -p "$(dirname "$OUT")"; : > "$OUT"
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NG" --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) tools/run_sharded.py "$@" 2>gpurun_out/p2p_err.log | grep -E '^\{|parity' >> "$OUT" || { echo "FAILED: $*" >> "$OUT"; tail -15 gpurun_out/p2p_err.log >> "$OUT"; }; }
echo "## config 5 parity, p2p groups=2" >> "$OUT"; run --steps 4 --check --exchange p2p --p2p-groups 2
for g in 1 2 4; do echo "## config 5, p2p groups=$g" >> "$OUT"; run --steps 50 --warmup 5 --exchange p2p --p2p-groups $g; done
grep -E "^##|parity|FAILED" "$OUT"; grep -o '"ms_per_step": [0-9.]*' "$OUT"
