#!/bin/bash
# A/B of the three exchange forms of the limb-sharded step (BASELINE config 5) on NG GPUs of one box.
#   tools/run_exchange_ab.sh NG OUTFILE
NG=${1:-2}; OUT=${2:-gpurun_out/exchange_ab.txt}
mkdir -p "$(dirname "$OUT")"; : > "$OUT"
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NG" --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) tools/run_sharded.py "$@" 2>gpurun_out/exchange_err.log | grep -E '^\{|parity' >> "$OUT" || { echo "FAILED: $*" >> "$OUT"; tail -15 gpurun_out/exchange_err.log >> "$OUT"; }; }
for ex in nccl pipelined p2p; do
  echo "## parity, N=2^13 12->8 limbs batch 4, exchange=$ex" >> "$OUT"
  run --logn 13 --lin 12 --lout 8 --batch 4 --check --steps 3 --exchange $ex
done
for ex in nccl pipelined p2p; do
  echo "## config 5 (N=2^17, 48->16 limbs, batch 16), exchange=$ex" >> "$OUT"
  run --steps 50 --warmup 5 --exchange $ex
done
echo "## config 5, pipelined groups=2" >> "$OUT"; run --steps 50 --warmup 5 --exchange pipelined --groups 2
echo "## config 5 with parity check, exchange=p2p" >> "$OUT"; run --steps 5 --check --exchange p2p
cat "$OUT"
