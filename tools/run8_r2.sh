#!/bin/bash
# N-GPU validation of round 2 (N = number of visible GPUs): config 5 at the full shape with --check on every exchange form
# (every rank's output against the CPU oracle), then the bench under torchrun.
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for ex in ipc p2p nccl; do
  extra=""; [ "$ex" = "p2p" ] && extra="--also-p2p-groups 2"
  timeout 600 $TR --master-port 29611 tools/run_sharded.py --exchange $ex --check $extra --steps 20 2>gpurun_out/sharded${N}_$ex.err | grep -E "^\{|parity" > gpurun_out/sharded${N}_$ex.txt
  echo "== $ex rc=$?"; cat gpurun_out/sharded${N}_$ex.txt | cut -c1-400
done
timeout 900 $TR --master-port 29612 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu > gpurun_out/bench${N}.json 2> gpurun_out/bench${N}.err; echo "bench rc=$?"
tail -c 400 gpurun_out/bench${N}.json
