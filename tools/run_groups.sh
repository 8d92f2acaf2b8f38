#!/bin/bash
# config 5 through the C ABI (mfhe_bconv_sharded_step) with 1, 2, 4 batch slices per step, N = visible GPUs, parity checked
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29613 tools/run_sharded.py --exchange ipc --check --ipc-groups 2,4,8,1 --steps 30 2>gpurun_out/groups${N}.err | grep -E "^\{|parity" > gpurun_out/groups${N}.txt
echo "rc=$?"; cut -c1-330 gpurun_out/groups${N}.txt
