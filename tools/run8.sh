set -x
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/run_sharded.py --steps 50 --warmup 5 --exchange ${EX:-p2p} $( [ "${EX:-p2p}" = p2p ] && echo --also-p2p-groups 2 ) > gpurun_out/sharded_${N}.json 2> gpurun_out/sharded_${N}.err; tail -c 500 gpurun_out/sharded_${N}.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/run_sharded.py --steps 2 --warmup 1 --batch 2 --check --exchange ${EX:-p2p} > gpurun_out/sharded_${N}_check.json 2> gpurun_out/sharded_${N}_check.err; tail -c 400 gpurun_out/sharded_${N}_check.json; tail -3 gpurun_out/sharded_${N}_check.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_${N}gpu_r1_16.json 2> gpurun_out/bench_${N}gpu_r1_16.err; tail -c 900 gpurun_out/bench_${N}gpu_r1_16.json
