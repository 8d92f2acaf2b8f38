#!/bin/bash
# Round-2 validation + measurement on one B200: GPU parity suite, smoke, both bench arms, the ncu launch list of the bench
# command and one `ncu --set full` capture of the flat-NTT kernels (summaries are copied to profiles/ by hand).
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_r2.log 2>&1; tail -3 gpurun_out/pytest_r2.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r2.log 2>&1; tail -2 gpurun_out/smoke_r2.log
python bench.py > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err; tail -c 300 gpurun_out/bench_r2.json
python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref_r2.json 2>&1; tail -c 300 gpurun_out/bench_ref_r2.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_r2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-baselines --e2e-steps 1 > gpurun_out/ncu_bench_r2.log 2>&1
wc -l gpurun_out/launches_r2.csv
ncu --set full --clock-control none --import-source on -k regex:"ntt_(cols|rows)" -s 12 -c 4 -o gpurun_out/prof_ntt_r2 -f \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-baselines --no-secondary --no-config5 --e2e-steps 1 > gpurun_out/ncu_full_r2.log 2>&1
ls -la gpurun_out/prof_ntt_r2.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:ct_matmul_umma -c 2 -o gpurun_out/prof_matmul_r2 -f \
    python tools/time_matmul.py > gpurun_out/ncu_matmul_r2.log 2>&1
ls -la gpurun_out/prof_matmul_r2.ncu-rep
