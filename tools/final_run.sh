set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_r1_15.log 2>&1; tail -3 gpurun_out/pytest_r1_15.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r1_15.log 2>&1; tail -2 gpurun_out/smoke_r1_15.log
python bench.py > gpurun_out/bench_r1_15.json 2> gpurun_out/bench_r1_15.err; tail -c 600 gpurun_out/bench_r1_15.json
python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref_r1_15.json 2>&1; tail -c 400 gpurun_out/bench_ref_r1_15.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r1_15.csv python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 1 > gpurun_out/ncu_bench_r1_15.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/pipe_launches_r1_15.csv python tools/prof_pipeline.py 11 > gpurun_out/ncu_pipe_r1_15.log 2>&1
wc -l gpurun_out/launches_r1_15.csv gpurun_out/pipe_launches_r1_15.csv
