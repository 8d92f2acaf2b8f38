for v in "$@"; do MFHE_LIB_PATH=/root/repo/matrix_fhe_gpu_b200/libmfhe_$v.so python bench.py --steps 30 --warmup 3 --no-secondary --no-cpu --no-baselines --no-config5 --e2e-steps 1 > gpurun_out/bench_$v.log 2>&1; python - "$v" <<'PY' || tail -5 gpurun_out/bench_$v.log
import json,sys
v=sys.argv[1]
l=[x for x in open(f"gpurun_out/bench_{v}.log") if x.startswith("{")]
d=json.loads(l[-1]); print(v, round(d["value"]), "fwd", round(d["roofline"]["ms_forward"],3), "inv", round(d["roofline"]["ms_inverse"],3))
PY
done
