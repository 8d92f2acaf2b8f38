#!/bin/bash
# compute-sanitizer over the small-shape part of the parity suite.  memcheck: out-of-bounds / misaligned
# global, shared and local accesses; racecheck: shared-memory hazards (the tcgen05 W-CRT's mbarrier ring and TMEM hand-back,
# the warp-private exchanges of the NTT row kernels, the staged digit tiles of the trace GEMMs).  Selected so that every
# kernel family runs at least once: flat NTT (N = 2^12 .. 2^14), n-point transforms, W-CRT on all three paths (CUDA-core,
# mma.sync, tcgen05 at phi = 512), fused row kernel, fp64 transforms, CRT, bconv (plain + scatter), trace GEMMs, matmul,
# rescale, samplers.  Usage: tools/run_sanitizer.sh memcheck|racecheck <logfile>
set -u
tool=${1:-memcheck}; log=${2:-gpurun_out/sanitizer_${tool}.log}
cd "$(dirname "$0")/.."
SEL='tests/test_sampler.py tests/test_modswitch_gpu.py tests/test_matmul_gpu.py::test_matmul_pieces_bit_exact_vs_oracle tests/test_scheme_gpu.py tests/test_sharded_gpu.py::test_limb_sharded_step_emulated tests/test_sharded_gpu.py::test_partial_scatter_matches_partial_plus_exchange tests/test_ntt_gpu.py::test_flat_forward_inverse_bit_exact tests/test_ntt_gpu.py::test_small_transforms_all_conventions tests/test_ntt_gpu.py::test_mul_intt_fused tests/test_ntt_gpu.py::test_flat_edge_inputs'
extra=""
[ "$tool" = "racecheck" ] && extra="--racecheck-report all"
timeout ${SAN_TIMEOUT:-1500} /usr/local/cuda/bin/compute-sanitizer --tool "$tool" $extra --print-limit 30 --error-exitcode 0 \
  python -m pytest $SEL -m gpu -q -x -p no:cacheprovider > "$log" 2>&1
echo "exit $?" >> "$log"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|exit " "$log" | tail -8
