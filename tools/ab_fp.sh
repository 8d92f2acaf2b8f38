# A/B of build variants of the FP64-pipe NTT kernels: tools/ab_fp.sh <variant>...  (libs built by tools/build_variants.sh)
for v in "$@"; do echo -n "$v: "; MFHE_LIB_PATH=/root/repo/matrix_fhe_gpu_b200/libmfhe_$v.so python tools/time_ntt50.py 2>&1 | tail -1; done
