# W-CRT limiter series: tools/wcrt_dbg_ab.sh <mask>...   (libs libmfhe_dbg<mask>.so built with -DMFHE_UMMA_DBG=<mask>; results are garbage, only time)
for v in "$@"; do echo -n "mask $v: "; MFHE_NO_CHECK=1 MFHE_LIB_PATH=/root/repo/matrix_fhe_gpu_b200/libmfhe_dbg$v.so python tools/prof_wcrt.py 11 4 2>&1 | tail -2 | tr '\n' ' '; echo; done
