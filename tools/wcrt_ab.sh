# W-CRT build variants: tools/wcrt_ab.sh <name>...  (libmfhe_<name>.so; b200 = the product build, checked; dbg* are timing-only)
for v in "$@"; do echo -n "$v: "; chk=1; [ "$v" = b200 ] || [ "$v" = dbgold ] && chk=""; MFHE_NO_CHECK=$chk MFHE_LIB_PATH=/root/repo/matrix_fhe_gpu_b200/libmfhe_$v.so timeout 120 python tools/prof_wcrt.py 11 4 2>&1 | tail -3 | tr '\n' ' '; echo; done
