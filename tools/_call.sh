#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/bitnet_unr.txt
run() {
echo "lib [$1] $2 $3" | tee -a gpurun_out/bitnet_unr.txt
LK_LIBRARY=$1 timeout 600 python tools/kq_bench.py --types BITNET_1_58 --shapes 4096x4096 11008x4096 4096x11008 8192x8192 $2 $3 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(' ', d['shape'], d['us_per_launch'], d['GBps'], d['hbm_frac'])
" | tee -a gpurun_out/bitnet_unr.txt
}
run ""
run tools/_ab/lib_bitnet_unr4.so
run tools/_ab/lib_bitnet_unr6.so
run "" --opt gemv_warps=16
run tools/_ab/lib_bitnet_unr4.so --opt gemv_warps=16
