#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/gemv_ab5.txt
: > $OUT
for rep in 1 2; do
for lib in "" tools/_ab/lib_gemv_nodetect.so tools/_ab/lib_gemv_nosites.so tools/_ab/lib_nofix.so; do
  echo "bench lib=$lib" | tee -a $OUT
  LK_LIBRARY=$lib timeout 600 python bench.py --steps 30 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['unit'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'])" | tee -a $OUT
done
done
