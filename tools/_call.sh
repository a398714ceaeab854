#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/c_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/c_plain.log; exit 1; }
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 300 -c 3 -f -o gpurun_out/r02h_gemv $CMD > gpurun_out/c_ncu.log 2>&1
echo "ncu rc $?"; ls -la gpurun_out/r02h_gemv.ncu-rep; tail -3 gpurun_out/c_ncu.log
