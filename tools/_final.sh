#!/bin/bash
# last check of the round's binary: tests, smoke, ncu capture for roofline.traffic, bench
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r02n_gpu_tests.txt 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r02n_gpu_tests.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02n_smoke.txt 2>&1; echo "smoke rc $?"
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/c_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 300 -c 3 -f -o gpurun_out/r02n_gemv $CMD > gpurun_out/c_ncu2.log 2>&1; echo "ncu gemv rc $?"
