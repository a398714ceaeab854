#!/bin/bash
# final evidence of the round: tests, smoke, bench (both arms), ncu launch list + full captures
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r02l_gpu_tests.txt 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r02l_gpu_tests.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02l_smoke.txt 2>&1; echo "smoke rc $?"
timeout 400 python bench.py > gpurun_out/r02l_bench.json 2> gpurun_out/r02l_bench.err; echo "bench rc $?"
timeout 400 python bench.py --impl reference > gpurun_out/r02l_bench_reference_arm.json 2>> gpurun_out/r02l_bench.err; echo "ref arm rc $?"
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/c_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/c_plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/r02l_launches_bench.csv $CMD > gpurun_out/c_ncu1.log 2>&1; echo "ncu list rc $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 300 -c 3 -f -o gpurun_out/r02l_gemv $CMD > gpurun_out/c_ncu2.log 2>&1; echo "ncu gemv rc $?"
LCMD="python tools/layer_bench.py --layers 8 --steps 3"
timeout 300 $LCMD > gpurun_out/c_layer_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemv_mchain -s 6 -c 1 -f -o gpurun_out/r02l_mchain $LCMD > gpurun_out/c_ncu3.log 2>&1; echo "ncu mchain rc $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 60 --csv --log-file gpurun_out/r02l_launches_layer_graph.csv $LCMD > gpurun_out/c_ncu4.log 2>&1; echo "ncu layer list rc $?"
ls -la gpurun_out/*.ncu-rep
