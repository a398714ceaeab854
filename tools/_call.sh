# scratch driver for one gpurun call (refresh of tests / bench / ncu evidence at HEAD)
set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/l_tests.txt
timeout 600 python bench.py > gpurun_out/l_bench.json 2> gpurun_out/l_bench.err
timeout 300 python tools/layer_bench.py 2>&1 | tail -1 > gpurun_out/l_layer.json
timeout 300 python tools/gemm_tc_bench.py --qtype Q8_0 2>&1 | tail -3 > gpurun_out/l_gemm.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/l_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/l_ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 300 -c 3 -o gpurun_out/l_gemv_full -f python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/l_ncu_full.log 2>&1
cat gpurun_out/l_tests.txt; cat gpurun_out/l_bench.json; cat gpurun_out/l_layer.json; cat gpurun_out/l_gemm.txt
