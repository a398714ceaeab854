set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kquants.py -x -q 2>&1 | tail -8 > gpurun_out/k_tests.txt
cat gpurun_out/k_tests.txt
timeout 600 python tools/kq_bench.py --types Q8_K Q4_K Q2_K > gpurun_out/k_bench.jsonl 2> gpurun_out/k_bench.err; cut -c1-200 gpurun_out/k_bench.jsonl; tail -3 gpurun_out/k_bench.err
