mkdir -p gpurun_out
# new parity cases first (each file under its own timeout: a hung kernel must not take the box down)
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tc or range or dynamic" > gpurun_out/a_tests_new.txt 2>&1; echo "new rc=$?" >> gpurun_out/a_tests_new.txt
tail -15 gpurun_out/a_tests_new.txt
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/a_tests_all.txt 2>&1; echo "all rc=$?" >> gpurun_out/a_tests_all.txt
tail -25 gpurun_out/a_tests_all.txt
for p in 0 1 2; do timeout 120 python tools/gemm_tc_bench.py --qtype Q8_0 --precision $p --shapes 11008x4096x512 4096x4096x2048 >> gpurun_out/a_gemm_tc.jsonl 2>gpurun_out/a_gemm_tc.err; done
timeout 60 python tools/gemm_tc_bench.py --qtype Q4_0 --shapes 11008x4096x512 >> gpurun_out/a_gemm_tc.jsonl 2>>gpurun_out/a_gemm_tc.err
timeout 60 python tools/gemm_tc_bench.py --qtype Q4_1 --shapes 11008x4096x512 >> gpurun_out/a_gemm_tc.jsonl 2>>gpurun_out/a_gemm_tc.err
cat gpurun_out/a_gemm_tc.jsonl
timeout 200 python bench.py --no-cpu-baseline > gpurun_out/a_bench.json 2>gpurun_out/a_bench.err; tail -c 1500 gpurun_out/a_bench.json; tail -3 gpurun_out/a_bench.err
