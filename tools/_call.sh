timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "prefill_gemm_tc" > gpurun_out/pytest_k2.log 2>&1; echo k2=$?; tail -3 gpurun_out/pytest_k2.log
for c in 1 2; do timeout 300 python tools/gemm_tc_bench.py --qtype Q8_0 --cluster $c --shapes 11008x4096x512 4096x4096x512 4096x11008x512 2>&1 | tail -3; done | tee gpurun_out/k2_bench.log
timeout 300 python tools/gemm_tc_bench.py --qtype Q4_0 --shapes 11008x4096x512 4096x4096x128 4096x4096x2048 2>&1 | tail -3 | tee -a gpurun_out/k2_bench.log
timeout 300 python tools/gemm_tc_bench.py --qtype Q4_1 --shapes 11008x4096x512 2>&1 | tail -1 | tee -a gpurun_out/k2_bench.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_r01c.json 2> gpurun_out/bench_r01c.err; echo bench=$?; cut -c1-900 gpurun_out/bench_r01c.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref=$?; cut -c1-500 gpurun_out/bench_ref.json
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
