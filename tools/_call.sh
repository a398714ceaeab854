timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "prefill_gemm_tc" > gpurun_out/pytest_k2.log 2>&1; echo k2=$?; tail -5 gpurun_out/pytest_k2.log
for c in 1 2 4; do timeout 300 python tools/gemm_tc_bench.py --qtype Q8_0 --cluster $c --shapes 11008x4096x512 4096x4096x512 4096x11008x512 2>&1 | tail -3; done | tee gpurun_out/k2_bench.log
timeout 300 python tools/gemm_tc_bench.py --qtype Q4_0 --shapes 11008x4096x512 2>&1 | tail -1 | tee -a gpurun_out/k2_bench.log
CMD="python tools/gemm_tc_bench.py --qtype Q8_0 --shapes 11008x4096x512 --reps 1 --rotate-mb 100"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --warp-sampling-interval 0 --clock-control none --import-source on -k regex:gemm_tc_kernel -s 4 -c 1 -o gpurun_out/gemm_tc_r01g $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
