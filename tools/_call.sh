mkdir -p gpurun_out
timeout 35 python bench.py --no-cpu-baseline > gpurun_out/v_bench_nocpu.json 2>/dev/null
tail -c 1500 gpurun_out/v_bench_nocpu.json
