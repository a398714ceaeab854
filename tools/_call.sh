mkdir -p gpurun_out
timeout 14 python bench.py --no-extras --no-cpu-baseline > gpurun_out/v_bench_spread.json 2>gpurun_out/v_bench_spread.err
tail -c 900 gpurun_out/v_bench_spread.json; tail -2 gpurun_out/v_bench_spread.err
