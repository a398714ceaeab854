timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo pytest=$?; tail -5 gpurun_out/pytest_gpu.log
python bench.py --no-cpu-baseline > gpurun_out/bench_overlap.log 2>&1; tail -1 gpurun_out/bench_overlap.log | cut -c1-400
python bench.py --no-cpu-baseline --opt gemv_overlap=0 > gpurun_out/bench_chain.log 2>&1; tail -1 gpurun_out/bench_chain.log | cut -c1-400
python tools/gemv_sweep.py --qtype Q4_0 --opts "" "gemv_overlap=0" > gpurun_out/sweep4.jsonl 2>gpurun_out/sweep4.err; cat gpurun_out/sweep4.jsonl
python tools/gemv_sweep.py --qtype Q8_0 --opts "" "gemv_overlap=0" >> gpurun_out/sweep4.jsonl 2>>gpurun_out/sweep4.err; python tools/gemv_sweep.py --qtype Q4_1 --opts "" "gemv_overlap=0" >> gpurun_out/sweep4.jsonl 2>>gpurun_out/sweep4.err; tail -12 gpurun_out/sweep4.jsonl
