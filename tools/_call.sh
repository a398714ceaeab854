mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/b_tests_all.txt 2>&1; echo "all rc=$?" >> gpurun_out/b_tests_all.txt
tail -25 gpurun_out/b_tests_all.txt
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b_bench.json 2>gpurun_out/b_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/b_bench.json; tail -5 gpurun_out/b_bench.err
timeout 100 python bench.py --no-cpu-baseline --no-extras --opt gemv_fix_gap=255 > gpurun_out/b_bench_nofix.json 2>gpurun_out/b_bench_nofix.err; python -c "
import json
for f in ('b_bench','b_bench_nofix'):
    d=json.load(open('gpurun_out/'+f+'.json')); print(f, d['value'], d['config']['dependent_chain_GBps'], d['e2e']['value'])
"
