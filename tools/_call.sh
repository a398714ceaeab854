mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -q -x --timeout=120 > gpurun_out/g_tests_all.txt 2>&1; echo "all rc=$?" >> gpurun_out/g_tests_all.txt
tail -8 gpurun_out/g_tests_all.txt
for v in cur r1 nofix cur r1 nofix; do
  if [ $v = cur ]; then unset LK_LIBRARY; else export LK_LIBRARY=$PWD/tools/_ab/lib_$v.so; fi
  timeout 60 python bench.py --no-cpu-baseline --no-extras --steps 200 > gpurun_out/g_bench_$v.json 2>gpurun_out/g_bench_$v.err
  python -c "
import json
d=json.load(open('gpurun_out/g_bench_$v.json')); print('$v', round(d['value'],1), round(d['config']['dependent_chain_GBps'],1), round(d['e2e']['value'],1))
" 2>&1 | tail -1
done
unset LK_LIBRARY
timeout 60 tools/overlap_timeline 24 2 > gpurun_out/g_overlap_timeline.txt 2>&1; tail -3 gpurun_out/g_overlap_timeline.txt
