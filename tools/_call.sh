mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_chain.py tests/test_gpu_kquants.py -m gpu -q -x --timeout=120 -k "gemv or chain or fused or layer or kq or matmul" > gpurun_out/n_tests.txt 2>&1; echo "rc=$?" >> gpurun_out/n_tests.txt
tail -3 gpurun_out/n_tests.txt
for v in cur r1 cur r1; do
  if [ $v = cur ]; then unset LK_LIBRARY; else export LK_LIBRARY=$PWD/tools/_ab/lib_$v.so; fi
  timeout 60 python bench.py --no-cpu-baseline --no-extras --steps 200 > gpurun_out/n_bench_$v.json 2>gpurun_out/n_bench_$v.err
  python -c "
import json
d=json.load(open('gpurun_out/n_bench_$v.json')); print('$v', round(d['value'],1), round(d['config']['dependent_chain_GBps'],1), round(d['e2e']['value'],1), d['roofline']['frac'])
" 2>&1 | tail -1
done
