set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_kquants.py -x -q 2>&1 | tail -8 > gpurun_out/f_tests.txt
cat gpurun_out/f_tests.txt
B="python bench.py --no-extras --no-cpu-baseline"
for i in 1 2; do
  timeout 300 $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('fold', round(d['value'],1), round(d['config']['dependent_chain_GBps'],1), round(d['e2e']['value'],1), d['check'])"
done
timeout 300 python tools/layer_bench.py --ctx 0 2>&1 | tail -1 | cut -c300-420
timeout 300 python tools/kq_bench.py --types Q4_K --shapes 4096x4096 11008x4096 2>/dev/null | cut -c1-160
