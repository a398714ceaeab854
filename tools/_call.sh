set -x
mkdir -p gpurun_out
rm -f gpurun_out/w_*.txt
timeout 300 python -m pytest tests/test_gpu_chain.py -x -q 2>&1 | tail -15 > gpurun_out/w_tests.txt
for c in 0 128 512 2048; do
for o in "max_chain=8" "max_chain=1"; do
  timeout 200 python tools/layer_bench.py --ctx $c --opt $o 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$o ctx$c', round(d['us_per_layer'],2), round(d['tok_s'],1), d['finite_output'])" >> gpurun_out/w_layer.txt
done
done
