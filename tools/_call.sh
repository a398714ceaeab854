mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_chain.py -m gpu -q -x --timeout=120 -k "fused or layer or rms_norm or softmax or chain" > gpurun_out/r_tests.txt 2>&1; echo "rc=$?" >> gpurun_out/r_tests.txt; tail -3 gpurun_out/r_tests.txt | cut -c1-300
for o in "" "--opt gemv_ctas_per_sm=2" "--opt gemv_warps=16" "--opt gemv_ctas_per_sm=2 --opt gemv_smem_budget=57344"; do
  echo "== layer_bench $o"; timeout 120 python tools/layer_bench.py --steps 30 $o 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['tok_s'],1), round(d['us_per_layer'],2), d['launches_per_token'])"
done
echo "== ctx 2048"; timeout 120 python tools/layer_bench.py --steps 30 --ctx 2048 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['tok_s'],1), round(d['us_per_layer'],2))"
LK_LIBRARY=$PWD/tools/_ab/lib_r1.so timeout 120 python tools/layer_bench.py --steps 30 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('r1', round(d['tok_s'],1), round(d['us_per_layer'],2))"
timeout 100 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); e=d['config']['other_paths']['elementwise']; print({k:(round(v['GBps']),round(v['roofline']['frac'],3)) for k,v in e.items()})"
