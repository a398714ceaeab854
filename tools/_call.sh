timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "fused or decoder_layer" > gpurun_out/pytest_fused.log 2>&1; echo fused=$?; tail -12 gpurun_out/pytest_fused.log
for o in "" "--opt nofuse=1"; do timeout 300 python tools/layer_bench.py $o 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['us_per_layer'],2),'us/layer', round(d['tok_s'],1))"; done
