timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fused or decoder_layer" 2>&1 | tail -12
R=$PWD
for rep in 1 2; do
for c in 43123ce cur; do
  if [ $c = cur ]; then d=$R; else d=$R/tools/_ab/$c; fi
  (cd $d && timeout 300 python bench.py --no-extras 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$c', round(d['value'],1), d['config'].get('dependent_chain_GBps'), round(d['e2e']['value'],1), d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'))")
done; done
for o in "" "--opt nofuse=1"; do timeout 300 python tools/layer_bench.py $o 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('layer $o', round(d['us_per_layer'],2), round(d['tok_s'],1))"; done
