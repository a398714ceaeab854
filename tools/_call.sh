R=$PWD
run() { d=$1; shift; (cd $d && timeout 300 python bench.py --no-extras --no-cpu-baseline "$@" 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$d $*', round(d['value'],1), d['config'].get('dependent_chain_GBps'), round(d['e2e']['value'],1), d['clocks']['sm_mhz'])"); }
for rep in 1 2; do
run $R
run $R/tools/_ab/tail2
run $R --opt gemv_stage_bytes=28672
run $R --opt gemv_warps=12
run $R/tools/_ab/tail2 --opt gemv_warps=12
done
timeout 300 python tools/layer_bench.py 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('layer', round(d['us_per_layer'],2), round(d['tok_s'],1))"
(cd tools/_ab/tail2; timeout 300 python tools/layer_bench.py 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('layer tail2', round(d['us_per_layer'],2), round(d['tok_s'],1))")
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
