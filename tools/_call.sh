set -x
mkdir -p gpurun_out
B="python bench.py --no-extras --no-cpu-baseline"
for i in 1 2; do
  timeout 300 $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('default  ', round(d['value'],1), round(d['config']['dependent_chain_GBps'],1), round(d['e2e']['value'],1), d['check']['ok'])"
  timeout 300 $B --opt gemv_resident=3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('resident3', round(d['value'],1), round(d['config']['dependent_chain_GBps'],1), round(d['e2e']['value'],1), d['check']['ok'])"
  timeout 300 $B --opt gemv_resident=3 --opt gemv_stage_bytes=16384 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('res3/16K ', round(d['value'],1), round(d['config']['dependent_chain_GBps'],1), round(d['e2e']['value'],1), d['check']['ok'])"
  timeout 300 $B --opt gemv_resident=3 --opt gemv_stage_bytes=24576 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('res3/24K ', round(d['value'],1), round(d['config']['dependent_chain_GBps'],1), round(d['e2e']['value'],1), d['check']['ok'])"
done 2>&1 | grep -v "^+" | tee gpurun_out/r_resident_ab.txt
