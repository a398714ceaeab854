#!/bin/bash
# layer graph per block type and chaining mode (one box): profiles/r02m_layer_types.txt
mkdir -p gpurun_out
out=gpurun_out/r02m_layer_types.txt
: > $out
for q in Q4_0 Q4_1 Q8_0; do
  for o in "" "--opt max_chain=1" "--opt max_chain=8 --opt gemv_chain_mma=0" "--opt max_chain=8 --opt gemv_chain_mma=1" "--opt gemv_fix_gap=255" "--opt max_chain=1 --opt gemv_fix_gap=255"; do
    echo -n "$q ctx 0 [$o]: " >> $out
    timeout 200 python tools/layer_bench.py --qtype $q $o 2>&1 | python -c "import sys,json
for l in sys.stdin:
    try:
        d=json.loads(l); print(round(d['us_per_layer'],2), 'us/layer', round(d['tok_s'],1), 'tok/s', round(d['hbm_frac'],3), 'of the HBM bound; launches', d['launches_per_token'], 'chained', d['chained_gemv_launches_per_token'], 'on K1m', d['of_those_on_the_mma_chain_kernel'], d['finite_output'])
    except Exception: print(l.rstrip()[-160:])" >> $out
  done
done
cat $out
