#!/bin/bash
# same-box A/B of library builds on the layer graph: tools/_call.sh "<label>=<lib path or empty>" ...
mkdir -p gpurun_out
out=gpurun_out/r02j_layer_ab.txt
: > $out
for rep in 1 2; do
for spec in "$@"; do
  label=${spec%%=*}; lib=${spec#*=}
  for ctx in 0 2048; do
    echo -n "$label ctx $ctx: " >> $out
    LK_LIBRARY=$lib timeout 300 python tools/layer_bench.py --ctx $ctx --opt max_chain=8 2>&1 | python -c "import sys,json
for l in sys.stdin:
    try:
        d=json.loads(l); print(round(d['us_per_layer'],2), 'us/layer', round(d['tok_s'],1), 'tok/s', round(d['hbm_frac'],3), d['finite_output'])
    except Exception: print(l.rstrip()[-200:])" >> $out
  done
done
done
cat $out
