#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/io_sched.txt
for c in "1,3,12,16" "1,1,2,4,8,16" "4,28" "1,3,28" "1,2,5,24" "2,30"; do
timeout 600 python bench.py --steps 50 --warmup 5 --no-extras --no-cpu-baseline --io-schedule $c 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('sched $c value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],4), 'serial', round(d['e2e']['serial_copies']['value'],1), round(d['e2e']['serial_copies']['ms_per_step'],4), 'dev ms', round(d['ms_per_step'],4))
" | tee -a gpurun_out/io_sched.txt
done
