# scratch driver for one gpurun call
set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/m_tests.txt
cat gpurun_out/m_tests.txt
timeout 300 python tools/attn_bench.py 2>&1 | tail -6 > gpurun_out/m_attn.jsonl
cat gpurun_out/m_attn.jsonl
for c in 0 128 2048; do timeout 300 python tools/layer_bench.py --ctx $c 2>&1 | tail -1; done > gpurun_out/m_layer.jsonl
cat gpurun_out/m_layer.jsonl
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 288 -c 400 --csv --log-file gpurun_out/m_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/m_ncu_launch.log 2>&1
tail -3 gpurun_out/m_launches_bench.csv | cut -c1-300
