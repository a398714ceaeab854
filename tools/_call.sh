set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py -x -q 2>&1 | tail -15 > gpurun_out/b_tests.txt
cat gpurun_out/b_tests.txt
timeout 300 python tools/attn_bench.py 2>&1 | tail -6 > gpurun_out/b_attn.jsonl; cut -c1-200 gpurun_out/b_attn.jsonl
for c in 0 128 2048; do timeout 300 python tools/layer_bench.py --ctx $c 2>&1 | tail -1; done > gpurun_out/b_layer.jsonl
cut -c300-520 gpurun_out/b_layer.jsonl
