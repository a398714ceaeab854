set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_kquants.py -x -q 2>&1 | tail -15 > gpurun_out/a_tests.txt
cat gpurun_out/a_tests.txt
for c in 0 128 2048; do timeout 300 python tools/layer_bench.py --ctx $c 2>&1 | tail -1; done > gpurun_out/a_layer.jsonl
cat gpurun_out/a_layer.jsonl
timeout 300 python tools/attn_bench.py 2>&1 | tail -6 > gpurun_out/a_attn.jsonl; cat gpurun_out/a_attn.jsonl
timeout 600 python tools/kq_bench.py > gpurun_out/a_kq_bench.jsonl 2> gpurun_out/a_kq_bench.err; cat gpurun_out/a_kq_bench.jsonl
