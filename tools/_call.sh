mkdir -p gpurun_out
timeout 420 python -m pytest tests/test_gpu_multi.py -m gpu -q -x --timeout=300 > gpurun_out/k_tests_multi.txt 2>&1; echo "multi rc=$?" >> gpurun_out/k_tests_multi.txt
tail -6 gpurun_out/k_tests_multi.txt | cut -c1-600
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 3 --no-extras > gpurun_out/k_bench_n2.json 2>gpurun_out/k_bench_n2.err; echo "bench2 rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/k_bench_n2.json')); print('n2', round(d['value'],1), d['config']['us_per_ffn'], d['config']['launches_per_step'], d['e2e']['value'], d['config']['other_exchanges'])
"
