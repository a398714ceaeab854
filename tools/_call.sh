set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "swiglu_sharded" 2>&1 | tail -5
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
rm -f gpurun_out/t_ffn70b.jsonl
timeout 400 $TR --nproc-per-node 2 --master-port 29621 tools/ffn70b_bench.py 2>gpurun_out/t_err1.txt | tail -1 >> gpurun_out/t_ffn70b.jsonl
timeout 400 $TR --nproc-per-node 2 --master-port 29622 tools/ffn70b_bench.py --fused 0 2>gpurun_out/t_err2.txt | tail -1 >> gpurun_out/t_ffn70b.jsonl
tail -5 gpurun_out/t_err1.txt
python -c "
import json
for l in open('gpurun_out/t_ffn70b.jsonl'):
    d=json.loads(l); print(d['n_gpus'], round(d['us_per_ffn'],1), round(d['aggregate_weight_GBps']), d['exchanges'][:40], d.get('spot_check_rel_err'), d.get('all_ranks_identical_y'), d.get('barrier_timed_out'))"
