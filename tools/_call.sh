set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "swiglu_sharded" 2>&1 | tail -2
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
rm -f gpurun_out/u_ffn70b.jsonl
for n in 8 4; do
  timeout 300 $TR --nproc-per-node $n --master-port 2963$n tools/ffn70b_bench.py 2>gpurun_out/u_err_$n.txt | tail -1 >> gpurun_out/u_ffn70b.jsonl
done
timeout 300 $TR --nproc-per-node 8 --master-port 29641 tools/ffn70b_bench.py --fused 0 2>/dev/null | tail -1 >> gpurun_out/u_ffn70b.jsonl
python -c "
import json
for l in open('gpurun_out/u_ffn70b.jsonl'):
    d=json.loads(l); print(d['n_gpus'], round(d['us_per_ffn'],1), round(d['aggregate_weight_GBps']), d['exchanges'][:40], d.get('spot_check_rel_err'), d.get('all_ranks_identical_y'), d.get('barrier_timed_out'))"
