timeout 100 python tools/ffn70b_bench.py 2> gpurun_out/ffn_g1.err | tee gpurun_out/ffn70b.jsonl
for g in 2 4 8; do for ex in push nccl; do
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2953$g tools/ffn70b_bench.py --exchange $ex 2> gpurun_out/ffn_g${g}_$ex.err | tee -a gpurun_out/ffn70b.jsonl; echo "g$g $ex rc=$?"
done; done
