set -x
python tools/gemv_sweep.py --qtype Q4_0 --opts "" "gemv_pdl=0" "gemv_warps=16" "gemv_warps=4" "gemv_stage_bytes=8192" "gemv_stage_bytes=40000,gemv_smem_budget=200000" "gemv_smem_budget=200000" "gemv_ctas_per_sm=2,gemv_smem_budget=100000" "gemv_ctas_per_sm=2,gemv_smem_budget=56000,gemv_warps=4" > gpurun_out/sweep1.jsonl 2> gpurun_out/sweep1.err
python tools/gemv_sweep.py --qtype Q4_0 --no-graph --opts "" "gemv_pdl=2" > gpurun_out/sweep1_nograph.jsonl 2>> gpurun_out/sweep1.err
CMD="python tools/gemv_sweep.py --qtype Q4_0 --reps 1 --rotate-mb 128"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_r1a.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 20 -c 2 -o gpurun_out/gemv_r1a $CMD > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
cat gpurun_out/sweep1.jsonl gpurun_out/sweep1_nograph.jsonl
