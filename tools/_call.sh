timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_r01b.json 2> gpurun_out/bench_r01b.err; echo bench=$?; cut -c1-600 gpurun_out/bench_r01b.json
CMD="python bench.py --steps 2 --warmup 3 --layers 4 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01b.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --warp-sampling-interval 0 --clock-control none --import-source on -k regex:gemv_kernel -s 30 -c 3 -o gpurun_out/gemv_r01b $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
