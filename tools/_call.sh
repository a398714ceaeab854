set -x
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:kq_gemv -s 3 -c 1 -o gpurun_out/kq4 -f python tools/kq_bench.py --types Q4_K --shapes 4096x4096 --reps 1 > gpurun_out/kq4.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:kq_gemv -s 3 -c 1 -o gpurun_out/kq8 -f python tools/kq_bench.py --types Q8_K --shapes 4096x4096 --reps 1 > gpurun_out/kq8.log 2>&1
ls -la gpurun_out | tail -5
