mkdir -p gpurun_out
rm -f gpurun_out/z_probe5.txt
for m in headline pinned; do timeout 30 python tools/_probe.py $m 2>&1 | tail -3 >> gpurun_out/z_probe5.txt; done
cat gpurun_out/z_probe5.txt
