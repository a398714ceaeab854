mkdir -p gpurun_out
timeout 40 python tools/_probe.py > gpurun_out/z_probe6.txt 2>&1
tail -8 gpurun_out/z_probe6.txt
