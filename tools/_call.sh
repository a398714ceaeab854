#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r02f_gpu_tests.txt
cat gpurun_out/r02f_gpu_tests.txt
