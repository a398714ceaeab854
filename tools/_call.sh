timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "decoder_layer" > gpurun_out/pytest_layer.log 2>&1; echo layer=$?; tail -8 gpurun_out/pytest_layer.log
timeout 300 python tools/layer_bench.py 2>&1 | tail -2 | tee gpurun_out/layer_bench.log
timeout 300 python tools/layer_bench.py --opt gemv_overlap=0 2>&1 | tail -1 | tee -a gpurun_out/layer_bench.log
timeout 300 python tools/layer_bench.py --qtype Q8_0 2>&1 | tail -1 | tee -a gpurun_out/layer_bench.log
