timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "gemv or matmul or chain or nan or inf or random" > gpurun_out/pytest_gpu.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_gpu.log
python bench.py --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['roofline']['frac'],4), 'dep', round(d['config']['dependent_chain_GBps'],1), 'e2e', round(d['e2e']['value'],1), d['check'], d['config']['activation_bits'])"
