python bench.py > gpurun_out/bench_r01e.json 2> gpurun_out/bench_r01e.err; echo bench=$?; wc -l gpurun_out/bench_r01e.json; python -c "
import json; d=json.load(open('gpurun_out/bench_r01e.json')); print(d['value'], d['roofline']['frac'], d['e2e']['value'], d['cpu_baseline']); print(json.dumps(d['config']['other_paths'], indent=1)[:1800])"; tail -3 gpurun_out/bench_r01e.err
