timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_fullsize.py tests/test_gpu_tc.py -m gpu -q 2>&1 | tail -3
timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-secondary > gpurun_out/bench_c4_c.json 2>gpurun_out/bench_c4_c.err; python -c "
import json
d=json.load(open('gpurun_out/bench_c4_c.json')); print(d['value'], d['ms_per_step'], {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items()}, d['roofline']['step']['frac'], d['roofline']['kernel'], d['roofline']['traffic'])"
