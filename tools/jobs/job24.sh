# parity of the configurations / multi-slab ring / full sizes, then the C4 step without the e2e and CPU legs
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_fullsize.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/gputest_cfg.log 2>&1; tail -5 gpurun_out/gputest_cfg.log | cut -c1-250
timeout 900 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-secondary > gpurun_out/bench_step.json 2> gpurun_out/bench_step.err; echo rc=$?
python -c "
import json
d=json.load(open('gpurun_out/bench_step.json')); print(d['ms_per_step'], {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items()}, d['roofline']['step']['frac'])"
