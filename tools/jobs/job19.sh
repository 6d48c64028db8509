timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/gputest_final.log 2>&1; tail -4 gpurun_out/gputest_final.log | cut -c1-200
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo refrc=$?; cut -c1-400 gpurun_out/bench_ref.json
timeout 1200 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo rc=$?; python -c "
import json
d=json.load(open('gpurun_out/bench_final.json')); print(d['value'], d['ms_per_step'], {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items()}, d['roofline']['step']['frac'], d['roofline']['kernel'], d['roofline']['frac'], d['e2e']['value'], d['e2e']['ms_per_step'], d['cpu_baseline']['value'], d['gpu_launches'], d['config']['secondary']['ms_per_step'])"
