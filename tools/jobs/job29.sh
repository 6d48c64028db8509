# A/B of the fused ring clear on the full default bench (same box)
for fc in 1 0 1 0; do
B2G_FUSED_CLEAR=$fc timeout 900 python bench.py --no-cpu --no-secondary > gpurun_out/bench_fc$fc.json 2> gpurun_out/bench_fc$fc.err
python -c "
import json
d=json.load(open('gpurun_out/bench_fc$fc.json')); print('fused_clear $fc', d['ms_per_step'], d['steps'], {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items() if k in ('prep','vis','fft')}, d['e2e']['ms_per_step'])"
done
