for fc in 0 1; do
B2G_FUSED_CLEAR=$fc timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-secondary > gpurun_out/bench_f$fc.json 2> gpurun_out/bench_f$fc.err
python -c "
import json
d=json.load(open('gpurun_out/bench_f$fc.json')); print('fused_clear $fc', d['ms_per_step'], {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items() if k in ('prep','vis','fft')})"
done
