# sweep of the chunk-class thresholds on the C4 step (degridding split: tensor-core / warp-run / one thread per sample)
for mr in 8 32 64; do
B2G_TC_MAX_RUNS=$mr timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-secondary > gpurun_out/bench_mr$mr.json 2> gpurun_out/bench_mr$mr.err
python -c "
import json
d=json.load(open('gpurun_out/bench_mr$mr.json')); print('max_runs $mr', d['ms_per_step'], {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items() if k in ('vis','fft')})"
done
for sr in 256 512; do
B2G_SPARSE_RUNS=$sr timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-secondary > gpurun_out/bench_sr$sr.json 2> gpurun_out/bench_sr$sr.err
python -c "
import json
d=json.load(open('gpurun_out/bench_sr$sr.json')); print('sparse_runs $sr', d['ms_per_step'], {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items() if k in ('vis','fft')})"
done
