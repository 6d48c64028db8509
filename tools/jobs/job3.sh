timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/gputest3.log 2>&1; tail -5 gpurun_out/gputest3.log | cut -c1-250
timeout 300 python tools/wshard_bench.py --config C4 --nchan 8 --steps 2 > gpurun_out/fft_hand.json 2>gpurun_out/fft_hand.err
python -c "
import json
for f in ('fft_hand',):
    d=json.load(open('gpurun_out/'+f+'.json')); print(f, d['ms_per_step'], d['ranks'][0]['stage_ms'], d['ranks'][0]['nplanes'])
"
timeout 900 python bench.py > gpurun_out/bench_c4_b.json 2> gpurun_out/bench_c4_b.err; echo benchrc=$?; tail -3 gpurun_out/bench_c4_b.err; python -c "
import json
d=json.load(open('gpurun_out/bench_c4_b.json')); print(d['value'], d['ms_per_step'], d['config']['stage_ms_per_step'], d['roofline']['step'], d['e2e'], d['config']['secondary'])
"
