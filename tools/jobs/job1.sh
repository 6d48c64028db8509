timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/gputest.log 2>&1; tail -8 gpurun_out/gputest.log
python tools/wshard_bench.py --config C4 --nchan 8 --steps 2 > gpurun_out/fft_hand.json 2>gpurun_out/fft_hand.err
B2G_NO_HFFT=1 python tools/wshard_bench.py --config C4 --nchan 8 --steps 2 > gpurun_out/fft_cufft.json 2>gpurun_out/fft_cufft.err
python -c "
import json
for f in ('fft_hand','fft_cufft'):
    d=json.load(open('gpurun_out/'+f+'.json')); print(f, d['ms_per_step'], d['ranks'][0]['stage_ms'], d['ranks'][0]['nplanes'])
"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hfft_ -s 40 -c 8 -o gpurun_out/r02_hfft_v1 python tools/wshard_bench.py --config C4 --nchan 8 --steps 1 --warmup 0 > gpurun_out/ncu_hfft.log 2>&1; tail -3 gpurun_out/ncu_hfft.log
