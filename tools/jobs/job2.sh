# v3 plane FFT: parity, shard test log, FFT timing (hand vs cuFFT), ncu of the four kernels
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_shard.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/gputest2.log 2>&1; tail -40 gpurun_out/gputest2.log | cut -c1-250
timeout 300 python tools/wshard_bench.py --config C4 --nchan 8 --steps 2 > gpurun_out/fft_hand.json 2>gpurun_out/fft_hand.err
python -c "
import json
for f in ('fft_hand',):
    d=json.load(open('gpurun_out/'+f+'.json')); print(f, d['ms_per_step'], d['ranks'][0]['stage_ms'], d['ranks'][0]['nplanes'])
"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hfft_ -s 20 -c 4 -o gpurun_out/r02_hfft_v3_inv python tools/wshard_bench.py --config C4 --nchan 8 --steps 1 --warmup 0 > gpurun_out/ncu_hfft.log 2>&1; tail -3 gpurun_out/ncu_hfft.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hfft_.*fwd -s 20 -c 4 -o gpurun_out/r02_hfft_v3_fwd python tools/wshard_bench.py --config C4 --nchan 8 --steps 1 --warmup 0 > gpurun_out/ncu_hfft2.log 2>&1; tail -3 gpurun_out/ncu_hfft2.log
