# full ncu capture of the four plane-transform kernels of the current build (all stall reasons)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hfft_ -s 20 -c 2 -o gpurun_out/r02_h_hfft_inv python tools/wshard_bench.py --config C4 --nchan 8 --steps 1 --warmup 0 > gpurun_out/ncu_hfft.log 2>&1; tail -3 gpurun_out/ncu_hfft.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hfft_.*fwd -s 20 -c 2 -o gpurun_out/r02_h_hfft_fwd python tools/wshard_bench.py --config C4 --nchan 8 --steps 1 --warmup 0 > gpurun_out/ncu_hfft2.log 2>&1; tail -3 gpurun_out/ncu_hfft2.log
ls -la gpurun_out/*.ncu-rep
