# full ncu capture of the gridding kernel of the current build on C2 (source-level stalls)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:^grid_warprun -s 0 -c 1 -o gpurun_out/r02_d_grid_c2 python tools/gpu_perf.py --config C2 --reps 1 > gpurun_out/ncu_grid.log 2>&1; tail -2 gpurun_out/ncu_grid.log
ls -la gpurun_out/*.ncu-rep
