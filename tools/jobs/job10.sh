# round-2 profile evidence on the headline command
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-secondary"
timeout 900 $CMD > gpurun_out/r02_bench_prof_plain.json 2> gpurun_out/r02_bench_prof_plain.err; echo plain rc=$?
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file gpurun_out/r02_launches_c4.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo launches rc=$?; tail -2 gpurun_out/ncu_launch.log | cut -c1-200
timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'warprun|degrid_tc|hfft_' -c 30000 --csv --log-file gpurun_out/r02_dram_c4.csv $CMD > gpurun_out/ncu_dram.log 2>&1; echo dram rc=$?; tail -2 gpurun_out/ncu_dram.log | cut -c1-200
wc -l gpurun_out/r02_launches_c4.csv gpurun_out/r02_dram_c4.csv
