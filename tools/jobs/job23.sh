timeout 1200 python bench.py > gpurun_out/bench_final2.json 2> gpurun_out/bench_final2.err; echo rc=$?; python -c "
import json
d=json.load(open('gpurun_out/bench_final2.json')); print(d['value'], d['ms_per_step'], {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items()}, d['roofline']['step']['frac'], d['roofline']['kernel'], d['roofline']['frac'], d['e2e']['value'], d['e2e']['ms_per_step'], d['cpu_baseline']['value'], d['gpu_launches'], d['config']['secondary']['ms_per_step'])"
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-secondary"
timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 30000 --csv --log-file gpurun_out/r02_c_launches_c4.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo launches rc=$?
python tools/ncu_launches.py gpurun_out/r02_c_launches_c4.csv --steps 4
