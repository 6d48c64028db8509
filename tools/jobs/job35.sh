# end-of-round evidence: full GPU suite, smoke, both bench arms, the launch list with DRAM bytes of the bench command
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/gputest_final.log 2>&1; tail -4 gpurun_out/gputest_final.log | cut -c1-200
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo refrc=$?
timeout 1200 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo rc=$?; python -c "
import json
d=json.load(open('gpurun_out/bench_final.json')); print(d['value'], d['ms_per_step'], {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items()}, d['roofline']['step']['frac'], d['roofline']['kernel'], d['roofline']['frac'], d['e2e']['value'], d['e2e']['ms_per_step'], d['cpu_baseline']['value'], d['gpu_launches'], d['config']['secondary']['ms_per_step'])"
timeout 600 python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-secondary > gpurun_out/bench_noe2e.json 2>/dev/null
timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 30000 --csv --log-file gpurun_out/launches_dram.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-secondary > gpurun_out/ncu_launches.log 2>&1; tail -1 gpurun_out/ncu_launches.log | cut -c1-200
gzip -f gpurun_out/launches_dram.csv; ls -la gpurun_out/launches_dram.csv.gz
timeout 600 python tools/e2e_cprofile.py --config C4 --nchan 256 > gpurun_out/e2e_cprofile.txt 2>&1; head -2 gpurun_out/e2e_cprofile.txt
