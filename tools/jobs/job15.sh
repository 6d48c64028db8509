for r in 16 32 64 128; do B2G_TC_MAX_RUNS=$r timeout 300 python tools/tc_perf.py --config C2 --modes 1 --steps 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('max_runs', $r, 'degrid ms', d['mode1_degrid_ms'])"; done
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:'degrid' -c 40 --csv --log-file gpurun_out/tc_launches.csv python tools/tc_perf.py --config C2 --modes 1 --steps 1 > /dev/null 2>&1
python tools/ncu_launches.py gpurun_out/tc_launches.csv --steps 3
grep -c degrid gpurun_out/tc_launches.csv
