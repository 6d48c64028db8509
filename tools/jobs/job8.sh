for r in 8 16 24 32 48 64; do B2G_TC_MAX_RUNS=$r timeout 300 python tools/tc_perf.py --config C2 --modes 1 --steps 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('max_runs', $r, 'degrid ms', d['mode1_degrid_ms'])"; done
