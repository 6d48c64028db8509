timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -m gpu -q 2>&1 | tail -3
for r in 100000 768 512 384 256 128 64; do B2G_SPARSE_RUNS=$r timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-secondary 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('sparse_runs', $r, 'ms', round(d['ms_per_step'],1), 'vis', round(d['config']['stage_ms_per_step']['vis'],1))"; done
