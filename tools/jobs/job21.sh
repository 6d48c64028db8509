timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_fullsize.py -m gpu -q 2>&1 | tail -3
timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-secondary 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('ms', round(d['ms_per_step'],1), {k: round(v,1) for k,v in d['config']['stage_ms_per_step'].items()})"
