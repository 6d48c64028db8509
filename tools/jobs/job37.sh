timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -q -x 2>&1 | tail -45 | cut -c1-300
