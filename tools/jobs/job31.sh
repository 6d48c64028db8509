timeout 1500 python -m pytest tests/test_gpu_configs.py -m gpu -q -x --durations=8 > gpurun_out/gputest_cfg.log 2>&1; tail -25 gpurun_out/gputest_cfg.log | cut -c1-250
