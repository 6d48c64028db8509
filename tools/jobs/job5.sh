timeout 600 python -m pytest tests/test_gpu_tc.py -m gpu -q -x > gpurun_out/gputest_tc.log 2>&1; tail -30 gpurun_out/gputest_tc.log | cut -c1-300
