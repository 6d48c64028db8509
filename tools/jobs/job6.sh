timeout 600 python -m pytest tests/test_gpu_tc.py -m gpu -q > gpurun_out/gputest_tc.log 2>&1; tail -5 gpurun_out/gputest_tc.log | cut -c1-300
timeout 600 python tools/tc_perf.py --config C2 > gpurun_out/tc_perf_c2.json 2> gpurun_out/tc_perf_c2.err; tail -3 gpurun_out/tc_perf_c2.err; cat gpurun_out/tc_perf_c2.json
