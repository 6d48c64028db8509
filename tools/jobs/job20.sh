# the allocation-free entry point: alone in a fresh process, then after the whole GPU suite (job19 = final verification)
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -4
compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -6
bash tools/jobs/job19.sh
tail -8 gpurun_out/gputest_final.log | cut -c1-200
