timeout 900 python -m pytest tests/test_gpu_wrapper.py tests/test_gpu_pipeline.py tests/test_gpu_skymodel.py tests/test_gpu_base.py -m gpu -q -x 2>&1 | tail -3
timeout 900 python tools/e2e_cprofile.py --config C4 --nchan 256 > gpurun_out/e2e_cprofile.txt 2>&1; head -42 gpurun_out/e2e_cprofile.txt | cut -c1-200
