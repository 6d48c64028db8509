timeout 900 python tools/e2e_cprofile.py --config C4 --nchan 256 > gpurun_out/e2e_cprofile.txt 2>&1; tail -60 gpurun_out/e2e_cprofile.txt | cut -c1-180
