# refreshes gpurun_out/r01_d_{clean,msclean,msmfs}_*.json (run on a B200: bash tools/clean_perf_all.sh)
python tools/clean_perf.py --npix 2048 --niter 2000 --cpu-niter 20 --json gpurun_out/r01_d_clean_2048.json | tail -1
python tools/clean_perf.py --npix 8192 --niter 500 --cpu-niter 2 --json gpurun_out/r01_d_clean_8192.json | tail -1
python tools/clean_perf.py --msclean --npix 2048 --niter 1000 --cpu-niter 10 --json gpurun_out/r01_d_msclean_2048.json | tail -1
python tools/clean_perf.py --msclean --npix 512 --niter 1000 --cpu-niter 50 --json gpurun_out/r01_d_msclean_512.json | tail -1
python tools/clean_perf.py --msmfs --npix 2048 --niter 500 --cpu-niter 0 --json gpurun_out/r01_d_msmfs_2048.json | tail -1
python tools/clean_perf.py --msmfs --npix 1024 --niter 500 --cpu-niter 5 --json gpurun_out/r01_d_msmfs_1024.json | tail -1
