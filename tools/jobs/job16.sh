timeout 600 python tools/sparse_perf.py --config C4 --nchan 256 --world 2 --rank 1 2>&1 | tail -2
timeout 600 python tools/sparse_perf.py --config C4 --nchan 256 --world 4 --rank 2 2>&1 | tail -2
