timeout 600 python tools/sparse_perf.py --config C4 --nchan 256 --world 8 --rank 7
timeout 600 python tools/sparse_perf.py --config C4 --nchan 256 --world 8 --rank 5
