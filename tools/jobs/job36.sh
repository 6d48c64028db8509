# two GPUs: the sharded drop-in tests, then the strong-scaling bench line at N = 2
timeout 900 python -m pytest tests/test_gpu_shard.py -m gpu -q 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu --no-secondary > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo rc=$?
python -c "
import json
d=json.load(open('gpurun_out/bench_n2.json')); print(d['n_gpus'], d['ms_per_step'], d['value'], d['scaling'], d['e2e']['value'] if d.get('e2e') else None)"
