# four GPUs: the strong-scaling bench line of the final build
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 4 --steps 3 --warmup 3 --no-cpu --no-secondary > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err; echo rc=$?
python -c "
import json
d=json.load(open('gpurun_out/bench_n4.json')); print(d['n_gpus'], d['ms_per_step'], d['value'], d['scaling'], d['e2e']['value'] if d.get('e2e') else None)"
