timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_c4_n2.json 2> gpurun_out/bench_c4_n2.err; echo rc=$?; tail -5 gpurun_out/bench_c4_n2.err | cut -c1-300; python -c "
import json
d=json.load(open('gpurun_out/bench_c4_n2.json')); print(d['value'], d['ms_per_step'], d['scaling'], d['config'].get('parallelism'), d['config'].get('collective'), d['roofline']['step'], d['e2e']['value'], d['config'].get('ranks'))
"
