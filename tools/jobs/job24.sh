timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29545 bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/bench_c4_n4.json 2> gpurun_out/bench_c4_n4.err; echo rc=$?; tail -3 gpurun_out/bench_c4_n4.err | cut -c1-300; python -c "
import json
d=json.load(open('gpurun_out/bench_c4_n4.json')); print(d['value'], d['ms_per_step'], d['roofline']['step']['frac'], d['e2e']['value'], d['config'].get('rebalance'))
for r in d['config']['ranks']: print(r['rank'], r['samples'], r['nplanes'], round(r['ms_per_step'],1), {k: round(v,1) for k,v in r['stage_ms'].items() if v>1})
"
