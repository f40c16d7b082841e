N=${NGPU:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err
tail -c 300 gpurun_out/r02_bench_${N}gpu.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_${N}gpu.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',{k:v for k,v in d['e2e'].items() if not k.endswith('note')})
m=d['matching']; print('match',m['value'],m.get('kernel_only_gmatch_s_per_gpu'),m.get('parity_n_shards'),m.get('ms_per_step'))
print(d.get('roofline_issue'))
PY
