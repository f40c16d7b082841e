N=${NGPU:-8}
timeout 600 python -m pytest tests/test_gpu_group.py -x -q 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err
tail -c 400 gpurun_out/r02_bench_${N}gpu.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_${N}gpu.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e'])
m=d['matching']; print('match',m['value'],m.get('kernel_only_gmatch_s_per_gpu'),m.get('parity_n_shards'),m.get('ms_per_step'))
for k,v in (d.get('configs') or {}).items(): print(k, v.get('value'), v.get('e2e',{}).get('value') if isinstance(v.get('e2e'),dict) else None)
PY
