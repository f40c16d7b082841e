set -x
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -5
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_b.json 2> gpurun_out/r02_bench_b.err; tail -c 600 gpurun_out/r02_bench_b.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_b.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d.get('stage_ms'), d['matching']['value'], d['matching'].get('kernel_only_gmatch_s_per_gpu'))
for k,v in d.get('configs',{}).items(): print(k, {a:b for a,b in v.items() if a in ('value','unit','us_per_call','us_per_call_no_pyramid','parity')})
PY
