timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_c.json 2> gpurun_out/r02_bench_c.err; tail -c 300 gpurun_out/r02_bench_c.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_c.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',{k:v for k,v in d['e2e'].items() if not k.endswith('note')})
print('roofline',{k:v for k,v in d['roofline'].items() if k in ('bound','kernel','achieved','peak','frac','traffic')})
print('issue',d['roofline_issue'])
print('pipe',d['roofline_pipeline'])
print('cpu',d['cpu_baseline'],d.get('cpu_baseline_cv2'))
m=d['matching']; print('match',m['value'],m.get('kernel_only_gmatch_s_per_gpu'),m['roofline'])
for k,v in (d.get('configs') or {}).items(): print(k, json.dumps(v)[:600])
PY
