timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_d.json 2> gpurun_out/r02_bench_d.err; tail -c 200 gpurun_out/r02_bench_d.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_d_ref.json 2>/dev/null; tail -c 600 gpurun_out/r02_bench_d_ref.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_d.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',{k:v for k,v in d['e2e'].items() if not k.endswith('note')})
print('issue',d['roofline_issue']['frac'],d['roofline_issue']['step_frac'],'stage_ms',d['roofline']['stage_ms'])
m=d['matching']; print('match',m['value'],m.get('kernel_only_gmatch_s_per_gpu'))
for k,v in (d.get('configs') or {}).items(): print(k, {a:b for a,b in v.items() if a in ('value','unit','us_per_call_with_pyramid_download','us_per_call_without_pyramid_download','ms_per_step')}, v.get('parity',{}).get('ok'), (v.get('e2e') or {}).get('value') if isinstance(v.get('e2e'),dict) else None)
PY
