set -x
( time python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_a.json 2> gpurun_out/r2_bench_a.err ); tail -5 gpurun_out/r2_bench_a.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_a.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'], 'stages', d['roofline']['stage_ms'])
for k,v in (d.get('configs') or {}).items():
    print(k, {kk:vv for kk,vv in v.items() if kk in ('value','unit','ms_per_step','stage_ms','parity','us_per_call_with_pyramid_download','us_per_call_without_pyramid_download','extraction_only_pairs_s','error','unavailable')}, 'e2e', (v.get('e2e') or {}).get('value'), 'frac', (v.get('roofline_pipeline') or {}).get('frac'), 'cpu', (v.get('cpu_reference') or {}).get('value'))
print('cv2', d.get('cpu_baseline_cv2'))
PY
