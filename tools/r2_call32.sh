timeout 600 python -m pytest tests/test_gpu_stereo_batch.py tests/test_gpu_extract.py -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-match 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'])
for k,v in d['configs'].items(): print(k,{a:b for a,b in v.items() if a in ('value','unit','us_per_call_with_pyramid_download','us_per_call_without_pyramid_download','parity','ms_per_step')})
"
