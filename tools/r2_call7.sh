set -x
timeout 900 python -m pytest tests/test_gpu_blur.py tests/test_gpu_extract.py tests/test_gpu_dropin.py -x -q 2>&1 | tail -12
timeout 300 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu --no-match 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e'])"
