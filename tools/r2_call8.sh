set -x
timeout 600 python -m pytest tests/test_gpu_match.py tests/test_gpu_group.py tests/test_gpu_window.py tests/test_gpu_projection.py tests/test_gpu_stereo_batch.py tests/test_gpu_vocabulary.py -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps 5 --warmup 3 --no-configs --frames 128 --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); m=d['matching']; print(m['value'], m['kernel_only_gmatch_s_per_gpu'], m['kernel_ms'], m['roofline']['frac'])"
