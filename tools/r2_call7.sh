set -x
timeout 900 python -m pytest tests/test_gpu_match.py tests/test_gpu_dropin.py tests/test_gpu_projection.py -x -q 2>&1 | tail -12

