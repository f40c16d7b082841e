set -x
timeout 600 python -m pytest tests/test_gpu_group.py tests/test_gpu_match.py -x -q 2>&1 | tail -4
