timeout 300 python tools/exp_e2e.py 2>&1 | grep -E "pipelined chunk"
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_fuzz.py -x -q 2>&1 | tail -3
