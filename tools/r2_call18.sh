timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_fuzz.py tests/test_gpu_blur.py tests/test_gpu_dropin.py -x -q 2>&1 | tail -4
ORBX_E2E_TRACE=3 timeout 300 python tools/exp_e2e.py 2>&1 | tail -30
