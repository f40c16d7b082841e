timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_fuzz.py tests/test_gpu_blur.py tests/test_gpu_dropin.py -x -q 2>&1 | tail -5
timeout 300 python tools/exp_stage.py 1024 textured 2>&1 | tail -2
timeout 300 python tools/exp_stage.py 1024 endoscopy 2>&1 | tail -2
