set -x
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_fuzz.py tests/test_gpu_blur.py -x -q 2>&1 | tail -4
timeout 300 python tools/exp_stage.py 1024 textured 2>&1 | tail -2 | cut -c1-230
timeout 300 python tools/exp_stage.py 1024 endoscopy 2>&1 | tail -2 | cut -c1-230
