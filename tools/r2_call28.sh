timeout 1200 python -m pytest tests/test_gpu_extract.py tests/test_gpu_fuzz.py tests/test_gpu_blur.py tests/test_gpu_dropin.py tests/test_gpu_stereo_batch.py -x -q 2>&1 | tail -5
timeout 300 python tools/exp_stage.py 1024 textured 2>&1 | tail -2 | sed 's/.*stages/stages/' | cut -c1-170
timeout 300 python tools/exp_stage.py 1024 endoscopy 2>&1 | tail -2 | sed 's/.*stages/stages/' | cut -c1-170
timeout 300 python tools/exp_latency.py 2>&1 | tail -3
