set -x
timeout 900 python -m pytest tests/test_gpu_extract.py tests/test_gpu_fuzz.py tests/test_gpu_dropin.py -x -q > gpurun_out/r2c3_tests.log 2>&1; tail -15 gpurun_out/r2c3_tests.log
for k in textured endoscopy; do
  timeout 120 python tools/exp_stage.py 1024 $k
done
timeout 120 python tools/exp_stage.py 1 textured
