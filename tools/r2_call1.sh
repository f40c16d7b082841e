set -x
timeout 600 python -m pytest tests/test_gpu_extract.py tests/test_gpu_fuzz.py -x -q > gpurun_out/r2c1_tests.log 2>&1; tail -15 gpurun_out/r2c1_tests.log
for k in textured endoscopy; do
  timeout 120 python tools/exp_stage.py 1024 $k
  ORBX_FAST_CELLS=1 timeout 120 python tools/exp_stage.py 1024 $k
done
