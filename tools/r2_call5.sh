set -x
timeout 600 python -m pytest tests/test_gpu_stereo_batch.py tests/test_gpu_frame_finish.py tests/test_gpu_extract.py -x -q > gpurun_out/r2c6_tests.log 2>&1; tail -6 gpurun_out/r2c6_tests.log
bash tools/r2_call4.sh
