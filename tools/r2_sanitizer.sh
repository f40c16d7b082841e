# one compute-sanitizer tool per gpurun call (B200_PROFILING.md); $1 = memcheck | racecheck
set -x
TOOL=$1
timeout 300 python -m pytest tests/test_gpu_extract.py tests/test_gpu_fuzz.py tests/test_gpu_window.py tests/test_gpu_match.py tests/test_gpu_blur.py tests/test_gpu_stereo_batch.py -x -q > gpurun_out/r2_san_plain_$TOOL.log 2>&1 || { tail -5 gpurun_out/r2_san_plain_$TOOL.log; exit 1; }
timeout 2400 compute-sanitizer --tool $TOOL --target-processes all --print-limit 20 --log-file gpurun_out/r2_sanitizer_$TOOL.txt python -m pytest tests/test_gpu_extract.py tests/test_gpu_fuzz.py tests/test_gpu_window.py tests/test_gpu_match.py tests/test_gpu_blur.py tests/test_gpu_stereo_batch.py -x -q > gpurun_out/r2_san_run_$TOOL.log 2>&1
tail -3 gpurun_out/r2_san_run_$TOOL.log; tail -5 gpurun_out/r2_sanitizer_$TOOL.txt
