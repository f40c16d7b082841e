set -x
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2c5_tests.log 2>&1; tail -6 gpurun_out/r2c5_tests.log
bash tools/r2_call4.sh
