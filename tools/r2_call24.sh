for k in 1 2 3 4 6; do ORBX_RESIDENT_SPLIT=$k timeout 300 python tools/exp_stage.py 1024 textured 2>&1 | tail -1 | sed 's/.*serial/serial/'; done
timeout 600 python -m pytest tests/test_gpu_dropin.py tests/test_gpu_extract.py -x -q 2>&1 | tail -2
