for v in 0 1 2 3; do ORBX_RZ=$v timeout 300 python tools/exp_stage.py 1024 textured 2>&1 | tail -1 | cut -c1-200; done
ORBX_NO_TMA_RESIZE=1 timeout 300 python tools/exp_stage.py 1024 textured 2>&1 | tail -1 | cut -c1-200
