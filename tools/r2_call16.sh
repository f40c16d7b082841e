for k in 1024 2048 4096; do ORBX_QT_SMEM_KEYS=$k timeout 300 python tools/exp_stage.py 1024 textured 2>&1 | tail -1 | cut -c1-190; done
