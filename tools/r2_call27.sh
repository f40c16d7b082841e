for r in 0 6 8 12 16; do echo "ratio4=$r"; ORBX_QP_RATIO4=$r timeout 300 python tools/exp_stage.py 1024 endoscopy 2>&1 | tail -2 | sed 's/.*stages/stages/' | cut -c1-150; done
ORBX_QT_NO_PYRAMID=1 timeout 300 python tools/exp_stage.py 1024 endoscopy 2>&1 | tail -2 | sed 's/.*stages/stages/' | cut -c1-150
