for k in 1 0 2; do ORBX_FORK_AT=$k timeout 300 python tools/exp_stage.py 1024 textured 2>&1 | tail -2 | head -1 | sed 's/.*serial/serial/'; done
