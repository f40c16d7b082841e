timeout 300 python tools/exp_e2e.py 2>&1 | grep -E "pipelined chunk 0"
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
