timeout 300 python tools/exp_c2.py 2>&1 | tail -3
