timeout 300 python tools/exp_latency.py 2>&1 | tail -4
