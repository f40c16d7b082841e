QP_NAMES=1 timeout 300 python tools/exp_qt_profile2.py 2>&1 | tail -13
