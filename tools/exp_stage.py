"""Experiment helper: per-stage and whole-batch device times of the extractor on synthetic frames.
    python tools/exp_stage.py [frames] [kind]     (env: ORBX_CHUNKS, ORBX_QT_SMEM_KEYS)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from orb_slam2_e_b200 import synth
from orb_slam2_e_b200.extractor import ORBextractor

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
kind = sys.argv[2] if len(sys.argv) > 2 else "textured"
cfg = (1000, 1.2, 8, 20, 7) if kind == "textured" else (1200, 1.1, 6, 24, 7)
h, w = 480, 640
host = synth.frame_batch(kind, n, h, w, 1000, distinct=min(n, 32))
dev = torch.from_numpy(host).cuda()
ext = ORBextractor(*cfg, device=0, max_width=w, max_height=h, max_batch=n)
for _ in range(3):
    ext.extract_batch_device(dev.data_ptr(), n, h, w, w, w * h)
ext.sync()
ext.set_stage_timing(True)
best = None
for _ in range(5):
    ext.extract_batch_device(dev.data_ptr(), n, h, w, w, w * h)
    ext.sync()
    st = ext.last_stage_ms()
    if best is None or sum(st.values()) < sum(best.values()):
        best = st
ext.set_stage_timing(False)
tot = []
for _ in range(8):
    ext.extract_batch_device(dev.data_ptr(), n, h, w, w, w * h)
    ext.sync()
    tot.append(ext.last_batch_ms()[0])
print("chunks=%s qt_keys=%s kind=%s n=%d stages(ms)=%s serial=%.3f batch(ms) min=%.3f med=%.3f -> %.0f frames/s" % (
    os.environ.get("ORBX_CHUNKS", "def"), os.environ.get("ORBX_QT_SMEM_KEYS", "def"), kind, n,
    {k: round(v, 3) for k, v in best.items()}, sum(best.values()), min(tot), sorted(tot)[len(tot) // 2], n / (min(tot) * 1e-3)))
try:
    import ctypes as C
    from orb_slam2_e_b200 import _lib
    lib = C.CDLL(os.path.join(os.path.dirname(_lib.__file__), "liborbfront.so"))
    hb = C.c_uint(0)
    lib.orbx_debug_qt_handed_back(C.byref(hb))
    ext.extract_batch_device(dev.data_ptr(), n, h, w, w, w * h); ext.sync()
    lib.orbx_debug_qt_handed_back(C.byref(hb))
    print("quadtree: %d of %d (frame, level) pairs handed to the sorting kernel" % (hb.value, n * cfg[2]))
except Exception as e:
    print("no handed-back counter:", e)
