"""Experiment: cycle stamps inside quadtree_select_kernel (level-0 CTA of frame 0).
Needs a developer build: ORBFRONT_NVCC_FLAGS=-DORBX_QT_PROFILE python -m orb_slam2_e_b200.build --force"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from orb_slam2_e_b200 import synth, _lib
from orb_slam2_e_b200.extractor import ORBextractor
lib = _lib.load()
names = ["attach", "radix", "to-smem", "bounds", "replay", "ordered", "leaf heads", "leaf walk", "leaf sort", "emit"]
for (h, w, cfg) in ((480, 640, (1000, 1.2, 8, 20, 7)), (376, 1241, (2000, 1.2, 8, 20, 7)), (2160, 3840, (8000, 1.2, 12, 20, 7))):
    img = synth.textured_frame(h, w, 3)
    ext = ORBextractor(*cfg)
    for _ in range(3): ext(img)
    out = (C.c_longlong * 32)()
    lib.orbx_debug_qt_prof(out)
    ext(img)
    lib.orbx_debug_qt_prof(out)
    t = list(out)
    print("%dx%d: n=%d L=%d d_bfs=%d d=%d dmax=%d nbits=%d ordered passes=%d first M=%d" % (w, h, t[10], t[11], t[12], t[13], t[14], t[15], t[19], t[20]))
    print("   cycles: " + ", ".join("%s %d" % (names[i], t[i + 1] - t[i]) for i in range(9)) + "; total %d" % (t[9] - t[0]))
    if t[19]: print("   first ordered pass: build %d sort %d rest %d" % (t[16] - t[4], t[17] - t[16], t[18] - t[17]))
