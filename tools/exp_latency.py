"""Experiment: single-frame latency through orbx_extract (host image in, host results out)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from orb_slam2_e_b200 import synth
from orb_slam2_e_b200.extractor import ORBextractor
for (h, w, cfg) in ((480, 640, (1000, 1.2, 8, 20, 7)), (376, 1241, (2000, 1.2, 8, 20, 7)), (2160, 3840, (8000, 1.2, 12, 20, 7))):
    img = synth.textured_frame(h, w, 3)
    ext = ORBextractor(*cfg)
    for _ in range(5): ext(img)
    ts = []
    for _ in range(50):
        t0 = time.perf_counter(); k, d = ext(img); ts.append(time.perf_counter() - t0)
    ts.sort()
    ext.set_stage_timing(True); ext(img); st = ext.last_stage_ms(); ext.set_stage_timing(False)
    print("%dx%d: %d kps, latency median %.0f us min %.0f us; device ms %.3f; stages %s" % (w, h, len(k), ts[25] * 1e6, ts[0] * 1e6, ext.last_batch_ms()[0], {a: round(b, 3) for a, b in st.items()}))
