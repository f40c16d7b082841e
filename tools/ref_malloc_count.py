"""how often does the reference's own ORBextractor.cc give a different kept set when its ExtractorNode pointer
order (the tie-break of ORBextractor.cc:684) comes from glibc malloc instead of the monotonic arena the parity oracle
uses?  Runs both builds (oracle/_ref/liborbref.so = arena, liborbref_fast.so = glibc malloc, -O3) on the frames the
parity tests and the bench use and counts frames / (frame, level) sets that differ."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib  # noqa: E402
from orb_slam2_e_b200 import synth  # noqa: E402

CASES = [("TUM textured 640x480", (1000, 1.2, 8, 20, 7), [synth.textured_frame(480, 640, 1000 * 1000 + i) for i in range(24)]),
         ("Hamlyn endoscopy 640x480", (1200, 1.1, 6, 24, 7), [synth.endoscopy_frame(480, 640, 3000 * 1000 + i) for i in range(24)]),
         ("KITTI 1241x376", (2000, 1.2, 8, 20, 7), [synth.stereo_pair(376, 1241, 2000 + i)[0] for i in range(12)])]
tot_f = tot_fl = bad_f = bad_fl = 0
for name, cfg, frames in CASES:
    a, b = oracle_lib.RefExtractor(*cfg), oracle_lib.RefExtractor(*cfg, fast=True)
    nf = nl = 0
    for img in frames:
        ka, _ = a.extract(img)
        kb, _ = b.extract(img)
        diff_levels = 0
        for l in range(cfg[2]):
            sa = set(map(tuple, np.stack([ka["x"], ka["y"]], 1)[ka["octave"] == l].tolist()))
            sb = set(map(tuple, np.stack([kb["x"], kb["y"]], 1)[kb["octave"] == l].tolist()))
            diff_levels += sa != sb
        nf += diff_levels > 0
        nl += diff_levels
    print("%-26s frames differing %d of %d, (frame, level) sets differing %d of %d" % (name, nf, len(frames), nl, len(frames) * cfg[2]))
    tot_f += len(frames); bad_f += nf; tot_fl += len(frames) * cfg[2]; bad_fl += nl
print("total: %d of %d frames, %d of %d (frame, level) kept sets differ between the arena build and the glibc-malloc build" % (bad_f, tot_f, bad_fl, tot_fl))
