"""Runs every small matcher kernel once at tracking-thread sizes (for ncu launch lists / captures) and
prints host-side call latency."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import match_cases as mc
from orb_slam2_e_b200 import matcher, synth, extractor as ex
from orb_slam2_e_b200.extractor import ORBextractor

def t(fn, reps=20):
    fn(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return sorted(ts)[len(ts) // 2] * 1e6

reps = int(os.environ.get("REPS", "20"))
frame, q, qd, claimed = mc.window_case(3, n=2000, nq=2000, th_px=15.0)
out = matcher.window_search(frame, q, qd, claimed, matcher.TH_HIGH, 0.0, True)
print(f"window_search claim=1  n=2000 nq=2000: {t(lambda: matcher.window_search(frame, q, qd, claimed, matcher.TH_HIGH, 0.0, True), reps):.0f} us, passes {out[2]}, accepted {out[1]}")
print(f"window_search claim=0  n=2000 nq=2000: {t(lambda: matcher.window_search(frame, q, qd, claimed, matcher.TH_HIGH, 0.8, False), reps):.0f} us")
case = mc.bow_case(3, n1=2000, n2=2000, n_nodes=100)
m = matcher.ORBmatcher(0.75, True)
print(f"SearchByBoW 2000x2000/100 nodes: {t(lambda: m.SearchByBoW(case['d1'], case['fv1'], case['v1'], case['d2'], case['fv2'], None, case['a1'], case['a2'], kf_kf=False), reps):.0f} us")
left, right = (np.ascontiguousarray(a) for a in synth.stereo_pair(376, 1241, 34))
g = ORBextractor(2000, 1.2, 8, 20, 7)
(kl, dl), (kr, dr) = g.extract_batch(np.stack([left, right]))
mbf, mb = 386.1448, 386.1448 / 718.856
ur, z = ex.stereo_matches(g, g, kl, dl, kr, dr, mb, mbf, 0, 1)
print(f"stereo_matches KITTI pair nl={len(kl)} nr={len(kr)}: {t(lambda: ex.stereo_matches(g, g, kl, dl, kr, dr, mb, mbf, 0, 1), reps):.0f} us, matched {(ur >= 0).sum()}")
