"""Experiment: per-frame tracking searches (SearchByProjection last-frame + local points) through the
product header on the GPU vs the reference's own ORBmatcher.cc (-O3) on one host core."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import scene_lib as sl
import projection_cases as pc

ref = sl.ref_side(fast=True)
got = sl.dropin_side()
for n_points, stereo in ((1500, False), (4000, False), (4000, True), (12000, False)):
    W = sl.World(7, n_points=n_points, stereo=stereo)
    rng = np.random.default_rng(1)
    Tl, Tc = sl.pose(rng), sl.pose(rng)
    last = W.observe(Tl, 11, assign=0.8, outliers=0.05)
    cur = W.observe(Tc, 12, assign=0.0)
    P = W.points(track=pc.frustum_fields(W, Tc, rng))
    order = rng.permutation(P.n)
    r = sl.time_tracking(ref, cur, last, P, order, mono=not stereo) if ref else (float("nan"),) * 2
    g = sl.time_tracking(got, cur, last, P, order, mono=not stereo)
    print(f"map {n_points:6d} pts, frame {cur.n:5d} kps, stereo={stereo}: last-frame GPU {g[0]:7.0f} us / ref CPU {r[0]:7.0f} us ; "
          f"local-points GPU {g[1]:7.0f} us / ref CPU {r[1]:7.0f} us")
