"""Experiment: latency of the small matcher entry points (host buffers in/out) vs the CPU oracle."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import match_cases as mc
from oracle_lib import oracle
from orb_slam2_e_b200 import matcher

def t(fn, reps=30):
    fn(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return sorted(ts)[len(ts) // 2] * 1e6

import ctypes as C
from orb_slam2_e_b200 import _lib
from orb_slam2_e_b200._lib import ptr, check
from orb_slam2_e_b200.matcher import _FlatFeatVec

L = oracle(fast=True)
lib = _lib.load()
for (n1, n2, nodes) in ((1000, 1000, 40), (2000, 2000, 80), (1000, 1200, 200), (1000, 1000, 1000)):
    case = mc.bow_case(3, n1=n1, n2=n2, n_nodes=nodes)
    # bare C calls on both sides: the Python flattening of the FeatureVector is not what is compared
    f1, f2 = _FlatFeatVec(case["fv1"]), _FlatFeatVec(case["fv2"])
    out = np.zeros(n2, np.int32)
    nm = C.c_int(0)
    def gpu():
        check(lib.orbm_search_bow(0, ptr(case["d1"]), n1, ptr(case["d2"]), n2, C.byref(f1.c), C.byref(f2.c), ptr(case["v1"]),
                                  None, ptr(case["a1"]), ptr(case["a2"]), 0.75, 1, 0, ptr(out), C.byref(nm)))
    nn1, off1, i1 = mc.flat_fv(case["fv1"])
    nn2, off2, i2 = mc.flat_fv(case["fv2"])
    out2 = np.zeros(n2, np.int32)
    def cpu():
        return L.orc_search_bow(case["d1"], n1, case["d2"], n2, nn1, off1, i1, len(nn1), nn2, off2, i2, len(nn2),
                                case["v1"], None, case["a1"], case["a2"], 0.75, 1, 0, out2)
    g = t(gpu); c = t(cpu)
    assert nm.value == cpu() and (out == out2).all()
    print(f"SearchByBoW n1={n1} n2={n2} nodes={nodes}: GPU {g:.0f} us  CPU oracle(-O3) {c:.0f} us  matches {nm.value}")
for (nl, nr) in ((2000, 2000),):
    kl, dl, kr, dr, scale, n_rows = mc.stereo_case(4, nl=nl, nr=nr)
    m = matcher.ORBmatcher()
    g = t(lambda: m.StereoSearch(kl, dl, kr, dr, scale, n_rows, 0.0, 80.0))
    c = t(lambda: mc.oracle_stereo(L, kl, dl, kr, dr, scale, n_rows, 0.0, 80.0))
    print(f"StereoSearch nl={nl} nr={nr}: GPU {g:.0f} us  CPU oracle {c:.0f} us")
