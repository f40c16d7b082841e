"""Latency of the batched SearchByBoW (K candidate keyframes against one frame, bare C calls) vs K sequential calls."""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import match_cases as mc
from orb_slam2_e_b200 import _lib
from orb_slam2_e_b200._lib import check, ptr
from orb_slam2_e_b200.matcher import _FlatFeatVec


def t(fn, reps=30):
    fn(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return sorted(ts)[len(ts) // 2] * 1e6


lib = _lib.load()
for K, n, nodes in ((30, 1000, 40), (10, 1000, 40), (30, 2000, 80)):
    base = mc.bow_case(3, n1=n, n2=n, n_nodes=nodes)
    f2 = _FlatFeatVec(base["fv2"])
    cases = [mc.bow_case(50 + k, n1=n, n2=n, n_nodes=nodes) for k in range(K)]
    f1 = [_FlatFeatVec(c["fv1"]) for c in cases]
    outs = [np.zeros(n, np.int32) for _ in range(K)]
    outs_seq = [np.zeros(n, np.int32) for _ in range(K)]
    sides = (_lib.BowSide * K)()
    for k, c in enumerate(cases):
        sides[k] = _lib.BowSide(ptr(c["d1"]).value, n, C.pointer(f1[k].c), ptr(c["v1"]).value, ptr(c["a1"]).value)
    side2 = _lib.BowSide(ptr(base["d2"]).value, n, C.pointer(f2.c), None, ptr(base["a2"]).value)
    op = (C.c_void_p * K)(*[ptr(o).value for o in outs])
    nm = np.zeros(K, np.int32)
    nm1 = C.c_int(0)

    def batch():
        check(lib.orbm_search_bow_batch(0, sides, K, C.byref(side2), 0.75, 1, 0, op, ptr(nm)))

    def seq():
        for k, c in enumerate(cases):
            check(lib.orbm_search_bow(0, ptr(c["d1"]), n, ptr(base["d2"]), n, C.byref(f1[k].c), C.byref(f2.c), ptr(c["v1"]), None,
                                      ptr(c["a1"]), ptr(base["a2"]), 0.75, 1, 0, ptr(outs_seq[k]), C.byref(nm1)))

    b, s = t(batch), t(seq)
    assert all(np.array_equal(a, c) for a, c in zip(outs, outs_seq))
    print(f"SearchByBoW K={K} {n}x{n} features, {nodes} nodes: batch {b:.0f} us, {K} sequential calls {s:.0f} us")
