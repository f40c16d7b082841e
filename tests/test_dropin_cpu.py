"""CPU-side checks of the boundary: the C-ABI library exports every symbol include/orbfront.h
declares, the C++ host classes compile against it, and -- without a GPU -- every device entry
point fails loudly (no CPU fallback) while the host-only ones work."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from orb_slam2_e_b200 import _lib
import dropin_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_all_exported_and_bound():
    text = open(os.path.join(ROOT, "include", "orbfront.h")).read()
    declared = set(re.findall(r"\b(orb[xm]?_[a-z0-9_]+)\s*\(", text))
    declared -= {"orbx_handle", "orbm_db"}
    L = _lib.load()
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in orbfront.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)


def test_dropin_harness_builds_and_exports():
    L = dropin_lib.load()
    for s in dropin_lib.SYMBOLS:
        assert hasattr(L, s)


def test_descriptor_distance_host_path():
    """ORBmatcher::DescriptorDistance (C++ class) == popcount of xor; host-only, no device."""
    L = dropin_lib.load()
    rng = np.random.default_rng(5)
    for _ in range(200):
        a = rng.integers(0, 256, 32, dtype=np.uint8)
        b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert L.dropin_distance(a, b) == int(np.unpackbits(a ^ b).sum())
    z = np.zeros(32, np.uint8)
    assert L.dropin_distance(z, z) == 0 and L.dropin_distance(z, ~z) == 256


def test_no_device_fails_loudly():
    L = _lib.load()
    if L.orb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    cfg = _lib.OrbxConfig(1000, 1.2, 8, 20, 7, 0, 0, 0, 0)
    h = C.c_void_p()
    assert L.orbx_create(C.byref(cfg), C.byref(h)) == _lib.ORB_E_NO_DEVICE
    assert b"device" in L.orb_last_error().lower()
    db = C.c_void_p()
    assert L.orbm_db_create(0, 0, C.byref(db)) == _lib.ORB_E_NO_DEVICE
    a = np.zeros((4, 32), np.uint8)
    out = np.zeros(4, np.int32)
    assert L.orbm_hamming_pairs(0, _lib.ptr(a), _lib.ptr(a), 4, _lib.ptr(out)) == _lib.ORB_E_NO_DEVICE
    # the C++ class refuses to construct (throws; the harness reports it)
    D = dropin_lib.load()
    assert D.dropin_create(1000, 1.2, 8, 20, 7) is None
    assert b"device" in D.dropin_error().lower()


def test_matcher_entry_points_have_no_cpu_path():
    """Every search entry point added for SURVEY 8f (windowed searches, triangulation, vocabulary,
    distinctive descriptors, stereo) validates its arguments and then needs a device: on a box
    without one each returns ORB_E_NO_DEVICE (bad arguments still report ORB_E_INVALID first), and
    the C++ drop-in methods throw instead of computing anything on the host."""
    import match_cases as mc
    import scene_lib as sl
    import projection_cases as pc
    from orb_slam2_e_b200 import matcher
    from orb_slam2_e_b200._lib import OrbFrontError
    L = _lib.load()
    if L.orb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    frame, q, qd, claimed = mc.window_case(1, n=200, nq=50)
    for call in (lambda: matcher.window_search(frame, q, qd, claimed),
                 lambda: matcher.window_candidates(frame, q, qd),
                 lambda: matcher.distinctive_descriptors(np.zeros((3, 32), np.uint8), [0, 3])):
        with pytest.raises(OrbFrontError) as e:
            call()
        assert e.value.code == _lib.ORB_E_NO_DEVICE
    with pytest.raises(OrbFrontError) as e:   # argument errors come first
        matcher.distinctive_descriptors(np.zeros((3, 32), np.uint8), [1, 3])
    assert e.value.code == _lib.ORB_E_INVALID
    parent = np.array([0, 0, 0], np.int32)
    h = C.c_void_p()
    assert L.orbm_vocab_create(0, 10, 1, 3, _lib.ptr(parent), _lib.ptr(np.array([0, 1, 1], np.uint8)), _lib.ptr(np.zeros((3, 32), np.uint8)),
                               _lib.ptr(np.ones(3)), C.byref(h)) == _lib.ORB_E_NO_DEVICE
    # the product header on top: SearchByProjection throws std::runtime_error (reported by the harness)
    W = sl.World(3, n_points=300)
    rng = np.random.default_rng(1)
    T = sl.pose(rng)
    F = W.observe(T, 5)
    P = W.points(track=pc.frustum_fields(W, T, rng))
    with pytest.raises(RuntimeError) as e:
        sl.dropin_side().sbp_local(F, P, np.arange(P.n), 1.0)
    assert "device" in str(e.value).lower()
