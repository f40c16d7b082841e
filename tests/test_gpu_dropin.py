"""GPU parity through the C++ host classes (include/ORBextractor.h, include/ORBmatcher.h), driven
like the reference's callers drive them, against the oracle -- and against the reference's own
ORBextractor.cc (oracle/_ref) when that build travelled with the snapshot."""
import numpy as np
import pytest

import dropin_lib
import match_cases as mc
import oracle_lib
from oracle_lib import PortExtractor, oracle
from orb_slam2_e_b200 import synth

pytestmark = pytest.mark.gpu


class DropinExtractor(oracle_lib._Ext):
    def __init__(self, nf, sf, nl, ini, mn):
        L = dropin_lib.load()
        super().__init__(L, "dropin", nf, sf, nl, ini, mn)
        assert self.h, L.dropin_error()

    def extract(self, img):  # strided input goes through as-is (cv::Mat with step)
        img = np.asarray(img, np.uint8)
        kps = np.zeros(self.cap, oracle_lib.KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = self.L.dropin_extract(self.h, img.ctypes.data, img.shape[1], img.shape[0], img.strides[0], kps.ctypes.data, desc, self.cap)
        assert n >= 0, (n, self.L.dropin_error())
        return kps[:n].copy(), desc[:n].copy()


def _same(a, b):
    (ak, ad), (bk, bd) = a, b
    assert len(ak) == len(bk)
    for name in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(ak[name], bk[name]), name
    assert np.abs(np.deg2rad(ak["angle"].astype(np.float64) - bk["angle"])).max(initial=0) <= 1e-3
    assert np.unpackbits(ad ^ bd).mean() <= 1e-4 if len(ad) else True


@pytest.mark.parametrize("cfg,shape,seed", [((1000, 1.2, 8, 20, 7), (480, 640), 3), ((1200, 1.1, 6, 24, 7), (480, 640), 12),
                                            ((2000, 1.2, 8, 20, 7), (376, 1241), 21)])
def test_operator_call_matches_reference(cfg, shape, seed):
    img = synth.textured_frame(*shape, seed) if cfg[0] != 1200 else synth.endoscopy_frame(*shape, seed)
    g = DropinExtractor(*cfg)
    got = g.extract(img)
    want_cls = oracle_lib.RefExtractor if oracle_lib.ref_lib() is not None else PortExtractor
    ref = want_cls(*cfg)
    _same(got, ref.extract(img))
    # getters and the public pyramid member, read like Frame.cc does
    assert g.L.dropin_levels(g.h) == cfg[2] and abs(g.L.dropin_scale_factor(g.h) - cfg[1]) < 1e-6
    port = PortExtractor(*cfg)
    port.extract(img)
    for l in range(cfg[2]):
        assert g.scale(l) == ref.scale(l)
        assert np.array_equal(g.level(l, True), port.level(l, True)), f"mvImagePyramid[{l}] incl. border"


def test_operator_call_edge_behaviour():
    g = DropinExtractor(500, 1.2, 3, 20, 7)
    k, d = g.extract(np.full((200, 260), 90, np.uint8))  # zero keypoints: descriptors released (checked in the harness)
    assert len(k) == 0
    big = synth.textured_frame(300, 420, 9)
    view = big[7:251, 11:390]  # non-contiguous rows
    _same(g.extract(view), PortExtractor(500, 1.2, 3, 20, 7).extract(np.ascontiguousarray(view)))


@pytest.mark.parametrize("kf_kf", [False, True])
@pytest.mark.parametrize("check_ori", [False, True])
def test_search_by_bow_through_cpp_class(kf_kf, check_ori):
    L = dropin_lib.load()
    for seed in (1, 2):
        case = mc.bow_case(seed)
        nm, out = mc.oracle_bow(oracle(), case, 0.75, check_ori, kf_kf)
        n1, off1, i1 = mc.flat_fv(case["fv1"])
        n2, off2, i2 = mc.flat_fv(case["fv2"])
        # validity -> MapPoint states: invalid entries alternate between NULL and bad MapPoints
        def states(v):
            s = v.astype(np.uint8).copy()
            bad = np.flatnonzero(v == 0)
            s[bad[::2]] = 2
            return s
        s1 = states(case["v1"])
        s2 = states(case["v2"]) if kf_kf else np.ones(len(case["d2"]), np.uint8)
        got = np.zeros(len(out), np.int32)
        gnm = L.dropin_search_bow(case["d1"], len(case["d1"]), case["d2"], len(case["d2"]), n1, off1, i1, len(n1), n2, off2, i2,
                                  len(n2), s1, s2, case["a1"], case["a2"], 0.75, int(check_ori), int(kf_kf), got)
        assert gnm == nm, (gnm, nm, L.dropin_error())
        assert np.array_equal(got, out)


@pytest.mark.parametrize("check_ori", [False, True])
def test_batched_search_by_bow_through_cpp_class(check_ori):
    """ORBmatcher::SearchByBoW(vector<KeyFrame*>, Frame&, ...) (the loop of Tracking.cc:1768-1795 as one device call) equals
    the reference's own SearchByBoW(KeyFrame*, Frame&, ...) run once per candidate (oracle/_ref/libmatchref.so when it
    travelled with the snapshot, else the restatement), with a bad keyframe among the candidates."""
    import ctypes as C
    L = dropin_lib.load()
    base = mc.bow_case(21, n1=500, n2=700, n_nodes=30)
    n2, off2, i2 = mc.flat_fv(base["fv2"])
    cases = mc.bow_candidates(base, 6, n1=350, seed=9)
    K = len(cases) + 1  # the last candidate is a bad keyframe
    def states(v):
        s = v.astype(np.uint8).copy()
        bad = np.flatnonzero(v == 0)
        s[bad[::2]] = 2
        return s
    arr = lambda xs: (C.c_void_p * K)(*[None if x is None else x.ctypes.data for x in xs])
    flat = [mc.flat_fv(c["fv1"]) for c in cases]
    st = [states(c["v1"]) for c in cases]
    outs = [np.zeros(700, np.int32) for _ in range(K)]
    nm = np.zeros(K, np.int32)
    d1 = [np.ascontiguousarray(c["d1"]) for c in cases]
    a1 = [np.ascontiguousarray(c["a1"], np.float32) for c in cases]
    rc = L.dropin_search_bow_batch(K, arr(d1 + [None]), np.array([len(x) for x in d1] + [0], np.int32), arr([f[0] for f in flat] + [None]),
                                   arr([f[1] for f in flat] + [None]), arr([f[2] for f in flat] + [None]),
                                   np.array([len(f[0]) for f in flat] + [0], np.int32), arr(st + [None]), arr(a1 + [None]),
                                   base["d2"], 700, n2, off2, i2, len(n2), np.ascontiguousarray(base["a2"], np.float32), 0.75,
                                   int(check_ori), arr(outs), nm)
    assert rc == 0, L.dropin_error()
    for k, c in enumerate(cases):
        want_nm, want = mc.oracle_bow(oracle(), c, 0.75, check_ori, False)
        assert nm[k] == want_nm and np.array_equal(outs[k], want), k
    assert nm[K - 1] == 0 and np.all(outs[K - 1] == -1)
    assert nm[:-1].sum() > 100


def test_stereo_search_through_cpp_class():
    L = dropin_lib.load()
    kl, dl, kr, dr, scale, n_rows = mc.stereo_case(4, nl=1500, nr=1600)
    bi, bd = mc.oracle_stereo(oracle(), kl, dl, kr, dr, scale, n_rows, 0.0, 80.0)
    gi = np.zeros(len(kl), np.int32); gd = np.zeros(len(kl), np.int32)
    kl = np.ascontiguousarray(kl); kr = np.ascontiguousarray(kr)
    assert L.dropin_stereo(kl.ctypes.data, dl, len(kl), kr.ctypes.data, dr, len(kr), np.ascontiguousarray(scale, np.float32),
                           len(scale), n_rows, 0.0, 80.0, gi, gd) == 0, L.dropin_error()
    assert np.array_equal(gd, bd) and np.array_equal(gi, bi)


def test_stereo_frame_through_cpp_classes_without_pyramid_download():
    """Two ORBextractor objects + ORBextractor::ComputeStereoMatches (whole Frame.cc:527-700 on the
    device-resident pyramids; SetPyramidDownload(false) on both) against the CPU restatement."""
    from orb_slam2_e_b200 import synth
    from orb_slam2_e_b200._lib import KEYPOINT_DTYPE
    from oracle_lib import PortExtractor
    L = dropin_lib.load()
    cfg = (1200, 1.2, 8, 20, 7)
    left, right = (np.ascontiguousarray(a) for a in synth.stereo_pair(240, 640, 12, max_disp=40.0))
    hl, hr = L.dropin_create(*cfg), L.dropin_create(*cfg)
    cap = 1500
    kps = np.zeros(cap, KEYPOINT_DTYPE); desc = np.zeros((cap, 32), np.uint8)
    ur = np.zeros(cap, np.float32); z = np.zeros(cap, np.float32)
    mbf, mb = 40.0, 0.08
    n = L.dropin_stereo_frame(hl, hr, left, right, 640, 240, 640, mb, mbf, kps.ctypes.data, desc, ur, z, cap)
    assert n > 0, L.dropin_error()
    pl, pr = PortExtractor(*cfg), PortExtractor(*cfg)
    okl, odl = pl.extract(left)
    okr, odr = pr.extract(right)
    assert kps[:n].tobytes() == okl.tobytes()
    scale = np.array([pl.scale(l) for l in range(cfg[2])], np.float32)
    our, oz = mc.oracle_stereo_matches(oracle(), okl, odl, okr, odr, scale, (np.float32(1.0) / scale).astype(np.float32),
                                       [pl.level(l) for l in range(cfg[2])], [pr.level(l) for l in range(cfg[2])], mb, mbf)
    assert (our >= 0).sum() > 100
    assert ur[:n].tobytes() == our.tobytes() and z[:n].tobytes() == oz.tobytes()
    L.dropin_destroy(hl); L.dropin_destroy(hr)
