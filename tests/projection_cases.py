"""Scene-level comparisons of two builds of ORBmatcher (see scene_lib.py): `ref` is the reference's own
src/ORBmatcher.cc, `got` the product header over the GPU library (GPU tests) or over the CPU
restatement (CPU tests).  Every output must be equal."""
import numpy as np

import scene_lib as sl


def frustum_fields(W, T, rng):
    """What Frame::isInFrustum (Frame.cc:283-340) leaves in the MapPoints before SearchLocalPoints."""
    u, v, z = W.project(T)
    Ow = -T[:3, :3].T @ T[:3, 3]
    dist = np.linalg.norm(W.pos - Ow, axis=1)
    level = np.clip(np.ceil(np.log(W.max_dist / dist) / np.log(W.scale)), 0, W.n_levels - 1)
    inside = (z > 0) & (u >= 0) & (u < W.width) & (v >= 0) & (v < W.height)
    return dict(in_view=inside & (rng.random(len(u)) < 0.9), proj_x=u, proj_y=v, proj_xr=u - W.cam["mbf"] / z, level=level,
                view_cos=np.where(rng.random(len(u)) < 0.5, 0.9995, rng.uniform(0.5, 0.998, len(u))))


def search_local_points(ref, got, stereo, th):
    W = sl.World(21 + int(th), stereo=stereo)
    rng = np.random.default_rng(5)
    T = sl.pose(rng)
    F = W.observe(T, 77, assign=0.15)
    P = W.points(track=frustum_fields(W, T, rng))
    order = rng.permutation(P.n)[:1300]
    a, b = ref.sbp_local(F, P, order, th), got.sbp_local(F, P, order, th)
    assert a["n"] > 150
    sl.assert_same(a, b, "SearchByProjection(F, vpMapPoints)")


def track_with_motion_model(ref, got, mono, stereo, th, motion):
    W = sl.World(31, stereo=stereo)
    rng = np.random.default_rng(int(th) + len(motion))
    Tl = sl.pose(rng)
    Tc = sl.pose(rng)
    if motion != "still":  # move along the optical axis by more than the baseline (bForward / bBackward, :1547-1548)
        Tc = Tl.copy()
        Tc[2, 3] += -0.5 if motion == "forward" else 0.5
    last = W.observe(Tl, 101, assign=0.8, outliers=0.05)
    cur = W.observe(Tc, 102, assign=0.1)
    P = W.points()
    a = ref.sbp_last_frame(cur, last, P, th, mono)
    b = got.sbp_last_frame(cur, last, P, th, mono)
    assert a["n"] > 100
    sl.assert_same(a, b, "SearchByProjection(CurrentFrame, LastFrame)")
    a = ref.sbp_last_frame(cur, last, P, th, mono, check_ori=False)
    b = got.sbp_last_frame(cur, last, P, th, mono, check_ori=False)
    sl.assert_same(a, b, "SearchByProjection(CurrentFrame, LastFrame), no orientation check")


def relocalization_refinement(ref, got, th, orb_dist):
    W = sl.World(41)
    rng = np.random.default_rng(9)
    kf = W.observe(sl.pose(rng), 201, assign=0.85)
    cur = W.observe(sl.pose(rng), 202, assign=0.2)
    found = rng.choice(W.pos.shape[0], 200, replace=False)
    P = W.points()
    a = ref.sbp_reloc_kf(cur, kf, P, found, th, orb_dist)
    b = got.sbp_reloc_kf(cur, kf, P, found, th, orb_dist)
    assert a["n"] > 50
    sl.assert_same(a, b, "SearchByProjection(CurrentFrame, pKF, sAlreadyFound)")


def loop_closure_sim3(ref, got, scale):
    W = sl.World(51)
    rng = np.random.default_rng(3)
    T = sl.pose(rng)
    kf = W.observe(T, 301, assign=0.3)
    S = T.copy()
    S[:3, :] *= np.float32(scale)  # Scw = [s R | s t]
    P = W.points()
    order = rng.permutation(P.n)[:1000]
    matched = kf.mp.copy()
    a = ref.sbp_sim3_kf(kf, S, P, order, matched, 10)
    b = got.sbp_sim3_kf(kf, S, P, order, matched, 10)
    assert a["n"] > 50
    sl.assert_same(a, b, "SearchByProjection(pKF, Scw)")


def map_wide_pose_candidate(ref, got, th):
    W = sl.World(61, n_points=2500)
    rng = np.random.default_rng(4)
    T = sl.pose(rng)
    F = W.observe(T, 401, assign=0.1)
    P = W.points()
    R, t = T[:3, :3].astype(np.float64), T[:3, 3].astype(np.float64)
    a, b = ref.sbp_map(F, P, R, t, th), got.sbp_map(F, P, R, t, th)
    assert a["n"] > 100
    sl.assert_same(a, b, "SearchByProjection(F, pMap, Rcw, tcw)")


def bow_and_distance(ref, got):
    """SearchByBoW x2 and DescriptorDistance of the product header against the reference's own code."""
    import match_cases as mc
    rng = np.random.default_rng(8)
    d = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    for i in range(0, 64, 2):
        assert ref._call("distance", d[i].ctypes.data, d[i + 1].ctypes.data) == got._call("distance", d[i].ctypes.data, d[i + 1].ctypes.data)
    for kf_kf in (0, 1):
        for check_ori in (0, 1):
            case = mc.bow_case(17 + kf_kf, n1=900, n2=950, n_nodes=60)
            n1, off1, i1 = mc.flat_fv(case["fv1"])
            n2, off2, i2 = mc.flat_fv(case["fv2"])
            st1 = np.where(case["v1"] > 0, 1, rng.integers(0, 3, len(case["v1"])) * 2 % 4 // 2 * 2).astype(np.uint8)
            st2 = np.where(case["v2"] > 0, 1, 0).astype(np.uint8)
            outs = []
            for side in (ref, got):
                out = np.zeros(len(case["d1"]) if kf_kf else len(case["d2"]), np.int32)
                p = lambda a: a.ctypes.data
                nm = side._call("search_bow", p(case["d1"]), len(case["d1"]), p(case["d2"]), len(case["d2"]), p(n1), p(off1), p(i1),
                                len(n1), p(n2), p(off2), p(i2), len(n2), p(st1), p(st2), p(case["a1"]), p(case["a2"]), 0.75,
                                check_ori, kf_kf, p(out))
                outs.append((nm, out))
            assert outs[0][0] == outs[1][0] and outs[0][0] > 50
            assert np.array_equal(outs[0][1], outs[1][1])


def fuse_neighbors(ref, got, stereo, th):
    """LocalMapping::SearchInNeighbors: the current keyframe's points into a neighbour (:557, :582)."""
    W = sl.World(71, stereo=stereo)
    rng = np.random.default_rng(12)
    kf = W.observe(sl.pose(rng), 501, assign=0.5)
    P = W.points()
    order = rng.permutation(P.n)[:1100]
    order[rng.random(len(order)) < 0.05] = -1          # NULL entries (:1044-1045)
    order[10:20] = order[0:10]                         # duplicates: the second visit finds the point in the keyframe
    a, b = ref.fuse(kf, P, order, th), got.fuse(kf, P, order, th)
    assert a["n"] > 100 and (a["replaced"] >= 0).sum() > 10
    sl.assert_same(a, b, "Fuse(pKF, vpMapPoints)")


def fuse_loop(ref, got, scale):
    """LoopClosing::SearchAndFuse (:599)."""
    W = sl.World(72)
    rng = np.random.default_rng(13)
    T = sl.pose(rng)
    kf = W.observe(T, 502, assign=0.4)
    S = T.copy()
    S[:3, :] *= np.float32(scale)
    P = W.points()
    order = rng.permutation(P.n)[:1000]
    a, b = ref.fuse_sim3(kf, S, P, order), got.fuse_sim3(kf, S, P, order)
    assert a["n"] > 100 and (a["replace"] >= 0).sum() > 10
    sl.assert_same(a, b, "Fuse(pKF, Scw)")


def sim3_mutual(ref, got, s12):
    """LoopClosing::ComputeSim3 (:323): KF1 and KF2 see the same map from two poses; T12 maps camera 2 to camera 1."""
    W = sl.World(73)
    rng = np.random.default_rng(14)
    T1, T2 = sl.pose(rng, 3.0, 0.2), sl.pose(rng, 3.0, 0.2)
    kf1 = W.observe(T1, 503, assign=0.8)
    kf2 = W.observe(T2, 504, assign=0.8)
    T12 = T1.astype(np.float64) @ np.linalg.inv(T2.astype(np.float64))
    R12, t12 = T12[:3, :3], T12[:3, 3] * s12
    P = W.points()
    m = np.full(kf1.n, -1, np.int32)
    pre = np.flatnonzero(kf1.mp >= 0)[:40]
    m[pre] = kf1.mp[pre]                                # already matched (:1336-1346)
    a = ref.search_by_sim3(kf1, kf2, P, m, s12, R12, t12)
    b = got.search_by_sim3(kf1, kf2, P, m, s12, R12, t12)
    assert a["n"] > 50
    sl.assert_same(a, b, "SearchBySim3")


def _feature_vector(truth, rng, n_nodes=90):
    """DBoW2-like FeatureVector: keypoints of the same map point mostly fall in the same node."""
    node = np.where(truth >= 0, (truth * 2654435761 % 1000003) % n_nodes, rng.integers(0, n_nodes, len(truth)))
    flip = rng.random(len(truth)) < 0.1
    node[flip] = rng.integers(0, n_nodes, int(flip.sum()))
    node = node * 5 + 2
    ids = np.unique(node)
    off, idx = [0], []
    for i in ids:
        members = np.flatnonzero(node == i)
        idx.extend(rng.permutation(members))
        off.append(len(idx))
    return np.ascontiguousarray(ids, np.uint32), np.ascontiguousarray(off, np.int32), np.ascontiguousarray(idx, np.uint32)


def triangulation(ref, got, stereo, only_stereo, check_ori):
    """LocalMapping::CreateNewMapPoints (:315): unmatched keypoints of two keyframes, epipolar gate from F12."""
    W = sl.World(81, stereo=stereo)
    rng = np.random.default_rng(15)
    T1, T2 = sl.pose(rng, 2.0, 0.3), sl.pose(rng, 2.0, 0.3)
    kf1 = W.observe(T1, 601, assign=0.3, bit_noise=40)
    truth1 = W.last_truth
    kf2 = W.observe(T2, 602, assign=0.3, bit_noise=40)
    truth2 = W.last_truth
    fv1, fv2 = _feature_vector(truth1, rng), _feature_vector(truth2, rng)
    # LocalMapping::ComputeF12 (LocalMapping.cc:596-614): F12 = K1^-T [t12]x R12 K2^-1
    R1, t1, R2, t2 = T1[:3, :3].astype(np.float64), T1[:3, 3].astype(np.float64), T2[:3, :3].astype(np.float64), T2[:3, 3].astype(np.float64)
    R12 = R1 @ R2.T
    t12 = -R1 @ R2.T @ t2 + t1
    tx = np.array([[0, -t12[2], t12[1]], [t12[2], 0, -t12[0]], [-t12[1], t12[0], 0]])
    K = np.array([[W.cam["fx"], 0, W.cam["cx"]], [0, W.cam["fy"], W.cam["cy"]], [0, 0, 1.0]])
    F12 = np.linalg.inv(K).T @ tx @ R12 @ np.linalg.inv(K)
    P = W.points()
    a = ref.search_triangulation(kf1, kf2, P, fv1, fv2, F12, only_stereo, check_ori)
    b = got.search_triangulation(kf1, kf2, P, fv1, fv2, F12, only_stereo, check_ori)
    assert a["n"] > 40, a["n"]
    sl.assert_same(a, b, "SearchForTriangulation")


def initialization(ref, got, window, ratio, check_ori):
    """Tracking::MonocularInitialization (Tracking.cc:718,791): level-0 keypoints of the initial frame
    searched around their previous position in the current frame; later queries steal keypoints."""
    W = sl.World(91, n_points=2500)
    rng = np.random.default_rng(16)
    f1 = W.observe(sl.pose(rng, 1.0, 0.05), 701, clutter=0.5)
    f2 = W.observe(sl.pose(rng, 1.0, 0.05), 702, clutter=0.5)
    f1.octave[:] = np.where(rng.random(f1.n) < 0.6, 0, f1.octave)   # plenty of level-0 keypoints on both sides
    f2.octave[:] = np.where(rng.random(f2.n) < 0.6, 0, f2.octave)
    prev = np.stack([f1.x, f1.y], 1) + rng.normal(0, 8.0, (f1.n, 2))
    a = ref.search_for_initialization(f1, f2, prev, window, ratio, check_ori)
    b = got.search_for_initialization(f1, f2, prev, window, ratio, check_ori)
    assert a["n"] > 30, a["n"]
    sl.assert_same(a, b, "SearchForInitialization")
