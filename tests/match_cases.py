"""Seeded inputs for the matcher parity tests (shared by the CPU-oracle and the GPU tests)."""
import numpy as np
from orb_slam2_e_b200 import synth


def bow_case(seed, n1=300, n2=320, n_nodes=40, shared=0.7, dup_frac=0.5, valid_frac=0.8):
    """Two feature sets with planted correspondences, random vocabulary-node assignment,
    validity flags and angles.  Returns dict of flat arrays + FeatureVector dicts."""
    rng = np.random.default_rng(seed)
    d1 = synth.random_descriptors(n1, seed * 3 + 1)
    d2 = synth.random_descriptors(n2, seed * 3 + 2)
    node1 = rng.integers(0, n_nodes, n1)
    node2 = rng.integers(0, n_nodes, n2)
    ang1 = rng.uniform(0, 360, n1).astype(np.float32)
    ang2 = rng.uniform(0, 360, n2).astype(np.float32)
    # plant matches: d2[j] = noisy copy of d1[i], same node, coherent rotation (+ outliers)
    m = int(min(n1, n2) * dup_frac)
    src = rng.choice(n1, m, replace=False)
    dst = rng.choice(n2, m, replace=False)
    for i, j in zip(src, dst):
        d2[j] = synth.flip_bits(d1[i], int(rng.integers(0, 50)), rng)
        if rng.random() < shared:
            node2[j] = node1[i]
        ang2[j] = np.float32((ang1[i] - (25.0 if rng.random() < 0.8 else rng.uniform(0, 360)) + rng.normal(0, 3)) % 360)
    # near-ties: a second copy of some d1 rows in the same node (ratio test / claim order)
    for i in src[: m // 5]:
        j = int(rng.integers(0, n2))
        d2[j] = synth.flip_bits(d1[i], int(rng.integers(0, 50)), rng)
        node2[j] = node1[i]
    fv1, fv2 = {}, {}
    for i in rng.permutation(n1):  # insertion order inside a node is arbitrary
        fv1.setdefault(int(node1[i]) * 7 + 3, []).append(int(i))
    for j in rng.permutation(n2):
        fv2.setdefault(int(node2[j]) * 7 + 3, []).append(int(j))
    # some nodes exist on one side only
    for k in list(fv1)[:3]:
        fv1[k + 1] = fv1.pop(k)
    v1 = (rng.random(n1) < valid_frac).astype(np.uint8)
    v2 = (rng.random(n2) < valid_frac).astype(np.uint8)
    return dict(d1=d1, d2=d2, fv1=fv1, fv2=fv2, v1=v1, v2=v2, a1=ang1, a2=ang2)


def bow_candidates(base, k, n1=400, seed=0):
    """k candidate keyframes for the FRAME of `base` (its side 2): each has planted noisy copies of frame descriptors in
    the frame feature's vocabulary node with a coherent rotation, plus unrelated features.  Returns bow_case-like dicts
    sharing d2 / fv2 / v2 / a2 with `base` (the K searches of one relocalization, Tracking.cc:1768-1795)."""
    node_of = {}
    for node, lst in base["fv2"].items():
        for j in lst:
            node_of[j] = node
    nodes = sorted(base["fv2"])
    n2 = len(base["d2"])
    out = []
    for c in range(k):
        rng = np.random.default_rng(seed * 1000 + c)
        n = n1 + 37 * c
        d1 = synth.random_descriptors(n, seed * 77 + c)
        node1 = [nodes[int(x)] for x in rng.integers(0, len(nodes), n)]
        ang1 = rng.uniform(0, 360, n).astype(np.float32)
        m = min(n, n2) // 2
        for i, j in zip(rng.choice(n, m, replace=False), rng.choice(n2, m, replace=False)):
            d1[i] = synth.flip_bits(base["d2"][j], int(rng.integers(0, 50)), rng)
            if rng.random() < 0.7:
                node1[i] = node_of[int(j)]
            ang1[i] = np.float32((base["a2"][j] + (25.0 if rng.random() < 0.8 else rng.uniform(0, 360)) + rng.normal(0, 3)) % 360)
        fv1 = {}
        for i in rng.permutation(n):
            fv1.setdefault(int(node1[i]), []).append(int(i))
        v1 = (rng.random(n) < 0.8).astype(np.uint8)
        out.append(dict(d1=d1, d2=base["d2"], fv1=fv1, fv2=base["fv2"], v1=v1, v2=base["v2"], a1=ang1, a2=base["a2"]))
    return out


def flat_fv(fv):
    nodes = sorted(fv)
    off = np.zeros(len(nodes) + 1, np.int32)
    off[1:] = np.cumsum([len(fv[k]) for k in nodes])
    idx = np.concatenate([np.asarray(fv[k], np.uint32) for k in nodes]) if nodes else np.zeros(0, np.uint32)
    return np.asarray(nodes, np.uint32), off, idx.astype(np.uint32)


def oracle_bow(L, case, ratio, check_ori, kf_kf):
    n1, off1, i1 = flat_fv(case["fv1"])
    n2, off2, i2 = flat_fv(case["fv2"])
    out = np.zeros(len(case["d1"]) if kf_kf else len(case["d2"]), np.int32)
    v2 = case["v2"].ctypes.data if kf_kf else None
    nm = L.orc_search_bow(case["d1"], len(case["d1"]), case["d2"], len(case["d2"]), n1, off1, i1, len(n1),
                          n2, off2, i2, len(n2), case["v1"], v2, case["a1"], case["a2"], ratio, int(check_ori),
                          int(kf_kf), out)
    return nm, out


def stereo_case(seed, nl=400, nr=420, n_rows=376, width=1241, n_levels=8):
    rng = np.random.default_rng(seed)
    from orb_slam2_e_b200._lib import KEYPOINT_DTYPE
    scale = (1.2 ** np.arange(n_levels)).astype(np.float32)
    kl = np.zeros(nl, KEYPOINT_DTYPE)
    kr = np.zeros(nr, KEYPOINT_DTYPE)
    for k, n in ((kl, nl), (kr, nr)):
        k["octave"] = rng.integers(0, n_levels, n)
        k["x"] = (rng.integers(19, width - 19, n) * 1.0).astype(np.float32) * 1.0
        k["y"] = rng.uniform(0, n_rows - 1, n).astype(np.float32)
    dl = synth.random_descriptors(nl, seed + 5)
    dr = synth.random_descriptors(nr, seed + 6)
    m = min(nl, nr) // 2
    for i, j in zip(rng.choice(nl, m, replace=False), rng.choice(nr, m, replace=False)):
        dr[j] = synth.flip_bits(dl[i], int(rng.integers(0, 110)), rng)
        kr["y"][j] = kl["y"][i] + np.float32(rng.normal(0, 1.0))
        kr["x"][j] = kl["x"][i] - np.float32(rng.uniform(-5, 90))
        kr["octave"][j] = np.clip(kl["octave"][i] + rng.integers(-2, 3), 0, n_levels - 1)
    kr["y"] = np.clip(kr["y"], 0, n_rows - 1)
    kl["x"][:3] = np.float32(-1.0)  # maxU < 0 case with min_d = 0
    return kl, dl, kr, dr, scale, n_rows


def oracle_stereo(L, kl, dl, kr, dr, scale, n_rows, min_d, max_d):
    bi = np.zeros(len(kl), np.int32)
    bd = np.zeros(len(kl), np.int32)
    c = np.ascontiguousarray
    L.orc_stereo_search(c(kl["x"]), c(kl["y"]), c(kl["octave"]), dl, len(kl), c(kr["x"]), c(kr["y"]), c(kr["octave"]),
                        dr, len(kr), scale, n_rows, min_d, max_d, bi, bd)
    return bi, bd


def window_case(seed, n=1500, nq=1200, width=640, height=480, n_levels=8, stereo=False, th_px=7.0, dup=0.6,
                clump=0.0, claimed_frac=0.1, nonblocking_frac=0.05, free_levels_frac=0.1):
    """A frame with a cell grid plus projected queries, shaped like Tracking's projection searches:
    most queries land near a keypoint that carries a noisy copy of their descriptor, several
    queries compete for the same keypoint (greedy claims), some keypoints are matched already.
    `clump` > 0 draws that fraction of the queries inside one small patch (long claim chains)."""
    from orb_slam2_e_b200.matcher import FrameGrid
    from orb_slam2_e_b200._lib import WINDOW_QUERY_DTYPE
    rng = np.random.default_rng(seed)
    x = rng.uniform(-3, width + 3, n).astype(np.float32)      # undistorted points may leave the image
    y = rng.uniform(-3, height + 3, n).astype(np.float32)
    octave = rng.integers(0, n_levels, n).astype(np.int32)
    desc = synth.random_descriptors(n, seed * 5 + 1)
    u_right = None
    if stereo:
        u_right = np.where(rng.random(n) < 0.7, x - rng.uniform(1, 40, n), -1).astype(np.float32)
    frame = FrameGrid(x, y, octave, desc, u_right, 0.0, 0.0, float(width), float(height))
    scale = (1.2 ** np.arange(n_levels)).astype(np.float32)
    q = np.zeros(nq, WINDOW_QUERY_DTYPE)
    qd = synth.random_descriptors(nq, seed * 5 + 2)
    target = rng.integers(0, n, nq)
    n_clump = int(nq * clump)
    if n_clump:
        cx, cy = rng.uniform(100, width - 100), rng.uniform(100, height - 100)
        near = np.argsort((x - cx) ** 2 + (y - cy) ** 2)[:40]
        target[:n_clump] = rng.choice(near, n_clump)
        rng.shuffle(target)
    for i in range(nq):
        t = int(target[i])
        lvl = int(np.clip(octave[t] + rng.integers(-1, 2), 0, n_levels - 1))
        q["x"][i] = x[t] + rng.normal(0, 2.0)
        q["y"][i] = y[t] + rng.normal(0, 2.0)
        q["r"][i] = np.float32(th_px) * scale[lvl]
        if rng.random() < free_levels_frac:
            q["min_level"][i], q["max_level"][i] = (0, -1) if rng.random() < 0.5 else (lvl, -1)
        else:
            q["min_level"][i], q["max_level"][i] = lvl - 1, lvl + int(rng.integers(0, 2))
        q["blocks"][i] = rng.random() >= nonblocking_frac
        if stereo and rng.random() < 0.8:
            q["u_right"][i] = (u_right[t] if u_right[t] > 0 else q["x"][i] - 10) + rng.normal(0, 3.0)
            q["er_max"][i] = q["r"][i]
        else:
            q["er_max"][i] = -1.0
        if rng.random() < dup:
            qd[i] = synth.flip_bits(desc[t], int(rng.integers(0, 70)), rng)
    # exact ties: the same descriptor on two keypoints of one window
    for i in range(0, nq, 17):
        t = int(target[i])
        d2 = (x - x[t]) ** 2 + (y - y[t]) ** 2
        nb = np.argsort(d2)[1:3]
        desc[nb[0]] = desc[t]
        octave[nb[0]] = octave[t]
    frame = FrameGrid(x, y, octave, desc, u_right, 0.0, 0.0, float(width), float(height))
    claimed = (rng.random(n) < claimed_frac).astype(np.uint8)
    return frame, q, qd, claimed


def oracle_window(L, frame, q, qd, claimed, th, ratio, claim, inv_level_sigma2=None, chi2_mono=5.99, chi2_stereo=7.8):
    from orb_slam2_e_b200._lib import WINDOW_RESULT_DTYPE
    out = np.zeros(len(q), WINDOW_RESULT_DTYPE)
    ur = None if frame.u_right is None else frame.u_right.ctypes.data
    cl = None if claimed is None else claimed.ctypes.data
    nm = L.orc_window_search(frame.x, frame.y, frame.octave, ur, frame.desc, frame.n, frame.min_x, frame.min_y,
                             frame.width_inv, frame.height_inv, frame.cols, frame.rows, frame.cell_off, frame.cell_idx,
                             q.ctypes.data, qd, len(q), cl, th, ratio, int(claim), out.ctypes.data,
                             None if inv_level_sigma2 is None else inv_level_sigma2.ctypes.data, chi2_mono, chi2_stereo)
    return nm, out


def oracle_stereo_matches(L, kl, dl, kr, dr, scale, inv_scale, levels_left, levels_right, mb, mbf):
    """Whole Frame::ComputeStereoMatches restatement on host pyramids (lists of 2-D u8 arrays)."""
    import ctypes as C
    n_levels = len(levels_left)
    PL = (C.c_void_p * n_levels)(*[a.ctypes.data for a in levels_left])
    PR = (C.c_void_p * n_levels)(*[a.ctypes.data for a in levels_right])
    for a, b in zip(levels_left, levels_right):
        assert a.shape == b.shape and a.strides == b.strides and a.strides[1] == 1
    step = np.array([a.strides[0] for a in levels_left], np.int32)
    w = np.array([a.shape[1] for a in levels_left], np.int32)
    h = np.array([a.shape[0] for a in levels_left], np.int32)
    f = lambda a: np.ascontiguousarray(a, np.float32)
    ur, z = np.zeros(len(kl), np.float32), np.zeros(len(kl), np.float32)
    L.orc_compute_stereo_matches(f(kl["x"]), f(kl["y"]), np.ascontiguousarray(kl["octave"], np.int32), dl, len(kl), f(kr["x"]),
                                 f(kr["y"]), np.ascontiguousarray(kr["octave"], np.int32), dr, len(kr), f(scale), f(inv_scale),
                                 n_levels, C.cast(PL, C.c_void_p), C.cast(PR, C.c_void_p), step, w, h, mb, mbf, ur, z)
    return ur, z
