"""Scenes for the ORBmatcher drop-in tests: a synthetic map, camera poses and frames, passed as flat
arrays to the C scene harness (tests/cpp/matcher_scenes.inc), which exists in three builds:

  dropin_*   tests/cpp/_build/libdropin.so      product header include/ORBmatcher.h -> liborbfront.so (GPU)
  refm_*     oracle/_ref/libmatchref.so          the reference's own src/ORBmatcher.cc, unmodified (CPU)
  portm_*    tests/cpp/_build/libdropin_port.so  product header over the CPU restatement (test-only backend)
"""
import ctypes as C
import os

import numpy as np

from orb_slam2_e_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
REF_PATH = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "libmatchref.so")
REF_FAST_PATH = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "libmatchref_fast.so")
PORT_PATH = os.path.join(HERE, "cpp", "_build", "libdropin_port.so")

_vp, _i, _f = C.c_void_p, C.c_int32, C.c_float


class ScnFrame(C.Structure):
    _fields_ = [("n", _i), ("x", _vp), ("y", _vp), ("octave", _vp), ("angle", _vp), ("u_right", _vp), ("desc", _vp),
                ("min_x", _f), ("min_y", _f), ("max_x", _f), ("max_y", _f), ("n_levels", _i), ("scale_factor", _f),
                ("fx", _f), ("fy", _f), ("cx", _f), ("cy", _f), ("mb", _f), ("mbf", _f), ("Tcw", _vp), ("mp", _vp),
                ("outlier", _vp)]


class ScnPoints(C.Structure):
    _fields_ = [("n", _i), ("pos", _vp), ("normal", _vp), ("desc", _vp), ("min_dist", _vp), ("max_dist", _vp),
                ("n_obs", _vp), ("bad", _vp), ("in_view", _vp), ("proj_x", _vp), ("proj_y", _vp), ("proj_xr", _vp),
                ("level", _vp), ("view_cos", _vp)]


def _p(a):
    return None if a is None else a.ctypes.data


class Frame:
    """Host-side description of one mock Frame; keeps the numpy arrays alive behind the C struct."""

    def __init__(self, x, y, octave, angle, desc, u_right=None, Tcw=None, mp=None, outlier=None, cam=None,
                 bounds=(0.0, 0.0, 640.0, 480.0), n_levels=8, scale_factor=1.2):
        self.x = np.ascontiguousarray(x, np.float32)
        self.y = np.ascontiguousarray(y, np.float32)
        self.octave = np.ascontiguousarray(octave, np.int32)
        self.angle = np.ascontiguousarray(angle, np.float32)
        self.desc = np.ascontiguousarray(desc, np.uint8)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        self.Tcw = None if Tcw is None else np.ascontiguousarray(Tcw, np.float32)
        self.mp = None if mp is None else np.ascontiguousarray(mp, np.int32)
        self.outlier = None if outlier is None else np.ascontiguousarray(outlier, np.uint8)
        cam = cam or dict(fx=520.0, fy=520.0, cx=320.0, cy=240.0, mb=0.08, mbf=41.6)
        self.cam = cam
        self.c = ScnFrame(len(self.x), _p(self.x), _p(self.y), _p(self.octave), _p(self.angle), _p(self.u_right), _p(self.desc),
                          bounds[0], bounds[1], bounds[2], bounds[3], n_levels, scale_factor, cam["fx"], cam["fy"], cam["cx"],
                          cam["cy"], cam["mb"], cam["mbf"], _p(self.Tcw), _p(self.mp), _p(self.outlier))
        self.n = len(self.x)


class Points:
    def __init__(self, pos, normal, desc, min_dist, max_dist, n_obs, bad, track=None):
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        self.pos, self.normal, self.desc = f32(pos), f32(normal), np.ascontiguousarray(desc, np.uint8)
        self.min_dist, self.max_dist = f32(min_dist), f32(max_dist)
        self.n_obs = np.ascontiguousarray(n_obs, np.int32)
        self.bad = np.ascontiguousarray(bad, np.uint8)
        self.n = len(self.pos)
        self.track = None
        t = [None] * 6
        if track is not None:
            self.track = (np.ascontiguousarray(track["in_view"], np.uint8), f32(track["proj_x"]), f32(track["proj_y"]),
                          f32(track["proj_xr"]), np.ascontiguousarray(track["level"], np.int32), f32(track["view_cos"]))
            t = [_p(a) for a in self.track]
        self.c = ScnPoints(self.n, _p(self.pos), _p(self.normal), _p(self.desc), _p(self.min_dist), _p(self.max_dist),
                           _p(self.n_obs), _p(self.bad), *t)


# ------------------------------------------------------------------------------------------------
def _rot(rng, deg):
    axis = rng.normal(size=3)
    axis /= np.linalg.norm(axis)
    a = np.deg2rad(deg)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * K @ K


def pose(rng, deg=2.0, trans=0.15):
    T = np.eye(4, dtype=np.float32)
    T[:3, :3] = _rot(rng, deg).astype(np.float32)
    T[:3, 3] = rng.normal(0, trans, 3).astype(np.float32)
    return T


class World:
    """Map points in front of the origin, each with a descriptor, a reference level and the scale
    invariance range MapPoint::UpdateNormalAndDepth would give it."""

    def __init__(self, seed, n_points=1500, n_levels=8, scale=1.2, stereo=False, width=640, height=480):
        self.rng = rng = np.random.default_rng(seed)
        self.n_levels, self.scale, self.stereo = n_levels, scale, stereo
        self.width, self.height = width, height
        self.cam = dict(fx=520.0, fy=520.0, cx=width / 2.0, cy=height / 2.0, mb=0.08, mbf=41.6)
        z = rng.uniform(3.0, 14.0, n_points)
        self.pos = np.stack([rng.uniform(-0.75, 0.75, n_points) * z, rng.uniform(-0.55, 0.55, n_points) * z, z], 1).astype(np.float32)
        d0 = np.linalg.norm(self.pos, axis=1)
        self.ref_level = rng.integers(0, n_levels, n_points)
        self.max_dist = (d0 * scale ** self.ref_level).astype(np.float32)
        self.min_dist = (self.max_dist / scale ** (n_levels - 1)).astype(np.float32)
        nrm = self.pos / d0[:, None] + rng.normal(0, 0.15, (n_points, 3))
        self.normal = (nrm / np.linalg.norm(nrm, axis=1)[:, None]).astype(np.float32)
        self.desc = synth.random_descriptors(n_points, seed * 7 + 1)
        # siblings: a third of the points sit next to the previous one and look like it, so windows
        # hold several plausible candidates and the order-dependent claims matter
        for i in range(1, n_points):
            if rng.random() < 0.33:
                self.pos[i] = self.pos[i - 1] + rng.normal(0, 0.03, 3).astype(np.float32)
                self.desc[i] = synth.flip_bits(self.desc[i - 1], int(rng.integers(3, 45)), rng)
                self.ref_level[i] = self.ref_level[i - 1]
        d0 = np.linalg.norm(self.pos, axis=1)
        self.max_dist = (d0 * scale ** self.ref_level).astype(np.float32)
        self.min_dist = (self.max_dist / scale ** (n_levels - 1)).astype(np.float32)
        self.n_obs = np.where(rng.random(n_points) < 0.06, 0, rng.integers(1, 9, n_points)).astype(np.int32)
        self.bad = (rng.random(n_points) < 0.04).astype(np.uint8)
        self.angle0 = rng.uniform(0, 360, n_points).astype(np.float32)

    def project(self, Tcw):
        pc = self.pos @ Tcw[:3, :3].T + Tcw[:3, 3]
        u = self.cam["fx"] * pc[:, 0] / pc[:, 2] + self.cam["cx"]
        v = self.cam["fy"] * pc[:, 1] / pc[:, 2] + self.cam["cy"]
        return u, v, pc[:, 2]

    def points(self, track=None):
        return Points(self.pos, self.normal, self.desc, self.min_dist, self.max_dist, self.n_obs, self.bad, track)

    def observe(self, Tcw, seed, frac=0.7, clutter=0.35, px_noise=1.5, bit_noise=55, rot=25.0, assign=0.0, outliers=0.0):
        """A frame that sees `frac` of the visible map points (noisy pixel position, noisy descriptor,
        level near the predicted one, coherent in-plane rotation `rot`) plus clutter keypoints.
        `assign`: fraction of the true observations whose MapPoint is already set in mvpMapPoints."""
        rng = np.random.default_rng(seed)
        u, v, z = self.project(Tcw)
        vis = np.flatnonzero((z > 0.5) & (u > 2) & (u < self.width - 2) & (v > 2) & (v < self.height - 2))
        seen = rng.choice(vis, int(len(vis) * frac), replace=False)
        n_cl = int(len(seen) * clutter)
        n = len(seen) + n_cl
        Ow = -Tcw[:3, :3].T @ Tcw[:3, 3]
        dist = np.linalg.norm(self.pos[seen] - Ow, axis=1)
        lvl = np.ceil(np.log(self.max_dist[seen] / dist) / np.log(self.scale))
        lvl = np.clip(lvl + rng.integers(-1, 2, len(seen)), 0, self.n_levels - 1)
        x = np.concatenate([u[seen] + rng.normal(0, px_noise, len(seen)), rng.uniform(0, self.width, n_cl)])
        y = np.concatenate([v[seen] + rng.normal(0, px_noise, len(seen)), rng.uniform(0, self.height, n_cl)])
        octave = np.concatenate([lvl, rng.integers(0, self.n_levels, n_cl)])
        desc = synth.random_descriptors(n, seed * 11 + 3)
        for k, p in enumerate(seen):
            desc[k] = synth.flip_bits(self.desc[p], int(rng.integers(0, bit_noise)), rng)
        for k in range(len(seen), n):  # clutter that imitates a nearby true observation
            if rng.random() < 0.6:
                j = int(rng.integers(0, len(seen)))
                x[k], y[k] = x[j] + rng.normal(0, 6), y[j] + rng.normal(0, 6)
                octave[k] = np.clip(octave[j] + rng.integers(-1, 2), 0, self.n_levels - 1)
                desc[k] = synth.flip_bits(desc[j], int(rng.integers(10, 90)), rng)
        ang = np.concatenate([(self.angle0[seen] - rot + rng.normal(0, 4, len(seen))) % 360, rng.uniform(0, 360, n_cl)])
        ang[rng.random(n) < 0.1] = rng.uniform(0, 360)
        ur = None
        if self.stereo:
            ur = np.concatenate([u[seen] - self.cam["mbf"] / z[seen] + rng.normal(0, 1.0, len(seen)), np.full(n_cl, -1.0)])
            ur[rng.random(n) < 0.25] = -1.0
        mp = np.full(n, -1, np.int32)
        k = rng.random(len(seen)) < assign
        mp[:len(seen)][k] = seen[k]
        perm = rng.permutation(n)  # keypoint order is unrelated to map order
        out = None
        if outliers > 0:
            out = (rng.random(n) < outliers).astype(np.uint8)
        self.last_truth = np.concatenate([seen, np.full(n_cl, -1)])[perm]  # map point behind each keypoint (-1: clutter)
        fr = Frame(x[perm], y[perm], octave[perm], ang[perm], desc[perm], None if ur is None else ur[perm], Tcw, mp[perm],
                   None if out is None else out[perm], self.cam, (0.0, 0.0, float(self.width), float(self.height)),
                   self.n_levels, self.scale)
        return fr


# ------------------------------------------------------------------------------------------------
def _declare(L, prefix):
    fp, pp = C.POINTER(ScnFrame), C.POINTER(ScnPoints)
    g = lambda name: getattr(L, prefix + name)
    g("scene_error").restype = C.c_char_p
    g("sbp_local").argtypes = [fp, pp, _vp, C.c_int, _f, _f, _vp, _vp, _vp]
    g("sbp_last_frame").argtypes = [fp, fp, pp, _f, C.c_int, _f, C.c_int, _vp]
    g("sbp_reloc_kf").argtypes = [fp, fp, pp, _vp, C.c_int, _f, C.c_int, _f, C.c_int, _vp]
    g("sbp_sim3_kf").argtypes = [fp, _vp, pp, _vp, C.c_int, _vp, C.c_int]
    g("sbp_map").argtypes = [fp, pp, _vp, _vp, _f, _f, _vp, _vp, _vp]
    g("distance").argtypes = [_vp, _vp]
    g("fuse").argtypes = [fp, pp, _vp, C.c_int, _f, _vp, _vp, _vp, _vp]
    g("fuse_sim3").argtypes = [fp, _vp, pp, _vp, C.c_int, _f, _vp, _vp, _vp]
    g("search_by_sim3").argtypes = [fp, fp, pp, _vp, _f, _vp, _vp, _f]
    g("search_for_initialization").argtypes = [fp, fp, _vp, _vp, C.c_int, _f, C.c_int]
    g("search_triangulation").argtypes = [fp, fp, pp, _vp, _vp, _vp, C.c_int, _vp, _vp, _vp, C.c_int, _vp, C.c_int, C.c_int,
                                          _vp, C.c_int]
    g("time_tracking_searches").argtypes = [fp, fp, pp, _vp, C.c_int, _f, _f, C.c_int, C.c_int, C.POINTER(C.c_double),
                                            C.POINTER(C.c_double)]
    g("time_tracking_searches").restype = C.c_double
    g("search_bow").argtypes = [_vp, C.c_int, _vp, C.c_int, _vp, _vp, _vp, C.c_int, _vp, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp,
                                _f, C.c_int, C.c_int, _vp]


class Side:
    """One build of the scene harness."""

    def __init__(self, lib, prefix):
        self.L, self.prefix = lib, prefix
        _declare(lib, prefix)

    def _call(self, name, *args):
        rc = getattr(self.L, self.prefix + name)(*args)
        if rc <= -1000000:
            raise RuntimeError(getattr(self.L, self.prefix + "scene_error")().decode())
        return rc

    def sbp_local(self, F, P, order, th, ratio=0.8):
        order = np.ascontiguousarray(order, np.int32)
        mp, inf, rad = np.zeros(F.n, np.int32), np.zeros(F.n, np.int32), np.zeros(F.n, np.float32)
        n = self._call("sbp_local", C.byref(F.c), C.byref(P.c), _p(order), len(order), th, ratio, _p(mp), _p(inf), _p(rad))
        return dict(n=n, mp=mp, in_frame=inf, radius=rad)

    def sbp_last_frame(self, cur, last, P, th, mono, ratio=0.9, check_ori=True):
        mp = np.zeros(cur.n, np.int32)
        n = self._call("sbp_last_frame", C.byref(cur.c), C.byref(last.c), C.byref(P.c), th, int(mono), ratio, int(check_ori), _p(mp))
        return dict(n=n, mp=mp)

    def sbp_reloc_kf(self, cur, kf, P, found, th, orb_dist, ratio=0.9, check_ori=True):
        found = np.ascontiguousarray(found, np.int32)
        mp = np.zeros(cur.n, np.int32)
        n = self._call("sbp_reloc_kf", C.byref(cur.c), C.byref(kf.c), C.byref(P.c), _p(found), len(found), th, orb_dist, ratio,
                       int(check_ori), _p(mp))
        return dict(n=n, mp=mp)

    def sbp_sim3_kf(self, kf, Scw, P, order, matched, th):
        order = np.ascontiguousarray(order, np.int32)
        m = np.ascontiguousarray(matched, np.int32).copy()
        S = np.ascontiguousarray(Scw, np.float32)
        n = self._call("sbp_sim3_kf", C.byref(kf.c), _p(S), C.byref(P.c), _p(order), len(order), _p(m), int(th))
        return dict(n=n, matched=m)

    def sbp_map(self, F, P, R, t, th, ratio=0.9):
        R = np.ascontiguousarray(R, np.float64)
        t = np.ascontiguousarray(t, np.float64)
        mp, mt, rel = np.zeros(F.n, np.int32), np.zeros(F.n, np.uint8), np.zeros((P.n, 4), np.float32)
        n = self._call("sbp_map", C.byref(F.c), C.byref(P.c), _p(R), _p(t), th, ratio, _p(mp), _p(mt), _p(rel))
        return dict(n=n, mp=mp, matched=mt, reloc=rel)


def time_tracking(side, cur, last, P, order, th_last=15.0, th_local=1.0, mono=True, reps=30):
    """(us SearchByProjection(cur, last), us SearchByProjection(cur, local points)), medians."""
    order = np.ascontiguousarray(order, np.int32)
    a, b = C.c_double(0), C.c_double(0)
    rc = getattr(side.L, side.prefix + "time_tracking_searches")(C.byref(cur.c), C.byref(last.c), C.byref(P.c), _p(order),
                                                                 len(order), th_last, th_local, int(mono), reps, C.byref(a), C.byref(b))
    if rc <= -1000000:
        raise RuntimeError("timing scene failed")
    return a.value, b.value


def _more_side_methods():
    def fuse(self, kf, P, order, th=3.0):
        order = np.ascontiguousarray(order, np.int32)
        kmp, rep, nobs, bad = np.zeros(kf.n, np.int32), np.zeros(P.n, np.int32), np.zeros(P.n, np.int32), np.zeros(P.n, np.uint8)
        n = self._call("fuse", C.byref(kf.c), C.byref(P.c), _p(order), len(order), th, _p(kmp), _p(rep), _p(nobs), _p(bad))
        return dict(n=n, kf_mp=kmp, replaced=rep, n_obs=nobs, bad=bad)

    def fuse_sim3(self, kf, Scw, P, order, th=4.0):
        order = np.ascontiguousarray(order, np.int32)
        S = np.ascontiguousarray(Scw, np.float32)
        kmp, rep, nobs = np.zeros(kf.n, np.int32), np.zeros(len(order), np.int32), np.zeros(P.n, np.int32)
        n = self._call("fuse_sim3", C.byref(kf.c), _p(S), C.byref(P.c), _p(order), len(order), th, _p(kmp), _p(rep), _p(nobs))
        return dict(n=n, kf_mp=kmp, replace=rep, n_obs=nobs)

    def search_by_sim3(self, kf1, kf2, P, matches12, s12, R12, t12, th=7.5):
        m = np.ascontiguousarray(matches12, np.int32).copy()
        R = np.ascontiguousarray(R12, np.float32)
        t = np.ascontiguousarray(t12, np.float32)
        n = self._call("search_by_sim3", C.byref(kf1.c), C.byref(kf2.c), C.byref(P.c), _p(m), s12, _p(R), _p(t), th)
        return dict(n=n, matches12=m)

    def search_triangulation(self, kf1, kf2, P, fv1, fv2, F12, only_stereo=False, check_ori=True):
        """fv = (node ids ascending u32, offsets i32, feature indices u32)."""
        F = np.ascontiguousarray(F12, np.float32)
        pairs = np.zeros((kf1.n, 2), np.int32)
        rc = self._call("search_triangulation", C.byref(kf1.c), C.byref(kf2.c), C.byref(P.c), _p(fv1[0]), _p(fv1[1]), _p(fv1[2]),
                        len(fv1[0]), _p(fv2[0]), _p(fv2[1]), _p(fv2[2]), len(fv2[0]), _p(F), int(only_stereo), int(check_ori),
                        _p(pairs), kf1.n)
        n, k = divmod(rc, 100000)
        return dict(n=n, pairs=pairs[:k].copy())

    Side.fuse, Side.fuse_sim3, Side.search_by_sim3 = fuse, fuse_sim3, search_by_sim3
    def search_for_initialization(self, f1, f2, prev, window=100, ratio=0.9, check_ori=True):
        pv = np.ascontiguousarray(prev, np.float32).copy()
        m = np.zeros(f1.n, np.int32)
        n = self._call("search_for_initialization", C.byref(f1.c), C.byref(f2.c), _p(pv), _p(m), int(window), ratio, int(check_ori))
        return dict(n=n, matches12=m, prev=pv)

    Side.search_triangulation = search_triangulation
    Side.search_for_initialization = search_for_initialization


_more_side_methods()


def ref_side(fast=False):
    path = REF_FAST_PATH if fast else REF_PATH
    if not os.path.exists(path):
        return None
    return Side(C.CDLL(path), "refm_")


def dropin_side():
    import dropin_lib
    dropin_lib.load()
    return Side(C.CDLL(dropin_lib.PATH), "dropin_")


def port_side():
    import subprocess
    subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "cpp"), "_build/libdropin_port.so"], stdout=subprocess.DEVNULL)
    return Side(C.CDLL(PORT_PATH), "portm_")


def assert_same(a, b, what=""):
    assert a.keys() == b.keys()
    for k in a:
        if isinstance(a[k], np.ndarray):
            assert a[k].shape == b[k].shape, (what, k, a[k].shape, b[k].shape)
            bad = np.flatnonzero(np.asarray(a[k] != b[k]).reshape(len(a[k]), -1).any(1))
            assert len(bad) == 0, (what, k, bad[:8], a[k][bad[:8]], b[k][bad[:8]])
        else:
            assert a[k] == b[k], (what, k, a[k], b[k])
