"""Host-side mirror of ORB_SLAM2::ORBmatcher's descriptor-search interface over the C ABI.

Names and argument meaning follow the reference (include/ORBmatcher.h:37-111); object graphs
(KeyFrame*, Frame&, MapPoint*) are replaced by the flat arrays they are read from: descriptor
matrices (N x 32 u8), DBoW2 FeatureVectors as {node_id: [feature indices]}, MapPoint validity
flags and keypoint angles.  Everything runs in liborbfront.so on the GPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import FeatVec, GridFrame, WindowParams, TOP2_DTYPE, WINDOW_QUERY_DTYPE, WINDOW_RESULT_DTYPE, check, ptr

TH_HIGH, TH_LOW, HISTO_LENGTH = 95, 45, 30  # ORBmatcher.cc:37-40


def _desc(a) -> np.ndarray:
    a = np.ascontiguousarray(a, np.uint8)
    if a.ndim == 1:
        a = a.reshape(1, -1)
    if a.ndim != 2 or a.shape[1] != 32:
        raise ValueError("descriptors must be N x 32 uint8 (cv::Mat CV_8U rows)")
    return a


class _FlatFeatVec:
    """DBoW2::FeatureVector flattened to CSR with ascending node ids (std::map order)."""

    def __init__(self, fv):
        if isinstance(fv, dict):
            nodes = sorted(fv)
            lists = [np.asarray(fv[k], np.uint32) for k in nodes]
        else:  # (node_ids, offsets, indices) already flat
            nodes, off, idx = fv
            self.node = np.ascontiguousarray(nodes, np.uint32)
            self.off = np.ascontiguousarray(off, np.int32)
            self.idx = np.ascontiguousarray(idx, np.uint32)
            if np.any(np.diff(self.node.astype(np.int64)) <= 0):
                raise ValueError("node ids must be strictly ascending")
            self._mk()
            return
        self.node = np.asarray(nodes, np.uint32)
        self.off = np.zeros(len(nodes) + 1, np.int32)
        if lists:
            self.off[1:] = np.cumsum([len(x) for x in lists])
        self.idx = np.concatenate(lists).astype(np.uint32) if lists else np.zeros(0, np.uint32)
        self._mk()

    def _mk(self):
        self.c = FeatVec(self.node.ctypes.data, self.off.ctypes.data, self.idx.ctypes.data, len(self.node))


class ORBmatcher:
    """ORBmatcher(nnratio=0.6, checkOri=True) -- include/ORBmatcher.h:41."""

    TH_LOW, TH_HIGH, HISTO_LENGTH = TH_LOW, TH_HIGH, HISTO_LENGTH

    def __init__(self, nnratio: float = 0.6, checkOri: bool = True, device: int = 0):
        self.mfNNratio = float(np.float32(nnratio))
        self.mbCheckOrientation = bool(checkOri)
        self.device = device

    @staticmethod
    def DescriptorDistance(a, b) -> int:
        """static int DescriptorDistance(const cv::Mat&, const cv::Mat&) -- ORBmatcher.cc:1848."""
        a = np.ascontiguousarray(a, np.uint8).reshape(-1)
        b = np.ascontiguousarray(b, np.uint8).reshape(-1)
        if a.size != 32 or b.size != 32:
            raise ValueError("descriptors are 32 bytes")
        return _lib.load().orbm_hamming(ptr(a), ptr(b))

    def DescriptorDistances(self, a, b) -> np.ndarray:
        """Row-wise distances of two N x 32 matrices (device)."""
        a, b = _desc(a), _desc(b)
        if a.shape != b.shape:
            raise ValueError("shape mismatch")
        out = np.zeros(len(a), np.int32)
        check(_lib.load().orbm_hamming_pairs(self.device, ptr(a), ptr(b), len(a), ptr(out)))
        return out

    def SearchByBoW(self, desc1, featvec1, valid1, desc2, featvec2, valid2=None, angles1=None, angles2=None,
                    kf_kf: bool = False):
        """Both SearchByBoW overloads (ORBmatcher.cc:360-489 KF-Frame, :723-856 KF-KF).

        Returns (nmatches, matches): KF-Frame: matches[f] = KeyFrame feature matched to frame
        feature f (the reference stores that feature's MapPoint*), or -1.  KF-KF: matches[i1] =
        feature of KF2, or -1.
        """
        d1, d2 = _desc(desc1) if len(desc1) else np.zeros((0, 32), np.uint8), \
            _desc(desc2) if len(desc2) else np.zeros((0, 32), np.uint8)
        f1, f2 = _FlatFeatVec(featvec1), _FlatFeatVec(featvec2)
        v1 = np.ascontiguousarray(valid1, np.uint8)
        v2 = None if valid2 is None else np.ascontiguousarray(valid2, np.uint8)
        if kf_kf and v2 is None:
            raise ValueError("KF-KF search needs valid2")
        a1 = None if angles1 is None else np.ascontiguousarray(angles1, np.float32)
        a2 = None if angles2 is None else np.ascontiguousarray(angles2, np.float32)
        if self.mbCheckOrientation and (a1 is None or a2 is None):
            raise ValueError("orientation check needs keypoint angles")
        out = np.zeros(len(d1) if kf_kf else len(d2), np.int32)
        nm = C.c_int(0)
        check(_lib.load().orbm_search_bow(
            self.device, ptr(d1), len(d1), ptr(d2), len(d2), C.byref(f1.c), C.byref(f2.c), ptr(v1),
            None if v2 is None else ptr(v2), None if a1 is None else ptr(a1), None if a2 is None else ptr(a2),
            self.mfNNratio, int(self.mbCheckOrientation), int(kf_kf), ptr(out), C.byref(nm)))
        return nm.value, out

    def SearchByBoWBatch(self, candidates, desc2, featvec2, valid2=None, angles2=None, kf_kf: bool = False):
        """SearchByBoW of K candidate keyframes against one Frame / KeyFrame in ONE call (Tracking.cc:1768-1795 loops
        over the candidates; orbm_search_bow_batch).  candidates: list of (desc1, featvec1, valid1, angles1).
        Returns a list of (nmatches, matches), equal to K SearchByBoW calls."""
        keep = []
        d2 = _desc(desc2) if len(desc2) else np.zeros((0, 32), np.uint8)
        f2 = _FlatFeatVec(featvec2)
        v2 = None if valid2 is None else np.ascontiguousarray(valid2, np.uint8)
        a2 = None if angles2 is None else np.ascontiguousarray(angles2, np.float32)
        if kf_kf and v2 is None:
            raise ValueError("KF-KF search needs valid2")
        side2 = _lib.BowSide(ptr(d2).value if len(d2) else None, len(d2), C.pointer(f2.c), None if v2 is None else ptr(v2).value,
                             None if a2 is None else ptr(a2).value)
        k = len(candidates)
        sides = (_lib.BowSide * max(k, 1))()
        outs, out_ptrs = [], (C.c_void_p * max(k, 1))()
        for p, (desc1, featvec1, valid1, angles1) in enumerate(candidates):
            d1 = _desc(desc1) if len(desc1) else np.zeros((0, 32), np.uint8)
            f1 = _FlatFeatVec(featvec1)
            v1 = np.ascontiguousarray(valid1, np.uint8)
            a1 = None if angles1 is None else np.ascontiguousarray(angles1, np.float32)
            if self.mbCheckOrientation and (a1 is None or a2 is None):
                raise ValueError("orientation check needs keypoint angles")
            keep.append((d1, f1, v1, a1))
            sides[p] = _lib.BowSide(ptr(d1).value if len(d1) else None, len(d1), C.pointer(f1.c), ptr(v1).value if len(v1) else None,
                                    None if a1 is None else ptr(a1).value)
            o = np.zeros(len(d1) if kf_kf else len(d2), np.int32)
            outs.append(o)
            out_ptrs[p] = ptr(o).value if len(o) else None
        nm = np.zeros(max(k, 1), np.int32)
        check(_lib.load().orbm_search_bow_batch(self.device, sides, k, C.byref(side2), self.mfNNratio, int(self.mbCheckOrientation),
                                                int(kf_kf), out_ptrs, ptr(nm)))
        return [(int(nm[p]), outs[p]) for p in range(k)]

    def StereoSearch(self, kps_left, desc_left, kps_right, desc_right, scale_factors, n_rows, min_d, max_d):
        """Descriptor part of Frame::ComputeStereoMatches (Frame.cc:527-610): (bestIdxR, bestDist)."""
        kl = np.ascontiguousarray(kps_left, _lib.KEYPOINT_DTYPE)
        kr = np.ascontiguousarray(kps_right, _lib.KEYPOINT_DTYPE)
        dl = _desc(desc_left) if len(kl) else np.zeros((0, 32), np.uint8)
        dr = _desc(desc_right) if len(kr) else np.zeros((0, 32), np.uint8)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        bi = np.zeros(len(kl), np.int32)
        bd = np.zeros(len(kl), np.int32)
        check(_lib.load().orbm_stereo_search(self.device, ptr(kl), ptr(dl), len(kl), ptr(kr), ptr(dr), len(kr), ptr(sf),
                                             len(sf), int(n_rows), float(min_d), float(max_d), ptr(bi), ptr(bd)))
        return bi, bd


class FrameGrid:
    """The searched side of a projection search as the reference's Frame holds it: undistorted
    keypoints, mvuRight, descriptors and the 64 x 48 cell grid (Frame::AssignFeaturesToGrid,
    Frame.cc:245-260, with PosInGrid, :397-410).  `cells` (list of index lists, column-major
    (ix, iy) -> ix*rows+iy) overrides the assignment, for KeyFrame grids of other sizes."""

    def __init__(self, x, y, octave, desc, u_right=None, min_x=0.0, min_y=0.0, max_x=640.0, max_y=480.0,
                 cols=64, rows=48, cells=None):
        self.x = np.ascontiguousarray(x, np.float32)
        self.y = np.ascontiguousarray(y, np.float32)
        self.octave = np.ascontiguousarray(octave, np.int32)
        self.desc = _desc(desc) if len(self.x) else np.zeros((0, 32), np.uint8)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        self.n = len(self.x)
        f32 = np.float32
        self.min_x, self.min_y = f32(min_x), f32(min_y)
        self.cols, self.rows = int(cols), int(rows)
        self.width_inv = f32(f32(cols) / (f32(max_x) - self.min_x))    # Frame.cc:106-107
        self.height_inv = f32(f32(rows) / (f32(max_y) - self.min_y))
        if cells is None:
            def rnd(v):  # C round(): half away from zero
                v = v.astype(np.float64)
                return np.where(v >= 0, np.floor(v + 0.5), np.ceil(v - 0.5)).astype(np.int64)
            px = rnd((self.x - self.min_x) * self.width_inv)
            py = rnd((self.y - self.min_y) * self.height_inv)
            ok = (px >= 0) & (px < cols) & (py >= 0) & (py < rows)
            cell = np.where(ok, px * rows + py, cols * rows)
            order = np.argsort(cell, kind="stable")  # push_back in keypoint order inside a cell
            order = order[: int(ok.sum())]
            counts = np.bincount(cell[ok], minlength=cols * rows)
            self.cell_idx = np.ascontiguousarray(order, np.uint32)
        else:
            assert len(cells) == cols * rows
            counts = np.array([len(c) for c in cells], np.int64)
            self.cell_idx = np.ascontiguousarray([i for c in cells for i in c], np.uint32)
        self.cell_off = np.zeros(cols * rows + 1, np.int32)
        np.cumsum(counts, out=self.cell_off[1:])
        self.c = GridFrame(ptr(self.x), ptr(self.y), ptr(self.octave), None if self.u_right is None else ptr(self.u_right),
                           ptr(self.desc), self.n, float(self.min_x), float(self.min_y), float(self.width_inv),
                           float(self.height_inv), self.cols, self.rows, ptr(self.cell_off), ptr(self.cell_idx))


def window_search(frame: FrameGrid, queries, query_desc, claimed=None, th: int = TH_HIGH, ratio: float = 0.0,
                  claim: bool = True, device: int = 0, inv_level_sigma2=None, chi2_mono: float = 5.99, chi2_stereo: float = 7.8):
    """orbm_window_search: the descriptor loop of the SearchByProjection overloads
    (ORBmatcher.cc:46-132, :134-220, :491-604, :1529-1800); with inv_level_sigma2 also the
    reprojection gate of Fuse (:1118-1146).  Returns (results, n_accepted, passes)."""
    q = np.ascontiguousarray(queries, WINDOW_QUERY_DTYPE)
    qd = _desc(query_desc) if len(q) else np.zeros((0, 32), np.uint8)
    assert len(qd) == len(q)
    cl = None if claimed is None else np.ascontiguousarray(claimed, np.uint8)
    out = np.zeros(len(q), WINDOW_RESULT_DTYPE)
    na, rounds = C.c_int(0), C.c_int(0)
    sig = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
    prm = WindowParams(int(th), float(ratio), int(bool(claim)), None if sig is None else ptr(sig), 0 if sig is None else len(sig),
                       float(chi2_mono), float(chi2_stereo))
    check(_lib.load().orbm_window_search(device, C.byref(frame.c), ptr(q), ptr(qd), len(q), None if cl is None else ptr(cl),
                                         C.byref(prm), ptr(out), C.byref(na), C.byref(rounds)))
    return out, na.value, rounds.value


def window_candidates(frame: FrameGrid, queries, query_desc, device: int = 0):
    """orbm_window_candidates: per query the keypoints GetFeaturesInArea returns, in its order, with
    their Hamming distances.  Returns a list of (keypoint indices, distances) per query."""
    q = np.ascontiguousarray(queries, WINDOW_QUERY_DTYPE)
    qd = _desc(query_desc) if len(q) else np.zeros((0, 32), np.uint8)
    off, cnt = np.zeros(len(q), np.int32), np.zeros(len(q), np.int32)
    cap = max(1024, 64 * len(q))
    L = _lib.load()
    while True:
        idx, dist = np.zeros(cap, np.int32), np.zeros(cap, np.int32)
        total = C.c_int(0)
        rc = L.orbm_window_candidates(device, C.byref(frame.c), ptr(q), ptr(qd), len(q), ptr(off), ptr(cnt), ptr(idx), ptr(dist), cap,
                                      C.byref(total))
        if rc == -4:  # ORB_E_CAPACITY
            cap = total.value
            continue
        check(rc)
        break
    return [(idx[o:o + c].copy(), dist[o:o + c].copy()) for o, c in zip(off, cnt)]


def distinctive_descriptors(desc, offset, device: int = 0) -> np.ndarray:
    """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:305-370) for a batch of map points: point p
    owns rows offset[p]..offset[p+1] of desc; returns the index (inside its range) of the descriptor
    with the least median distance to the rest, -1 for an empty range."""
    off = np.ascontiguousarray(offset, np.int32)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    best = np.zeros(len(off) - 1, np.int32)
    check(_lib.load().orbm_distinctive_descriptors(device, ptr(d) if len(d) else None, ptr(off), len(off) - 1, ptr(best)))
    return best


class DescriptorDB:
    """One GPU-resident shard of the relocalization descriptor database:
    rows [row_base, row_base + n) of the global N x 32 matrix."""

    def __init__(self, device: int = 0, row_base: int = 0):
        self._h = C.c_void_p()
        self.device, self.row_base = device, row_base
        check(_lib.load().orbm_db_create(device, row_base, C.byref(self._h)))

    def append(self, rows):
        rows = _desc(rows)
        check(_lib.load().orbm_db_append(self._h, ptr(rows), len(rows), 0))

    def append_device(self, dev_ptr: int, n: int):
        check(_lib.load().orbm_db_append(self._h, C.c_void_p(dev_ptr), n, 1))

    def save(self, path: str):
        """Flat interchange file (16-byte header + n x 32 bytes), orbm_db_save."""
        check(_lib.load().orbm_db_save(self._h, str(path).encode()))

    def load(self, path: str, first: int = 0, count: int = -1):
        """Appends rows [first, first + count) of an interchange file (count < 0: to its end)."""
        check(_lib.load().orbm_db_load(self._h, str(path).encode(), int(first), int(count)))

    def __len__(self):
        n = C.c_int64()
        check(_lib.load().orbm_db_size(self._h, C.byref(n)))
        return n.value

    def top2(self, queries) -> np.ndarray:
        """Brute-force scan (ORBmatcher.cc:417-426 rule): structured array (best1, best2, idx)."""
        q = _desc(queries)
        out = np.zeros(len(q), TOP2_DTYPE)
        check(_lib.load().orbm_top2_search(self._h, ptr(q), len(q), ptr(out), 0))
        return out

    def top2_device(self, q_ptr: int, nq: int, out_ptr: int):
        """Asynchronous scan with device-resident queries/results (on the shard's own stream)."""
        check(_lib.load().orbm_top2_search(self._h, C.c_void_p(q_ptr), nq, C.c_void_p(out_ptr), 1))

    def sync(self):
        check(_lib.load().orbm_db_sync(self._h))

    def cuda_stream(self) -> int:
        s = C.c_void_p()
        check(_lib.load().orbm_db_get_stream(self._h, C.byref(s)))
        return s.value or 0

    def last_ms(self):
        ms, n = C.c_float(), C.c_int()
        check(_lib.load().orbm_db_last_ms(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def close(self):
        if self._h:
            _lib.load().orbm_db_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def merge_top2(shard_results, ratio: float = 0.75, th: int = TH_LOW):
    """Exact merge of per-shard top-2 results (shards ordered by row_base) + acceptance rule
    (ORBmatcher.cc:429-431).  Returns (merged structured array, accept flags)."""
    arr = np.ascontiguousarray(np.stack([np.ascontiguousarray(s, TOP2_DTYPE) for s in shard_results]))
    n_shards, nq = arr.shape
    out = np.zeros(nq, TOP2_DTYPE)
    acc = np.zeros(nq, np.uint8)
    check(_lib.load().orbm_merge_top2(ptr(arr), n_shards, nq, float(ratio), int(th), ptr(out), ptr(acc)))
    return out, acc.astype(bool)


def popc_peak(device: int = 0) -> float:
    v = C.c_double()
    check(_lib.load().orbm_popc_peak(device, C.byref(v)))
    return v.value


class ShardGroup:
    """The sharded search behind the C ABI (orbm_group_*): scan -> ncclAllGather -> device merge + acceptance ->
    one small transfer, all inside liborbfront.so.  Build with `from_torch` (one rank per process; the NCCL unique id
    travels over the existing torch.distributed group) or `local` (one process, one shard per device)."""

    def __init__(self, handle, dbs):
        self._h, self._dbs = handle, dbs  # keeps the shards alive

    @staticmethod
    def _nccl_first():
        """liborbfront.so binds NCCL with dlopen("libnccl.so.2"); PyTorch links its own bundled libnccl.so.2.  Whichever is
        loaded first serves both (same SONAME), and PyTorch cannot run on an older system NCCL: load PyTorch's first."""
        try:
            import torch  # noqa: F401
        except Exception:
            pass

    @staticmethod
    def unique_id() -> np.ndarray:
        ShardGroup._nccl_first()
        uid = np.zeros(128, np.uint8)
        check(_lib.load().orbm_group_unique_id(ptr(uid)))
        return uid

    @classmethod
    def create(cls, db: "DescriptorDB", rank: int, world: int, uid: np.ndarray) -> "ShardGroup":
        cls._nccl_first()
        h = C.c_void_p()
        uid = np.ascontiguousarray(uid, np.uint8)
        check(_lib.load().orbm_group_create(db._h, rank, world, ptr(uid), C.byref(h)))
        return cls(h, [db])

    @classmethod
    def from_torch(cls, db: "DescriptorDB", group=None) -> "ShardGroup":
        import torch
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        dev = f"cuda:{db.device}" if dist.get_backend(group) == "nccl" else "cpu"
        t = torch.from_numpy(cls.unique_id() if rank == 0 else np.zeros(128, np.uint8)).to(dev)
        dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        return cls.create(db, rank, world, t.cpu().numpy())

    @classmethod
    def local(cls, dbs) -> "ShardGroup":
        cls._nccl_first()
        h = C.c_void_p()
        arr = (C.c_void_p * len(dbs))(*[d._h for d in dbs])
        check(_lib.load().orbm_group_create_local(arr, len(dbs), C.byref(h)))
        return cls(h, list(dbs))

    def search(self, queries, ratio: float = 0.75, th: int = TH_LOW):
        q = _desc(queries)
        out = np.zeros(len(q), TOP2_DTYPE)
        acc = np.zeros(len(q), np.uint8)
        check(_lib.load().orbm_group_search(self._h, ptr(q), len(q), float(ratio), int(th), ptr(out), ptr(acc)))
        return out, acc.astype(bool)

    def last_ms(self):
        a, b = C.c_float(), C.c_float()
        check(_lib.load().orbm_group_last_ms(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def close(self):
        if self._h:
            _lib.load().orbm_group_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

