"""Host-side mirror of ORB_SLAM2::ORBextractor (include/ORBextractor.h:46-112) over the C ABI.

Same constructor arguments, same call (image, mask) -> (keypoints, descriptors), same getters,
and `mvImagePyramid` as a readable member.  Keypoints come back as a structured array with
cv::KeyPoint's field layout.  All work happens in liborbfront.so on the GPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KEYPOINT_DTYPE, OrbxConfig, check, ptr


class ORBextractor:
    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures: int, scaleFactor: float, nlevels: int, iniThFAST: int, minThFAST: int,
                 device: int = 0, max_width: int = 0, max_height: int = 0, max_batch: int = 1):
        self._h = C.c_void_p()
        self.nfeatures, self.nlevels = int(nfeatures), int(nlevels)
        self.scaleFactor = float(np.float32(scaleFactor))
        self.iniThFAST, self.minThFAST = int(iniThFAST), int(minThFAST)
        cfg = OrbxConfig(self.nfeatures, scaleFactor, self.nlevels, self.iniThFAST, self.minThFAST, device,
                         max_width, max_height, max_batch)
        check(_lib.load().orbx_create(C.byref(cfg), C.byref(self._h)))
        self._L = _lib.load()
        self._n_frames = 0

    # ---- reference getters (ORBextractor.h:64-84)
    def GetLevels(self) -> int:
        return self.nlevels

    def GetScaleFactor(self) -> float:
        return self.scaleFactor

    def _table(self, which):
        out = np.zeros(self.nlevels, np.float32)
        check(self._L.orbx_get_scale_table(self._h, which, ptr(out)))
        return out

    def GetScaleFactors(self):
        return self._table(0)

    def GetInverseScaleFactors(self):
        return self._table(1)

    def GetScaleSigmaSquares(self):
        return self._table(2)

    def GetInverseScaleSigmaSquares(self):
        return self._table(3)

    @property
    def mnFeaturesPerLevel(self):
        out = np.zeros(self.nlevels, np.int32)
        check(self._L.orbx_get_level_quota(self._h, ptr(out)))
        return out

    @property
    def out_cap(self) -> int:
        """Upper bound of keypoints per frame (sum over levels of quota + 4*roots + 3)."""
        q = self.mnFeaturesPerLevel
        return int(q.sum()) + 67 * self.nlevels

    # ---- operator()
    def __call__(self, image, mask=None):
        """operator()(image, mask, keypoints, descriptors) -- ORBextractor.cc:1051-1113.
        mask is ignored, as in the reference (ORBextractor.h:59)."""
        if image is None or image.size == 0:
            return np.zeros(0, KEYPOINT_DTYPE), np.zeros((0, 32), np.uint8)
        if image.dtype != np.uint8 or image.ndim != 2:
            raise AssertionError("image.type() == CV_8UC1")  # ORBextractor.cc:1058
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        cap = self.out_cap
        kps = np.zeros(cap, KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        check(self._L.orbx_extract(self._h, ptr(image), image.shape[1], image.shape[0], image.strides[0],
                                   ptr(kps), ptr(desc), cap, C.byref(n)))
        self._n_frames = 1
        return kps[:n.value].copy(), desc[:n.value].copy()

    # ---- batched form (frames are independent)
    def extract_batch(self, frames: np.ndarray, read: bool = True):
        """frames: (n,h,w) u8 host array.  Returns list of (keypoints, descriptors) per frame."""
        frames = np.ascontiguousarray(frames, np.uint8)
        n, h, w = frames.shape
        check(self._L.orbx_extract_batch(self._h, ptr(frames), n, w, h, frames.strides[1], frames.strides[0], 0))
        self._n_frames = n
        return self.read_batch() if read else None

    def extract_batch_device(self, dev_ptr: int, n: int, h: int, w: int, stride: int, frame_stride: int):
        """Frames already resident in device memory (asynchronous; results stay on the device)."""
        check(self._L.orbx_extract_batch(self._h, C.c_void_p(dev_ptr), n, w, h, stride, frame_stride, 1))
        self._n_frames = n

    def extract_batch_host_ptr(self, host_ptr: int, n: int, h: int, w: int, stride: int, frame_stride: int):
        """Frames in (pinned) host memory given by address; H2D copy inside the call (asynchronous)."""
        check(self._L.orbx_extract_batch(self._h, C.c_void_p(host_ptr), n, w, h, stride, frame_stride, 0))
        self._n_frames = n

    def frame_capacity(self, w: int, h: int) -> int:
        """Exact per-frame output rows needed for w x h frames (<= out_cap)."""
        cap = C.c_int(0)
        check(self._L.orbx_frame_capacity(self._h, w, h, C.byref(cap)))
        return cap.value

    def extract_batch_host(self, frames: np.ndarray, chunk_frames: int = 0):
        """Host batch through the pipelined upload | extract | download path; returns the list of
        (keypoints, descriptors) per frame like extract_batch."""
        frames = np.ascontiguousarray(frames, np.uint8)
        n, h, w = frames.shape
        cap = self.frame_capacity(w, h)
        kps = np.zeros((n, cap), KEYPOINT_DTYPE)
        desc = np.zeros((n, cap, 32), np.uint8)
        counts = np.zeros(n, np.int32)
        self.extract_batch_host_ptr_into(frames.ctypes.data, n, h, w, frames.strides[1], frames.strides[0],
                                         kps.ctypes.data, desc.ctypes.data, cap, counts, chunk_frames)
        return [(kps[f, :counts[f]].copy(), desc[f, :counts[f]].copy()) for f in range(n)]

    def extract_batch_host_ptr_into(self, host_ptr: int, n: int, h: int, w: int, stride: int, frame_stride: int,
                                    kps_ptr: int, desc_ptr: int, cap: int, counts: np.ndarray, chunk_frames: int = 0):
        """orbx_extract_batch_host on raw (ideally pinned) host addresses; blocks until results are on the host."""
        check(self._L.orbx_extract_batch_host(self._h, C.c_void_p(host_ptr), n, w, h, stride, frame_stride,
                                              C.c_void_p(kps_ptr), C.c_void_p(desc_ptr), cap, ptr(counts), chunk_frames))
        self._n_frames = n

    def counts(self) -> np.ndarray:
        out = np.zeros(self._n_frames, np.int32)
        if self._n_frames:
            check(self._L.orbx_batch_counts(self._h, ptr(out)))
        return out

    def read_batch(self):
        n = self._n_frames
        cap = self.out_cap
        kps = np.zeros((n, cap), KEYPOINT_DTYPE)
        desc = np.zeros((n, cap, 32), np.uint8)
        counts = np.zeros(n, np.int32)
        if n:
            check(self._L.orbx_batch_read_all(self._h, ptr(kps), ptr(desc), cap, ptr(counts)))
        return [(kps[f, :counts[f]].copy(), desc[f, :counts[f]].copy()) for f in range(n)]

    def read_batch_begin(self, kps_ptr: int, desc_ptr: int, cap: int, counts_ptr: int):
        """orbx_batch_read_begin: queue the download of every frame's results (page-locked destinations) and return."""
        check(self._L.orbx_batch_read_begin(self._h, C.c_void_p(kps_ptr), C.c_void_p(desc_ptr), cap, C.c_void_p(counts_ptr)))

    def read_batch_end(self):
        check(self._L.orbx_batch_read_end(self._h))

    def read_batch_into(self, kps_ptr: int, desc_ptr: int, cap: int, counts: np.ndarray):
        check(self._L.orbx_batch_read_all(self._h, C.c_void_p(kps_ptr), C.c_void_p(desc_ptr), cap, ptr(counts)))

    def sync(self):
        check(self._L.orbx_sync(self._h))

    STAGES = ("pyramid", "fast", "quadtree", "blur", "describe")

    def cuda_stream(self) -> int:
        """Address of the handle's cudaStream_t (wrap with torch.cuda.ExternalStream to time on it)."""
        s = C.c_void_p()
        check(self._L.orbx_get_stream(self._h, C.byref(s)))
        return s.value or 0

    def set_stage_timing(self, enabled: bool):
        check(self._L.orbx_set_stage_timing(self._h, int(enabled)))

    def last_stage_ms(self) -> dict:
        ms = np.zeros(5, np.float32)
        check(self._L.orbx_last_stage_ms(self._h, ptr(ms)))
        return dict(zip(self.STAGES, ms.tolist()))

    def last_batch_ms(self):
        ms, n = C.c_float(), C.c_int()
        check(self._L.orbx_last_batch_ms(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # ---- pyramid access (public member mvImagePyramid, read by Frame.cc:534,624-641)
    def level(self, l: int, frame: int = 0, with_border: bool = False) -> np.ndarray:
        w, h = C.c_int(), C.c_int()
        check(self._L.orbx_level_size(self._h, l, C.byref(w), C.byref(h)))
        b = 19 if with_border else 0
        out = np.zeros((h.value + 2 * b, w.value + 2 * b), np.uint8)
        check(self._L.orbx_get_level(self._h, frame, l, ptr(out), out.strides[0], int(with_border)))
        return out

    def blur_level(self, l: int, frame: int = 0) -> np.ndarray:
        """7x7 sigma-2 Gaussian of level l as the descriptor stage sees it (ORBextractor.cc:1093-1094)."""
        w, h = C.c_int(), C.c_int()
        check(self._L.orbx_level_size(self._h, l, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), np.uint8)
        check(self._L.orbx_get_blur_level(self._h, frame, l, ptr(out), out.strides[0]))
        return out

    @property
    def mvImagePyramid(self):
        """Level images of the last frame as views into their 19-px-bordered buffers, like the
        reference's cv::Mat ROIs."""
        return [self.level(l, 0, True)[19:-19, 19:-19] for l in range(self.nlevels)]

    def candidates(self, l: int, frame: int = 0) -> np.ndarray:
        """FAST candidates (x, y, response) of a level, border-relative, unspecified order."""
        cap = 1 << 22
        n = C.c_int()
        buf = np.zeros((cap, 3), np.int32)
        check(self._L.orbx_get_candidates(self._h, frame, l, ptr(buf), cap, C.byref(n)))
        return buf[:n.value].copy()

    def close(self):
        if self._h:
            self._L.orbx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def stereo_matches(left: "ORBextractor", right: "ORBextractor", keys_left, desc_left, keys_right, desc_right, mb: float,
                   mbf: float, frame_left: int = 0, frame_right: int = 0):
    """Frame::ComputeStereoMatches, whole (Frame.cc:527-700), on the pyramids the two extractors hold on
    the device: returns (mvuRight, mvDepth), -1 where unmatched.  `left` and `right` may be the same
    extractor holding both images as one batch (frame_left / frame_right)."""
    kl = np.ascontiguousarray(keys_left, _lib.KEYPOINT_DTYPE)
    kr = np.ascontiguousarray(keys_right, _lib.KEYPOINT_DTYPE)
    dl = np.ascontiguousarray(desc_left, np.uint8).reshape(-1, 32)
    dr = np.ascontiguousarray(desc_right, np.uint8).reshape(-1, 32)
    ur = np.zeros(len(kl), np.float32)
    z = np.zeros(len(kl), np.float32)
    check(_lib.load().orbx_stereo_matches(left._h, frame_left, right._h, frame_right, ptr(kl), ptr(dl), len(kl), ptr(kr), ptr(dr),
                                          len(kr), float(mb), float(mbf), ptr(ur), ptr(z)))
    return ur, z


def stereo_matches_batch(ext: "ORBextractor", n_pairs: int, mb: float, mbf: float):
    """Frame::ComputeStereoMatches for pairs (2p, 2p+1) of `ext`'s last batch, read in place on the device
    (orbx_stereo_matches_batch): returns (mvuRight, mvDepth) as (n_pairs, cap) float arrays, row i = left keypoint i."""
    cap = ext.out_cap
    ur = np.zeros((n_pairs, cap), np.float32)
    z = np.zeros((n_pairs, cap), np.float32)
    check(_lib.load().orbx_stereo_matches_batch(ext._h, n_pairs, float(mb), float(mbf), ptr(ur), ptr(z), cap))
    return ur, z
