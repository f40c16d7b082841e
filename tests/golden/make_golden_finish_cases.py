"""Seeded inputs of the frame-finish fixtures (shared by the generator and the tests; no reference
needed).  Camera models are the reference's own settings files (values copied as test data):
Examples/RGB-D/TUM1.yaml:8-26, TUM2.yaml:8-26, TUM3.yaml (no distortion), Examples/Monocular/EuRoC.yaml:8-16."""
import numpy as np

# name -> (width, height, (fx, fy, cx, cy), (k1, k2, p1, p2[, k3]), bf)
CAMERAS = {
    "tum1": (640, 480, (517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314), 40.0),
    "tum2": (640, 480, (520.908620, 521.007327, 325.141442, 249.701764), (0.231222, -0.784899, -0.003257, -0.000105, 0.917205), 40.0),
    "tum3": (640, 480, (535.4, 539.2, 320.1, 247.6), (0.0, 0.0, 0.0, 0.0), 40.0),
    "euroc": (752, 480, (458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05), 47.90639384423901),
    # strong barrel distortion: corners leave the image by a wide margin, icdist changes sign far outside
    "wide": (640, 480, (300.0, 300.0, 320.0, 240.0), (-0.35, 0.12, 0.001, -0.002, -0.02), 30.0),
}


def keypoints(name, n, seed):
    """n keypoint positions as the extractor emits them: level-0 coordinates of scaled integer level
    positions inside the 16-px border, plus a few on the exact border and duplicates."""
    w, h = CAMERAS[name][:2]
    rng = np.random.default_rng(seed)
    lvl = rng.integers(0, 8, n)
    scale = np.float32(1.2) ** lvl.astype(np.float32)
    lw, lh = np.round(w / scale).astype(np.int64), np.round(h / scale).astype(np.int64)
    x = (rng.integers(16, np.maximum(lw - 16, 17)).astype(np.float32) * scale).astype(np.float32)
    y = (rng.integers(16, np.maximum(lh - 16, 17)).astype(np.float32) * scale).astype(np.float32)
    x = np.minimum(x, np.float32(w - 1))
    y = np.minimum(y, np.float32(h - 1))
    if n >= 8:
        x[:4] = [0, w - 1, 0, w - 1]
        y[:4] = [0, 0, h - 1, h - 1]
        x[4:8] = x[8:12] if n >= 12 else x[4:8]
        y[4:8] = y[8:12] if n >= 12 else y[4:8]
    return np.ascontiguousarray(x, np.float32), np.ascontiguousarray(y, np.float32)


def depth_image(name, seed):
    """CV_32F depth map with holes (0), negative junk and a smooth surface in between."""
    w, h = CAMERAS[name][:2]
    rng = np.random.default_rng(seed + 1000)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    d = (np.float32(1.5) + np.float32(0.002) * xx + np.float32(0.8) * np.sin(yy / np.float32(37.0))).astype(np.float32)
    d[rng.random((h, w)) < 0.15] = 0.0
    d[rng.random((h, w)) < 0.03] = -1.0
    return np.ascontiguousarray(d, np.float32)


# name -> (camera, n keypoints, seed, with depth)
FINISH_CASES = {
    "tum1_rgbd": ("tum1", 1000, 11, True),
    "tum2_rgbd": ("tum2", 1373, 12, True),
    "tum3_rgbd_nodist": ("tum3", 997, 13, True),
    "euroc_mono": ("euroc", 1200, 14, False),
    "wide_mono": ("wide", 2000, 15, False),
    "tum1_tiny": ("tum1", 3, 16, True),
    "euroc_dense": ("euroc", 8000, 17, False),
}


def random_points(name, n, seed):
    """Arbitrary float points in and around the image (cv2 pin of cv::undistortPoints)."""
    w, h = CAMERAS[name][:2]
    rng = np.random.default_rng(seed)
    p = np.stack([rng.uniform(-40, w + 40, n), rng.uniform(-40, h + 40, n)], 1).astype(np.float32)
    p[:4] = [[0, 0], [w, 0], [0, h], [w, h]]
    return np.ascontiguousarray(p)
