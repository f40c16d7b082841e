"""Extended parity fuzz, run by hand on a GPU box (not part of the test suite: minutes of oracle time):

  python tools/fuzz_extended.py single N     N more random single-frame cases of tests/test_gpu_fuzz.py (seeds 36 ..) + 60 full-size frames
  python tools/fuzz_extended.py batch N      N more random batches large enough to select every stage's batch kernels
  python tools/fuzz_extended.py host N       N random batch sizes through the upload | extract | download pipeline == the resident path

Run each also with ORBX_QP_RATIO4=0 (the sort-free quadtree kernel tries every level: midway hand-backs) and
ORBX_QT_NO_PYRAMID=1 (sorting kernel only).  Round 2: 400 + 110 + 60 cases in all modes; two bugs found and fixed
(profiles/r02_history.md section 8)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

from orb_slam2_e_b200 import synth  # noqa: E402
from orb_slam2_e_b200.extractor import ORBextractor  # noqa: E402


def single(n):
    import test_gpu_extract as E
    import test_gpu_fuzz as F
    bad = 0
    for seed in range(36, 36 + n):
        try:
            F.test_random_case(seed)
        except AssertionError as e:
            bad += 1
            print("FAIL seed", seed, str(e)[:200])
    for seed in range(20):
        for frame, cfg in ((synth.textured_frame(480, 640, 5000 + seed), E.TUM), (synth.endoscopy_frame(480, 640, 5000 + seed), E.HAMLYN),
                           (synth.stereo_pair(376, 1241, 6000 + seed)[0], E.KITTI)):
            try:
                E.compare(frame, cfg)
            except AssertionError as e:
                bad += 1
                print("FAIL full-size", seed, cfg, str(e)[:200])
    return bad


def batch(n):
    import test_gpu_fuzz as F
    bad = 0
    for seed in range(6, 6 + n):
        try:
            F.test_random_batch_takes_the_batch_kernels(seed)
        except AssertionError as e:
            bad += 1
            print("FAIL batch seed", seed, str(e)[:200])
    return bad


def host(n):
    rng = np.random.default_rng(1)
    bad = 0
    for it in range(n):
        h, w = int(rng.integers(100, 200)), int(rng.integers(120, 260))
        nfr = int(rng.choice([1, 2, 7, 31, 32, 33, 127, 128, 129, 200, 255, 256, 257, 300, 511, 512, 513, 640, 777, 1000]))
        base = [synth.textured_frame(h, w, 9000 + it * 7 + i) for i in range(5)]
        frames = np.ascontiguousarray(np.stack([np.roll(base[i % 5], (i // 5) * 3, axis=1) for i in range(nfr)]))
        cfg = (int(rng.integers(50, 600)), 1.2, 2, 20, 7)
        ra = ORBextractor(*cfg, max_width=w, max_height=h, max_batch=nfr).extract_batch(frames)
        rb = ORBextractor(*cfg, max_width=w, max_height=h, max_batch=nfr).extract_batch_host(frames, int(rng.choice([0, 0, 0, 5, 64])))
        if not (len(ra) == len(rb) and all(np.array_equal(x[0], y[0]) and np.array_equal(x[1], y[1]) for x, y in zip(ra, rb))):
            bad += 1
            print("MISMATCH", it, h, w, nfr, cfg)
    return bad


if __name__ == "__main__":
    mode, n = sys.argv[1], int(sys.argv[2])
    print("extended fuzz (%s, %d): %d failures" % (mode, n, {"single": single, "batch": batch, "host": host}[mode](n)))
