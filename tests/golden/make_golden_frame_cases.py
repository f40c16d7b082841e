"""Case table shared by make_golden_frame.py (generator) and tests/test_golden_frame.py."""
import numpy as np

from orb_slam2_e_b200 import synth

STEREO_CASES = {  # name -> ((h, w), seed, extractor config, max disparity, mb, mbf)
    "kitti": ((376, 1241), 34, (2000, 1.2, 8, 20, 7), 60.0, 386.1448 / 718.856, 386.1448),
    "small": ((240, 640), 12, (1200, 1.2, 8, 20, 7), 40.0, 0.08, 40.0),
    "coarse": ((480, 640), 9, (1000, 1.5, 5, 20, 7), 30.0, 0.12, 47.9),
}


def stereo_pair(shape, seed, max_disp):
    return tuple(np.ascontiguousarray(a) for a in synth.stereo_pair(shape[0], shape[1], seed, max_disp=max_disp))


def distinctive_case():
    """Observations of 400 map points (1..60 each, at least one): noisy copies of one descriptor per point, with exact
    duplicates (median ties)."""
    rng = np.random.default_rng(15)
    counts = np.concatenate([[1, 2, 2, 3, 4, 33, 64], rng.integers(1, 60, 393)]).astype(np.int32)
    offset = np.zeros(len(counts) + 1, np.int32)
    np.cumsum(counts, out=offset[1:])
    desc = np.zeros((int(offset[-1]), 32), np.uint8)
    for p in range(len(counts)):
        base = synth.random_descriptors(1, 500 + p)[0]
        for i in range(offset[p], offset[p + 1]):
            desc[i] = synth.flip_bits(base, int(rng.integers(0, 15)) * 4, rng)
        if counts[p] > 3:
            desc[offset[p] + 2] = desc[offset[p]]
    return desc, offset
