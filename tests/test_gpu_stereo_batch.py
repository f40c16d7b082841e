"""orbx_stereo_matches_batch (pairs of one handle's batch, everything on the device) == one orbx_stereo_matches call per
pair (which tests/test_gpu_frame_finish.py and test_golden_frame.py pin on the reference's own Frame.cc)."""
import numpy as np
import pytest

from orb_slam2_e_b200 import synth
from orb_slam2_e_b200.extractor import ORBextractor, stereo_matches, stereo_matches_batch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape,cfg,n_pairs", [((376, 1241), (2000, 1.2, 8, 20, 7), 3), ((240, 320), (500, 1.2, 4, 20, 7), 5)])
def test_batch_equals_per_pair_calls(shape, cfg, n_pairs):
    h, w = shape
    frames = np.empty((2 * n_pairs, h, w), np.uint8)
    for p in range(n_pairs):
        frames[2 * p], frames[2 * p + 1] = synth.stereo_pair(h, w, 900 + p)
    frames[2 * (n_pairs - 1) + 1] = 0  # a right image without keypoints: every left keypoint stays unmatched
    ext = ORBextractor(*cfg)
    res = ext.extract_batch(frames)
    mbf, mb = 386.1448, 386.1448 / 718.856
    ur_b, z_b = stereo_matches_batch(ext, n_pairs, mb, mbf)
    total = 0
    for p in range(n_pairs):
        (kl, dl), (kr, dr) = res[2 * p], res[2 * p + 1]
        ur, z = stereo_matches(ext, ext, kl, dl, kr, dr, mb, mbf, frame_left=2 * p, frame_right=2 * p + 1)
        n = len(kl)
        assert np.array_equal(ur_b[p, :n], ur) and np.array_equal(z_b[p, :n], z), p
        assert np.all(ur_b[p, n:] == -1) and np.all(z_b[p, n:] == -1)
        total += int((z > 0).sum())
    assert total > 50 * (n_pairs - 1)
    assert np.all(z_b[n_pairs - 1] == -1)
