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


def test_read_back_beside_the_stereo_kernels_equals_blocking_read():
    """orbx_batch_read_begin / _end: the download of keypoints and descriptors queued before orbx_stereo_matches_batch
    and finished after it returns the same bytes as the blocking orbx_batch_read_all, and the stereo results are
    unchanged; the next extract call waits for a read that was never ended."""
    import torch
    from orb_slam2_e_b200 import _lib
    h, w, n_pairs = 240, 320, 4
    cfg = (500, 1.2, 4, 20, 7)
    frames = np.empty((2 * n_pairs, h, w), np.uint8)
    for p in range(n_pairs):
        frames[2 * p], frames[2 * p + 1] = synth.stereo_pair(h, w, 950 + p)
    ext = ORBextractor(*cfg)
    n = 2 * n_pairs
    mbf, mb = 386.1448, 386.1448 / 718.856
    ext.extract_batch(frames, read=False)
    cap = ext.frame_capacity(w, h)
    want_kp = np.zeros((n, cap), _lib.KEYPOINT_DTYPE)
    want_de = np.zeros((n, cap, 32), np.uint8)
    want_cnt = np.zeros(n, np.int32)
    ext.read_batch_into(want_kp.ctypes.data, want_de.ctypes.data, cap, want_cnt)
    want_ur, want_z = stereo_matches_batch(ext, n_pairs, mb, mbf)

    pin_kp = torch.zeros(n * cap * 28, dtype=torch.uint8).pin_memory()
    pin_de = torch.zeros(n * cap * 32, dtype=torch.uint8).pin_memory()
    pin_cnt = torch.zeros(n, dtype=torch.int32).pin_memory()
    ext.extract_batch(frames, read=False)
    ext.read_batch_begin(pin_kp.data_ptr(), pin_de.data_ptr(), cap, pin_cnt.data_ptr())
    ur, z = stereo_matches_batch(ext, n_pairs, mb, mbf)
    ext.read_batch_end()
    assert np.array_equal(pin_cnt.numpy(), want_cnt)
    got_kp = pin_kp.numpy().view(_lib.KEYPOINT_DTYPE).reshape(n, cap)
    got_de = pin_de.numpy().reshape(n, cap, 32)
    for f in range(n):
        c = int(want_cnt[f])
        assert np.array_equal(got_kp[f, :c], want_kp[f, :c]) and np.array_equal(got_de[f, :c], want_de[f, :c])
    assert np.array_equal(ur, want_ur) and np.array_equal(z, want_z)
    # a read that is never ended: the next extraction must not overwrite the results under the copy
    ext.read_batch_begin(pin_kp.data_ptr(), pin_de.data_ptr(), cap, pin_cnt.data_ptr())
    ext.extract_batch(frames[:2], read=False)
    ext.read_batch_end()
    assert np.array_equal(pin_cnt.numpy(), want_cnt)
