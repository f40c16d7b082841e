"""The blur stage itself (ORBextractor.cc:1093-1094): every blurred level of the CUDA path, pixel for pixel, against the
oracle's cv::GaussianBlur restatement (orc_blur7, pinned to cv2 4.13 in test_golden.py / test_oracle_prims.py) applied
to the CUDA path's own pyramid level -- including levels whose width is not a multiple of 4 or 128 (edge lanes of the
BORDER_REFLECT_101 PRMT selectors), tiles cut by the level's bottom, and both tile heights (batch / few frames).
Also: a rejected frame size must not leave the handle in a state where the next good frame is processed with a
half-built geometry (round-1 advisor finding)."""
import numpy as np
import pytest

import oracle_lib
from orb_slam2_e_b200 import synth
from orb_slam2_e_b200._lib import OrbFrontError
from orb_slam2_e_b200.extractor import ORBextractor

pytestmark = pytest.mark.gpu


def _blur_ref(img):
    L = oracle_lib.oracle()
    img = np.ascontiguousarray(img)
    dst = np.zeros_like(img)
    L.orc_blur7(img, img.shape[1], img.shape[0], img.strides[0], dst, dst.strides[0])
    return dst


@pytest.mark.parametrize("shape,cfg,batch", [
    ((480, 640), (1000, 1.2, 8, 20, 7), 1),
    ((480, 640), (1000, 1.2, 8, 20, 7), 40),     # 56-row tiles (batch path)
    ((376, 1241), (2000, 1.2, 8, 20, 7), 2),     # odd width: partial last word, right-edge reflection inside a lane
    ((131, 197), (300, 1.3, 3, 20, 7), 1),       # small levels, widths 197 / 152 / 117
    ((101, 129), (200, 1.1, 2, 20, 7), 3),       # one pixel past a 128-column tile: a tile that is 1 px wide
])
def test_blurred_levels_equal_oracle(shape, cfg, batch):
    h, w = shape
    frames = np.stack([synth.textured_frame(h, w, 4200 + i) for i in range(min(batch, 3))])
    frames = np.ascontiguousarray(np.tile(frames, ((batch + len(frames) - 1) // len(frames), 1, 1))[:batch])
    ext = ORBextractor(*cfg)
    ext.extract_batch(frames, read=False)
    for f in sorted({0, batch - 1}):
        for l in range(cfg[2]):
            level = ext.level(l, f)
            got = ext.blur_level(l, f)
            want = _blur_ref(level)
            assert got.shape == want.shape
            bad = np.argwhere(got != want)
            assert len(bad) == 0, (f, l, level.shape, bad[:5])


def test_rejected_size_does_not_poison_the_handle():
    cfg = (1000, 1.2, 8, 20, 7)
    img = synth.textured_frame(480, 640, 77)
    ext = ORBextractor(*cfg)
    k0, d0 = ext(img)
    with pytest.raises(OrbFrontError):
        ext(synth.textured_frame(40, 40, 1))  # level 0 is too small for the 30-px cell grid
    k1, d1 = ext(img)
    assert np.array_equal(k0, k1) and np.array_equal(d0, d1)
    pk, pd = oracle_lib.PortExtractor(*cfg).extract(img)
    for name in ("x", "y", "octave", "response", "size"):
        assert np.array_equal(k1[name], pk[name]), name
    assert np.array_equal(d1, pd)


def test_limits_given_at_creation_are_enforced():
    ext = ORBextractor(500, 1.2, 4, 20, 7, max_width=320, max_height=240, max_batch=2)
    ext(synth.textured_frame(240, 320, 3))
    with pytest.raises(OrbFrontError):
        ext(synth.textured_frame(480, 640, 3))
    with pytest.raises(OrbFrontError):
        ext.extract_batch(np.stack([synth.textured_frame(240, 320, 3)] * 3))


@pytest.mark.parametrize("shape,cfg,batch", [
    ((480, 640), (1000, 1.2, 8, 20, 7), 64),
    ((480, 640), (1200, 1.1, 6, 24, 7), 64),
    ((376, 1241), (2000, 1.2, 8, 20, 7), 48),   # odd width and pitch
    ((240, 322), (500, 1.2, 8, 20, 7), 200),     # small levels, many frames
])
def test_batch_pyramid_equals_single_frame_pyramid(shape, cfg, batch):
    """Batches large enough to fill the GPU take the shared-memory-fed resize kernel (pyramid.cu
    resize_rows_tma_kernel); one frame takes the register-fed one, which test_gpu_extract.py compares with the
    oracle level by level.  Both must give the same pixels."""
    h, w = shape
    distinct = [synth.textured_frame(h, w, 4300 + i) for i in range(3)]
    frames = np.ascontiguousarray(np.stack([distinct[i % 3] for i in range(batch)]))
    big = ORBextractor(*cfg, max_width=w, max_height=h, max_batch=batch)
    big.extract_batch(frames, read=False)
    one = ORBextractor(*cfg)
    for f in (0, batch // 2 + 1, batch - 1):
        one(frames[f])
        for l in range(cfg[2]):
            assert np.array_equal(big.level(l, f), one.level(l, 0)), (f, l)
            assert np.array_equal(big.blur_level(l, f), one.blur_level(l, 0)), (f, l)


def test_pyramid_staging_survives_a_host_batch_on_the_same_handle():
    """The page-locked pyramid staging of the single-frame drop-in path (orbx_set_pyramid_staging /
    orbx_get_pyramid_host) and its side stream belong to the handle until orbx_destroy: a host-buffer batch in
    between (which grows other page-locked buffers) must leave them usable."""
    import ctypes as C
    from orb_slam2_e_b200 import _lib
    L = _lib.load()
    cfg = (500, 1.2, 4, 20, 7)
    h, w = 240, 320
    ext = ORBextractor(*cfg, max_width=w, max_height=h, max_batch=40)
    _lib.check(L.orbx_set_pyramid_staging(ext._h, 1))

    def staged_level(level):
        p, lw, lh, st = C.c_void_p(), C.c_int(), C.c_int(), C.c_size_t()
        _lib.check(L.orbx_get_pyramid_host(ext._h, level, C.byref(p), C.byref(lw), C.byref(lh), C.byref(st)))
        rows = (C.c_uint8 * (st.value * lh.value)).from_address(p.value)
        return np.frombuffer(rows, np.uint8).reshape(lh.value, st.value)[:, :lw.value].copy()

    a, b = synth.textured_frame(h, w, 4400), synth.textured_frame(h, w, 4401)
    ext(a)
    for l in range(cfg[2]):
        assert np.array_equal(staged_level(l), ext.level(l, 0))
    batch = np.ascontiguousarray(np.stack([synth.textured_frame(h, w, 4410 + i % 3) for i in range(40)]))
    ext.extract_batch_host(batch)     # upload | extract | download pipeline: grows its own page-locked buffers
    ext(b)
    for l in range(cfg[2]):
        assert np.array_equal(staged_level(l), ext.level(l, 0))
    ext(a)
    assert np.array_equal(staged_level(0), a)
