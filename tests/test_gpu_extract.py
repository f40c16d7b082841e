"""GPU parity: the CUDA extractor (through the C ABI) vs the CPU oracle on the same frames.

Bar (BASELINE.json north_star): pyramid pixels, candidate sets, kept keypoints (octave, x, y,
response, size) and their ORDER bit-exact; angles within 1e-3 rad (in practice identical floats);
descriptors bit-exact up to <= 1e-4 mismatching bits (rounding-boundary samples)."""
import numpy as np
import pytest

from oracle_lib import PortExtractor
from orb_slam2_e_b200 import synth
from orb_slam2_e_b200.extractor import ORBextractor

pytestmark = pytest.mark.gpu

TUM = (1000, 1.2, 8, 20, 7)
KITTI = (2000, 1.2, 8, 20, 7)
HAMLYN = (1200, 1.1, 6, 24, 7)


def compare(img, cfg, gpu=None, frame_result=None):
    port = PortExtractor(*cfg)
    pk, pd = port.extract(img)
    if frame_result is None:
        gpu = gpu or ORBextractor(*cfg)
        gk, gd = gpu(img)
        for l in range(cfg[2]):
            assert np.array_equal(gpu.level(l, 0, False), port.level(l, False)), f"pyramid level {l}"
            assert np.array_equal(gpu.level(l, 0, True), port.level(l, True)), f"bordered level {l}"
            got = gpu.candidates(l)
            want = port.stage(0, l)
            key = lambda a: a[np.lexsort((a[:, 0], a[:, 1]))]
            assert len(got) == len(want), f"level {l}: {len(got)} candidates vs {len(want)}"
            assert np.array_equal(key(got), key(want)), f"candidates level {l}"
    else:
        gk, gd = frame_result
    assert len(gk) == len(pk), (len(gk), len(pk))
    for name in ("octave", "x", "y", "response", "size", "class_id"):
        assert np.array_equal(gk[name], pk[name]), name
    dang = np.abs(np.deg2rad(gk["angle"].astype(np.float64) - pk["angle"].astype(np.float64)))
    dang = np.minimum(dang, 2 * np.pi - dang)
    assert dang.max(initial=0) <= 1e-3
    same_angle = gk["angle"] == pk["angle"]
    bits = np.unpackbits(gd ^ pd, axis=1)
    rate = bits.sum() / max(bits.size, 1)
    assert rate <= 1e-4, f"descriptor bit mismatch rate {rate}"
    return dict(n=len(pk), angle_equal=float(same_angle.mean()) if len(pk) else 1.0, bit_rate=rate)


def test_getters_match_oracle():
    g = ORBextractor(*TUM)
    p = PortExtractor(*TUM)
    assert [p.quota(l) for l in range(8)] == list(g.mnFeaturesPerLevel)
    assert [p.scale(l) for l in range(8)] == list(g.GetScaleFactors())
    assert g.GetLevels() == 8 and abs(g.GetScaleFactor() - 1.2) < 1e-6
    s = g.GetScaleFactors()
    assert np.array_equal(g.GetInverseScaleFactors(), (np.float32(1) / s).astype(np.float32))
    assert np.array_equal(g.GetScaleSigmaSquares(), s * s)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_tum_textured(seed):
    r = compare(synth.textured_frame(480, 640, seed), TUM)
    assert r["n"] >= 1000


@pytest.mark.parametrize("seed", [11, 12])
def test_hamlyn_low_texture(seed):
    compare(synth.endoscopy_frame(480, 640, seed), HAMLYN)


def test_kitti_stereo_shapes():
    left, right = synth.stereo_pair(376, 1241, 21)
    g = ORBextractor(*KITTI)
    compare(left, KITTI, g)
    compare(right, KITTI, g)


def test_edge_cases():
    g = ORBextractor(500, 1.2, 3, 20, 7)
    k, d = g(np.zeros((0, 0), np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
    k, d = g(np.full((200, 260), 90, np.uint8))  # flat: no corners anywhere
    assert len(k) == 0
    with pytest.raises(AssertionError):
        g(np.zeros((100, 100, 3), np.uint8))
    # strided (non-contiguous rows) input, odd sizes, small quota
    big = synth.textured_frame(300, 420, 9)
    compare(big[7:251, 11:390], (300, 1.2, 4, 20, 7))
    compare(synth.textured_frame(123, 157, 5), (150, 1.3, 3, 15, 5))
    # a level too small for the 30-px cell grid is rejected loudly
    from orb_slam2_e_b200._lib import OrbFrontError
    with pytest.raises(OrbFrontError):
        ORBextractor(500, 1.2, 8, 20, 7)(synth.textured_frame(100, 120, 1))


def test_frame_size_change_and_reuse():
    g = ORBextractor(*TUM)
    compare(synth.textured_frame(480, 640, 4), TUM, g)
    compare(synth.textured_frame(360, 480, 5), TUM, g)
    compare(synth.textured_frame(480, 640, 6), TUM, g)


def test_batch_equals_single_frames():
    frames = np.stack([synth.textured_frame(480, 640, 20), synth.endoscopy_frame(480, 640, 21),
                       synth.textured_frame(480, 640, 22), np.full((480, 640), 7, np.uint8),
                       synth.endoscopy_frame(480, 640, 23)])
    g = ORBextractor(*HAMLYN, max_width=640, max_height=480, max_batch=5)
    res = g.extract_batch(frames)
    assert len(res) == 5 and len(res[3][0]) == 0
    for f in range(5):
        compare(frames[f], HAMLYN, frame_result=res[f])


def test_large_frame_global_sort_path():
    """1280x720 with a small smem key budget forces the global-memory quadtree path."""
    import os
    os.environ["ORBX_QT_SMEM_KEYS"] = "1024"
    try:
        compare(synth.textured_frame(720, 1280, 31), (3000, 1.2, 8, 20, 7))
    finally:
        del os.environ["ORBX_QT_SMEM_KEYS"]


def test_config5_4k_frame():
    """BASELINE config 5: 3840x2160, nFeatures=8000, 12 levels (large-image stress of pyramid + quadtree)."""
    cfg = (8000, 1.2, 12, 20, 7)
    img = synth.textured_frame(2160, 3840, 51)
    port = PortExtractor(*cfg)
    port.cap = 9000
    pk, pd = port.extract(img)
    g = ORBextractor(*cfg)
    gk, gd = g(img)
    assert len(pk) >= 8000
    compare(img, cfg, frame_result=(gk, gd))
    for l in (0, 5, 11):
        assert np.array_equal(g.level(l, 0, False), port.level(l, False)), f"pyramid level {l}"
        got, want = g.candidates(l), port.stage(0, l)
        key = lambda a: a[np.lexsort((a[:, 0], a[:, 1]))]
        assert np.array_equal(key(got), key(want)), f"candidates level {l}"


def test_config2_stereo_extract_then_search():
    """BASELINE config 2: KITTI-shape pair, 2000 features per image, extraction of both images on the
    GPU followed by the descriptor search of Frame::ComputeStereoMatches on the GPU's own output."""
    import match_cases as mc
    from oracle_lib import oracle
    from orb_slam2_e_b200 import matcher
    left, right = synth.stereo_pair(376, 1241, 33)
    gl, gr = ORBextractor(*KITTI), ORBextractor(*KITTI)  # two extractor objects, as Frame.cc:78-81
    kl, dl = gl(left)
    kr, dr = gr(right)
    pl, pr = PortExtractor(*KITTI), PortExtractor(*KITTI)
    pl.cap = pr.cap = 2200
    okl, odl = pl.extract(left)
    okr, odr = pr.extract(right)
    assert kl.tobytes() == okl.tobytes() and kr.tobytes() == okr.tobytes()
    assert np.array_equal(dl, odl) and np.array_equal(dr, odr)
    scale = gl.GetScaleFactors()
    bi, bd = mc.oracle_stereo(oracle(), okl, odl, okr, odr, scale, 376, 0.0, 80.0)
    gi, gd = matcher.ORBmatcher().StereoSearch(kl, dl, kr, dr, scale, 376, 0.0, 80.0)
    assert np.array_equal(gd, bd) and np.array_equal(gi, bi)
    assert ((bd >= 0) & (bd < 70)).sum() > 200  # plenty of candidates pass to the SAD refinement


@pytest.mark.parametrize("one_handle", [False, True])
def test_config2_whole_stereo_matching_on_device(one_handle):
    """BASELINE config 2 end to end: both images extracted on the GPU, then Frame::ComputeStereoMatches
    (Frame.cc:527-700: row search, 11x11 sliding-window refinement, parabola fit, median rejection)
    on the pyramids still resident in HBM.  mvuRight / mvDepth must equal the CPU restatement run on
    the oracle's pyramids bit for bit (integer patch distances, identical float expressions)."""
    import match_cases as mc
    from oracle_lib import oracle
    from orb_slam2_e_b200 import extractor as ex
    left, right = synth.stereo_pair(376, 1241, 34)
    mbf, mb = 386.1448, 386.1448 / 718.856  # KITTI 00-02
    if one_handle:  # one extractor, the pair as a batch of two
        g = ORBextractor(*KITTI)
        (kl, dl), (kr, dr) = g.extract_batch(np.stack([left, right]))
        ur, z = ex.stereo_matches(g, g, kl, dl, kr, dr, mb, mbf, 0, 1)
    else:           # two extractor objects, as Frame.cc:78-81
        gl, gr = ORBextractor(*KITTI), ORBextractor(*KITTI)
        kl, dl = gl(left)
        kr, dr = gr(right)
        ur, z = ex.stereo_matches(gl, gr, kl, dl, kr, dr, mb, mbf)
        g = gl
    pl, pr = PortExtractor(*KITTI), PortExtractor(*KITTI)
    pl.cap = pr.cap = 2200
    okl, odl = pl.extract(left)
    okr, odr = pr.extract(right)
    assert kl.tobytes() == okl.tobytes() and kr.tobytes() == okr.tobytes()
    levels_l = [pl.level(l) for l in range(KITTI[2])]
    levels_r = [pr.level(l) for l in range(KITTI[2])]
    our, oz = mc.oracle_stereo_matches(oracle(), okl, odl, okr, odr, g.GetScaleFactors(), g.GetInverseScaleFactors(), levels_l,
                                       levels_r, mb, mbf)
    assert (our >= 0).sum() > 150, (our >= 0).sum()
    assert ur.tobytes() == our.tobytes() and z.tobytes() == oz.tobytes()
    # disparities of the synthetic pair are 0..60 px: depths follow
    d = kl["x"][our >= 0] - our[our >= 0]
    assert d.min() >= 0 and d.max() < 80


def test_pipelined_host_batch_equals_single_frames():
    """orbx_extract_batch_host (chunked upload | extract | download) == per-frame results, ragged last chunk."""
    frames = np.stack([synth.textured_frame(240, 320, 40 + i) if i % 3 else synth.endoscopy_frame(240, 320, 40 + i) for i in range(7)])
    cfg = (400, 1.2, 4, 20, 7)
    g = ORBextractor(*cfg)
    assert g.frame_capacity(320, 240) <= g.out_cap
    for chunk in (3, 64):
        res = g.extract_batch_host(frames, chunk_frames=chunk)
        assert len(res) == 7
        for f in range(7):
            compare(frames[f], cfg, frame_result=res[f])
    res2 = g.extract_batch(frames)  # the two-call path still agrees
    for f in range(7):
        assert res2[f][0].tobytes() == res[f][0].tobytes() and np.array_equal(res2[f][1], res[f][1])
