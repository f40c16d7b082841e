"""Pins of the restatement on the reference's OWN src/Frame.cc and src/MapPoint.cc, compiled
unmodified into oracle/_ref/libframeref.so (oracle/ref_frame_harness.cc):
  Frame::ComputeStereoMatches            == matchport::compute_stereo_matches   (bit-equal floats)
  Frame::AssignFeaturesToGrid + GetFeaturesInArea == FrameGrid + matchport::features_in_area (same lists, same order)
  MapPoint::ComputeDistinctiveDescriptors == matchport::distinctive_descriptor
and of the constants the product header uses on ORBmatcher.cc's.  CPU only; needs /root/reference at
build time (skipped otherwise -- the GPU box checks the committed fixtures instead, test_golden.py)."""
import numpy as np
import pytest

import match_cases as mc
import oracle_lib
from oracle_lib import PortExtractor, oracle
from orb_slam2_e_b200 import synth

FR = oracle_lib.frame_ref_lib()
pytestmark = pytest.mark.skipif(FR is None, reason="oracle/_ref/libframeref.so not built (needs /root/reference)")


def test_constants_are_the_compiled_reference_ones():
    from orb_slam2_e_b200 import matcher
    c = np.zeros(4, np.int32)
    FR.frameref_constants(c)
    assert (int(c[0]), int(c[1])) == (matcher.TH_HIGH, matcher.TH_LOW)
    assert (int(c[2]), int(c[3])) == (64, 48)


def stereo_inputs(shape, seed, cfg, max_disp):
    left, right = (np.ascontiguousarray(a) for a in synth.stereo_pair(shape[0], shape[1], seed, max_disp=max_disp))
    pl, pr = PortExtractor(*cfg), PortExtractor(*cfg)
    kl, dl = pl.extract(left)
    kr, dr = pr.extract(right)
    scale = np.array([pl.scale(l) for l in range(cfg[2])], np.float32)
    return kl, dl, kr, dr, scale, [pl.level(l) for l in range(cfg[2])], [pr.level(l) for l in range(cfg[2])]


@pytest.mark.parametrize("shape,seed,cfg,max_disp,mb,mbf", [
    ((240, 640), 12, (1200, 1.2, 8, 20, 7), 40.0, 0.08, 40.0),
    ((376, 1241), 5, (2000, 1.2, 8, 20, 7), 60.0, 0.537, 386.1),     # KITTI calibration (config 2)
    ((240, 320), 3, (500, 1.2, 8, 20, 7), 25.0, 0.1, 20.0),
    ((480, 640), 9, (1000, 1.5, 5, 20, 7), 30.0, 0.12, 47.9),
])
def test_compute_stereo_matches_equals_reference_frame_cc(shape, seed, cfg, max_disp, mb, mbf):
    kl, dl, kr, dr, scale, ll, lr = stereo_inputs(shape, seed, cfg, max_disp)
    inv = (np.float32(1.0) / scale).astype(np.float32)
    our, oz = mc.oracle_stereo_matches(oracle(), kl, dl, kr, dr, scale, inv, ll, lr, mb, mbf)
    rur, rz = oracle_lib.ref_stereo_matches(FR, cfg, kl, dl, kr, dr, ll, lr, mb, mbf)
    assert (rur >= 0).sum() > 50
    assert our.tobytes() == rur.tobytes() and oz.tobytes() == rz.tobytes()


@pytest.mark.parametrize("seed,bounds", [(1, (0.0, 640.0, 0.0, 480.0)), (2, (-12.5, 655.25, -8.0, 490.5)), (3, (0.0, 1241.0, 0.0, 376.0))])
def test_frame_grid_and_features_in_area_equal_reference(seed, bounds):
    from orb_slam2_e_b200.matcher import FrameGrid
    rng = np.random.default_rng(seed)
    min_x, max_x, min_y, max_y = bounds
    n, nq = 3000, 1500
    x = rng.uniform(min_x - 2, max_x + 2, n).astype(np.float32)
    y = rng.uniform(min_y - 2, max_y + 2, n).astype(np.float32)
    # cell-boundary cases: coordinates on half-cell positions, where round() decides the cell
    x[:200] = (min_x + (rng.integers(0, 129, 200) * 0.5) * (max_x - min_x) / 64).astype(np.float32)
    y[:200] = (min_y + (rng.integers(0, 97, 200) * 0.5) * (max_y - min_y) / 48).astype(np.float32)
    octave = rng.integers(0, 8, n).astype(np.int32)
    g = FrameGrid(x, y, octave, synth.random_descriptors(n, seed), None, min_x, min_y, max_x, max_y)
    qx = rng.uniform(min_x - 30, max_x + 30, nq).astype(np.float32)
    qy = rng.uniform(min_y - 30, max_y + 30, nq).astype(np.float32)
    qr = rng.uniform(0.5, 60, nq).astype(np.float32)
    qmin = rng.integers(-1, 5, nq).astype(np.int32)
    qmax = np.where(rng.random(nq) < 0.4, -1, qmin + rng.integers(0, 4, nq)).astype(np.int32)
    cap = 1 << 22
    ro, ri = np.zeros(nq + 1, np.int32), np.zeros(cap, np.int32)
    nr = FR.frameref_features_in_area(x, y, octave, n, min_x, max_x, min_y, max_y, qx, qy, qr, qmin, qmax, nq, ro, ri, cap)
    oo, oi = np.zeros(nq + 1, np.int32), np.zeros(cap, np.int32)
    no = oracle().orc_features_in_area(g.x, g.y, g.octave, n, g.min_x, g.min_y, g.width_inv, g.height_inv, g.cols, g.rows,
                                       g.cell_off, g.cell_idx, qx, qy, qr, qmin, qmax, nq, oo, oi, cap)
    assert nr == no and nr > 10000
    assert np.array_equal(ro, oo) and np.array_equal(ri[:nr], oi[:no])


def test_distinctive_descriptor_equals_reference_mappoint_cc():
    rng = np.random.default_rng(4)
    counts = np.concatenate([[1, 2, 2, 3, 4], rng.integers(1, 40, 300)]).astype(np.int32)
    offset = np.zeros(len(counts) + 1, np.int32)
    np.cumsum(counts, out=offset[1:])
    desc = np.zeros((int(offset[-1]), 32), np.uint8)
    for p in range(len(counts)):  # observations of one point: noisy copies of one descriptor, with exact duplicates and median ties
        base = synth.random_descriptors(1, 100 + p)[0]
        for i in range(offset[p], offset[p + 1]):
            desc[i] = synth.flip_bits(base, int(rng.integers(0, 12)) * 4, rng)
        if counts[p] > 3:
            desc[offset[p] + 1] = desc[offset[p]]
    best = np.zeros(len(counts), np.int32)
    oracle().orc_distinctive_descriptors(desc, offset, len(counts), best)
    out = np.zeros((len(counts), 32), np.uint8)
    FR.frameref_distinctive(desc, offset, None, len(counts), out)
    assert np.array_equal(out, desc[offset[:-1] + best])
    # the choice is by index, not only by value: the reference keeps the FIRST row with the lowest median
    for p in np.flatnonzero(counts > 3)[:50]:
        rows = desc[offset[p]:offset[p + 1]]
        first = int(np.flatnonzero((rows == out[p]).all(axis=1))[0])
        assert first == best[p] or (rows[best[p]] == rows[first]).all()
