"""orbm_window_search on the GPU against the CPU restatement (oracle/match_port.hpp window_search,
which follows ORBmatcher.cc:46-132 over Frame::GetFeaturesInArea, Frame.cc:342-395).
Integer / index work: every field of every result must be equal."""
import numpy as np
import pytest

import match_cases as mc
from oracle_lib import oracle
from orb_slam2_e_b200 import matcher
from orb_slam2_e_b200._lib import OrbFrontError, WINDOW_QUERY_DTYPE

pytestmark = pytest.mark.gpu

FIELDS = ("idx", "best_idx", "best_dist", "best_dist2", "best_level", "best_level2")


def _check(frame, q, qd, claimed, th, ratio, claim, sigma=None):
    want_n, want = mc.oracle_window(oracle(), frame, q, qd, claimed, th, ratio, claim, sigma)
    got, got_n, passes = matcher.window_search(frame, q, qd, claimed, th, ratio, claim, inv_level_sigma2=sigma)
    for f in FIELDS:
        bad = np.flatnonzero(got[f] != want[f])
        assert len(bad) == 0, (f, bad[:5], got[bad[:5]], want[bad[:5]])
    assert got_n == want_n
    return passes, want_n


@pytest.mark.parametrize("claim", [True, False])
@pytest.mark.parametrize("ratio", [0.0, 0.8])
@pytest.mark.parametrize("kw", [dict(), dict(stereo=True), dict(th_px=15.0, n=2500, nq=3000),
                                dict(clump=0.6, dup=0.9), dict(n=60, nq=40), dict(claimed_frac=0.0, nonblocking_frac=0.0)])
def test_window_search_matches_oracle(kw, ratio, claim):
    frame, q, qd, claimed = mc.window_case(11, **kw)
    passes, n = _check(frame, q, qd, claimed, matcher.TH_HIGH, ratio, claim)
    assert n > 0
    assert passes == 1 if not claim else passes >= 2


def test_claim_chain_takes_many_passes_and_stays_exact():
    """Every query wants the same few keypoints with the same descriptor: query k can only be
    resolved after query k-1 (worst case for the parallel passes)."""
    rng = np.random.default_rng(3)
    n, nq = 48, 64
    x = (300 + 2.0 * np.arange(n) % 9).astype(np.float32)
    y = (200 + np.arange(n) // 9).astype(np.float32)
    from orb_slam2_e_b200 import synth
    base = synth.random_descriptors(1, 5)[0]
    desc = np.stack([synth.flip_bits(base, k, rng) for k in range(n)])  # distance k to the query descriptor (nested flips not needed)
    frame = matcher.FrameGrid(x, y, np.zeros(n, np.int32), desc)
    q = np.zeros(nq, WINDOW_QUERY_DTYPE)
    q["x"], q["y"], q["r"], q["er_max"], q["min_level"], q["max_level"], q["blocks"] = 304, 202, 30, -1, 0, 0, 1
    qd = np.repeat(base[None, :], nq, 0)
    passes, n_acc = _check(frame, q, qd, None, 256, 0.0, True)
    assert n_acc == n and passes > 10


def test_window_outside_grid_and_empty_inputs():
    frame, q, qd, claimed = mc.window_case(5, n=300, nq=50)
    q["x"][:10] = -500.0      # window left of the grid: Frame.cc:351-353 returns nothing
    q["x"][10:20] = 5000.0    # right of it: :347-349
    q["y"][20:30] = -400.0
    q["y"][30:40] = 4000.0
    _check(frame, q, qd, claimed, matcher.TH_HIGH, 0.8, True)
    got, n, _ = matcher.window_search(frame, q[:0], qd[:0], claimed)
    assert len(got) == 0 and n == 0
    empty = matcher.FrameGrid(np.zeros(0), np.zeros(0), np.zeros(0), np.zeros((0, 32), np.uint8))
    got, n, _ = matcher.window_search(empty, q, qd, None)
    assert n == 0 and (got["best_idx"] == -1).all() and (got["best_dist"] == 256).all()


def test_keyframe_sized_grid_and_custom_cells():
    frame, q, qd, claimed = mc.window_case(9, n=800, nq=600, width=1241, height=376)
    _check(frame, q, qd, claimed, matcher.TH_LOW, 0.0, True)


def test_rejects_keypoint_in_two_cells():
    cells = [[] for _ in range(64 * 48)]
    cells[5] = [0, 1]
    cells[6] = [1]
    with pytest.raises(OrbFrontError):
        f = matcher.FrameGrid(np.ones(2), np.ones(2), np.zeros(2), np.zeros((2, 32), np.uint8), cells=cells)
        q = np.zeros(1, WINDOW_QUERY_DTYPE)
        matcher.window_search(f, q, np.zeros((1, 32), np.uint8))


@pytest.mark.parametrize("stereo", [False, True])
def test_reprojection_gate_of_fuse(stereo):
    """ORBmatcher.cc:1118-1146: chi-square gate on the reprojection error, per keypoint level."""
    frame, q, qd, claimed = mc.window_case(13, stereo=stereo, th_px=6.0)
    sigma = (1.0 / (1.2 ** np.arange(8)) ** 2).astype(np.float32)
    q["er_max"] = -1.0
    if stereo:
        q["u_right"] = q["x"] - 12.0
    _, n_gate = _check(frame, q, qd, None, matcher.TH_LOW, 0.0, False, sigma)
    _, n_free = _check(frame, q, qd, None, matcher.TH_LOW, 0.0, False, None)
    assert 0 < n_gate < n_free


@pytest.mark.parametrize("kw", [dict(th_px=40.0, n=1500, nq=400), dict(n=50, nq=30), dict(th_px=3.0)])
def test_window_candidates_lists(kw):
    """orbm_window_candidates: every candidate of every window, in GetFeaturesInArea's order, with its
    distance -- against a brute-force numpy model of Frame.cc:342-395."""
    frame, q, qd, _ = mc.window_case(17, **kw)
    got = matcher.window_candidates(frame, q, qd)
    pop = np.array([bin(i).count("1") for i in range(256)], np.int32)
    cell_of = np.full(frame.n, -1)
    for c in range(frame.cols * frame.rows):
        cell_of[frame.cell_idx[frame.cell_off[c]:frame.cell_off[c + 1]]] = c
    pos_of = np.empty(frame.n, np.int64)
    pos_of[frame.cell_idx] = np.arange(len(frame.cell_idx))
    total = 0
    for i in range(len(q)):
        x, y, r = q["x"][i], q["y"][i], q["r"][i]
        f32 = np.float32
        cx0 = max(0, int(np.floor((x - frame.min_x - r) * frame.width_inv)))
        cx1 = min(frame.cols - 1, int(np.ceil((x - frame.min_x + r) * frame.width_inv)))
        cy0 = max(0, int(np.floor((y - frame.min_y - r) * frame.height_inv)))
        cy1 = min(frame.rows - 1, int(np.ceil((y - frame.min_y + r) * frame.height_inv)))
        ok = (cell_of >= 0) & (cell_of // frame.rows >= cx0) & (cell_of // frame.rows <= cx1) & (cell_of % frame.rows >= cy0) & \
             (cell_of % frame.rows <= cy1)
        lo, hi = q["min_level"][i], q["max_level"][i]
        if lo > 0 or hi >= 0:
            ok &= frame.octave >= lo
            if hi >= 0:
                ok &= frame.octave <= hi
        ok &= (np.abs(frame.x - x) < r) & (np.abs(frame.y - y) < r)
        want = np.flatnonzero(ok)
        want = want[np.argsort(pos_of[want])]
        assert np.array_equal(got[i][0], want), i
        assert np.array_equal(got[i][1], pop[frame.desc[want] ^ qd[i]].sum(1))
        total += len(want)
    assert total > 0
