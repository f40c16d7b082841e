"""GPU parity: CUDA matcher kernels (through the C ABI) vs the CPU oracle, bit-exact."""
import numpy as np
import pytest

from oracle_lib import oracle
from orb_slam2_e_b200 import matcher, synth, _lib
import match_cases as mc

pytestmark = pytest.mark.gpu


def _oracle_top2(q, db):
    L = oracle(fast=True)
    d1 = np.zeros(len(q), np.int32); i1 = np.zeros(len(q), np.int64); d2 = np.zeros(len(q), np.int32)
    L.orc_top2_mt(q, len(q), db, len(db), d1, i1, d2, 8)
    return d1, d2, i1


def test_library_loaded_is_in_tree():
    import os
    assert os.path.exists(_lib.lib_path())
    assert _lib.load().orb_device_count() >= 1


def test_descriptor_distance():
    m = matcher.ORBmatcher(0.75, True)
    a = synth.random_descriptors(2000, 1)
    b = synth.random_descriptors(2000, 2)
    b[:10] = a[:10]
    b[10:20] = ~a[10:20]
    got = m.DescriptorDistances(a, b)
    L = oracle()
    want = np.array([L.orc_hamming(a[i], b[i]) for i in range(len(a))], np.int32)
    assert np.array_equal(got, want)
    assert matcher.ORBmatcher.DescriptorDistance(a[3], b[30]) == L.orc_hamming(a[3], b[30])


@pytest.mark.parametrize("n_db,n_q", [(0, 5), (1, 3), (255, 33), (256, 1024), (257, 1025), (40000, 300), (300000, 1000)])
def test_top2_matches_oracle(n_db, n_q):
    db, q = synth.reloc_workload(max(n_db, 1), n_q, 11 + n_db, n_shards=4)
    db = db[:n_db]
    shard = matcher.DescriptorDB(0, row_base=0)
    if n_db:
        shard.append(db[: n_db // 2])
        shard.append(db[n_db // 2:])  # growth path
    assert len(shard) == n_db
    got = shard.top2(q)
    if n_db == 0:
        assert np.all(got["best1"] == 256) and np.all(got["idx"] == -1)
        return
    d1, d2, i1 = _oracle_top2(q, db)
    assert np.array_equal(got["best1"], d1)
    assert np.array_equal(got["best2"], d2)
    assert np.array_equal(got["idx"], i1)


def test_sharded_merge_equals_single_scan():
    """Config 4 semantics on one GPU: R shards scanned separately, merged, == one scan."""
    db, q = synth.reloc_workload(200000, 1000, 5, n_shards=8)
    d1, d2, i1 = _oracle_top2(q, db)
    for R in (2, 8):
        res = []
        for r in range(R):
            lo, hi = len(db) * r // R, len(db) * (r + 1) // R
            s = matcher.DescriptorDB(0, row_base=lo)
            s.append(db[lo:hi])
            res.append(s.top2(q))
        merged, acc = matcher.merge_top2(res, 0.75, 45)
        assert np.array_equal(merged["best1"], d1) and np.array_equal(merged["best2"], d2) and np.array_equal(merged["idx"], i1)
        want = (d1 <= 45) & (d1.astype(np.float32) < np.float32(0.75) * d2.astype(np.float32))
        assert np.array_equal(acc, want) and acc.sum() > 100


@pytest.mark.parametrize("kf_kf", [False, True])
@pytest.mark.parametrize("check_ori", [False, True])
def test_search_by_bow(kf_kf, check_ori):
    for seed in (1, 2, 3):
        case = mc.bow_case(seed, n1=900 if seed == 3 else 300, n2=1000 if seed == 3 else 320, n_nodes=25 if seed == 3 else 40)
        nm, out = mc.oracle_bow(oracle(), case, 0.75, check_ori, kf_kf)
        m = matcher.ORBmatcher(0.75, check_ori)
        gnm, gout = m.SearchByBoW(case["d1"], case["fv1"], case["v1"], case["d2"], case["fv2"],
                                  case["v2"] if kf_kf else None, case["a1"], case["a2"], kf_kf=kf_kf)
        assert gnm == nm and np.array_equal(gout, out)
        assert nm > 20


@pytest.mark.parametrize("kf_kf", [False, True])
@pytest.mark.parametrize("check_ori", [False, True])
def test_search_by_bow_batch_equals_sequential_calls(kf_kf, check_ori):
    """K candidate keyframes against one frame in one call == the oracle's (and orbm_search_bow's) K sequential searches;
    includes a candidate without features, one without a common node, and one whose list repeats a feature (serial fallback)."""
    base = mc.bow_case(7, n1=600, n2=640, n_nodes=30)
    cands, want = [], []
    for seed, c in enumerate(mc.bow_candidates(base, 8, n1=300, seed=5)):
        if seed == 5:  # a feature listed under two nodes: nodes are no longer independent
            k0, k1 = sorted(c["fv1"])[:2]
            c["fv1"][k1] = list(c["fv1"][k1]) + [c["fv1"][k0][0]]
        cands.append((c["d1"], c["fv1"], c["v1"], c["a1"]))
        want.append(mc.oracle_bow(oracle(), c, 0.75, check_ori, kf_kf))
    cands.append((np.zeros((0, 32), np.uint8), {}, np.zeros(0, np.uint8), np.zeros(0, np.float32)))
    want.append((0, np.full(0 if kf_kf else 640, -1, np.int32)))
    cands.append((base["d1"][:5], {10 ** 6: [0, 1, 2]}, np.ones(5, np.uint8), np.zeros(5, np.float32)))
    want.append((0, np.full(5 if kf_kf else 640, -1, np.int32)))
    m = matcher.ORBmatcher(0.75, check_ori)
    got = m.SearchByBoWBatch(cands, base["d2"], base["fv2"], base["v2"] if kf_kf else None, base["a2"], kf_kf=kf_kf)
    assert len(got) == len(want)
    for p, ((gnm, gout), (nm, out)) in enumerate(zip(got, want)):
        assert gnm == nm and np.array_equal(gout, out), p
    assert sum(nm for nm, _ in want) > 100


def test_search_by_bow_empty():
    m = matcher.ORBmatcher(0.75, False)
    d = synth.random_descriptors(10, 1)
    nm, out = m.SearchByBoW(d, {1: [0, 1]}, np.ones(10, np.uint8), d, {2: [0, 1]}, None, None, None)
    assert nm == 0 and np.all(out == -1)


def test_stereo_search():
    for seed in (4, 5):
        kl, dl, kr, dr, scale, n_rows = mc.stereo_case(seed, nl=2000, nr=2100)
        bi, bd = mc.oracle_stereo(oracle(), kl, dl, kr, dr, scale, n_rows, 0.0, 80.0)
        m = matcher.ORBmatcher()
        gi, gd = m.StereoSearch(kl, dl, kr, dr, scale, n_rows, 0.0, 80.0)
        assert np.array_equal(gd, bd) and np.array_equal(gi, bi)
        assert (bd == -1).sum() >= 3 and ((bd >= 0) & (bd < 95)).sum() > 50


def test_popc_peak_is_plausible():
    v = matcher.popc_peak(0)
    assert 1e12 < v < 2e13


def test_distinctive_descriptors_batch():
    """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:305-370), a keyframe's worth of map points at once."""
    rng = np.random.default_rng(6)
    sizes = np.concatenate([[0, 1, 2, 3, 32, 33, 64, 200], rng.integers(1, 40, 500)])
    off = np.zeros(len(sizes) + 1, np.int32)
    np.cumsum(sizes, out=off[1:])
    desc = synth.random_descriptors(int(off[-1]), 8)
    for p in range(len(sizes)):  # observations of one point look alike; several are identical (median ties)
        base = desc[off[p]] if sizes[p] else None
        for j in range(off[p] + 1, off[p + 1]):
            desc[j] = base if rng.random() < 0.15 else synth.flip_bits(base, int(rng.integers(0, 60)), rng)
    want = np.zeros(len(sizes), np.int32)
    oracle().orc_distinctive_descriptors(desc, off, len(sizes), want)
    got = matcher.distinctive_descriptors(desc, off)
    assert np.array_equal(got, want)
    assert got[0] == -1 and (got[1:] >= 0).all() and len(np.unique(got[8:])) > 5


def test_descriptor_db_file_round_trip(tmp_path):
    db, q = synth.reloc_workload(5000, 50, 4, n_shards=2)
    a = matcher.DescriptorDB(0, 0)
    a.append(db)
    path = tmp_path / "db.orbd"
    a.save(path)
    raw = path.read_bytes()
    assert raw[:4] == b"ORBD" and int.from_bytes(raw[4:8], "little") == 1 and int.from_bytes(raw[8:16], "little") == 5000
    assert raw[16:] == db.tobytes()
    b = matcher.DescriptorDB(0, 0)
    b.load(path)
    assert len(b) == 5000 and b.top2(q).tobytes() == a.top2(q).tobytes()
    # two shards straight from the file == the single scan
    s0, s1 = matcher.DescriptorDB(0, 0), matcher.DescriptorDB(0, 2500)
    s0.load(path, 0, 2500)
    s1.load(path, 2500, -1)
    merged, _ = matcher.merge_top2([s0.top2(q), s1.top2(q)])
    whole = a.top2(q)
    for f in ("best1", "best2", "idx"):
        assert np.array_equal(merged[f], whole[f])
