"""orbm_group_search (scan -> ncclAllGather -> device merge + acceptance, inside liborbfront.so) == one scan of the
whole database followed by orbm_merge_top2 (whose rule tests/test_oracle_match.py and test_gpu_match.py pin)."""
import numpy as np
import pytest

from orb_slam2_e_b200 import matcher, reloc, synth

pytestmark = pytest.mark.gpu


def _device_count():
    import torch
    return torch.cuda.device_count()


def _check(n_shards, n_db=60000, nq=300):
    db, q = synth.reloc_workload(n_db, nq, 11, n_shards=max(n_shards, 2))
    whole = matcher.DescriptorDB(0, 0)
    whole.append(db)
    want, want_acc = matcher.merge_top2([whole.top2(q)], 0.75, 45)
    shards = []
    for r in range(n_shards):
        lo, hi = reloc.shard_range(n_db, r, n_shards)
        s = matcher.DescriptorDB(r, row_base=lo)
        s.append(db[lo:hi])
        shards.append(s)
    g = matcher.ShardGroup.local(shards)
    for _ in range(2):  # the second call reuses every buffer
        got, acc = g.search(q, 0.75, 45)
        assert np.array_equal(got, want) and np.array_equal(acc, want_acc)
    scan_ms, total_ms = g.last_ms()
    assert 0 < scan_ms <= total_ms
    # fewer queries than the buffers hold, and an odd count
    got, acc = g.search(q[:37], 0.75, 45)
    assert np.array_equal(got, want[:37]) and np.array_equal(acc, want_acc[:37])
    g.close()


def test_group_of_one_shard():
    _check(1)


@pytest.mark.parametrize("n_shards", [2, 4, 8])
def test_group_over_devices(n_shards):
    if _device_count() < n_shards:
        pytest.skip("needs %d GPUs" % n_shards)
    _check(n_shards)


def test_two_shards_on_one_device_are_rejected():
    a, b = matcher.DescriptorDB(0, 0), matcher.DescriptorDB(0, 10)
    with pytest.raises(Exception):
        matcher.ShardGroup.local([a, b])
