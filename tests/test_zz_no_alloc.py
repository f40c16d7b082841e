"""orbm_top2_search_noalloc (scratch in a static device array, no allocation in the call) == orbm_top2_search
(the existing entry point, whose results tests/test_gpu_match.py pins against the oracle), bit for bit, on two
calls in a row that share the scratch.  Collected last on purpose (the file name).
"""
import ctypes as C

import numpy as np
import pytest

from orb_slam2_e_b200 import _lib, matcher, synth

pytestmark = pytest.mark.gpu


def _existing(db_rows, q, row_base):
    db = matcher.DescriptorDB(0, row_base=row_base)
    db.append(db_rows)
    out = db.top2(q)
    db.close() if hasattr(db, "close") else None
    return out


def _noalloc(L, d_rows, n_rows, row_base, d_q, nq, stream=0):
    import torch
    d_out = torch.empty(max(nq, 1) * _lib.TOP2_DTYPE.itemsize, dtype=torch.uint8, device="cuda:0")
    rc = L.orbm_top2_search_noalloc(0, C.c_void_p(d_rows.data_ptr()), n_rows, row_base, C.c_void_p(d_q.data_ptr()), nq,
                                    C.c_void_p(d_out.data_ptr()), C.c_void_p(stream))
    _lib.check(rc)
    torch.cuda.synchronize(0)
    return d_out.cpu().numpy().view(_lib.TOP2_DTYPE)[:nq]


def test_no_alloc_search_equals_existing_entry_point_twice_in_a_row():
    import torch
    L = _lib.load()
    # first call: several query blocks and many row chunks; second: fewer of both, so the scratch still holds the
    # first call's partials beside the second's
    cases = [(70000, 2500, 123456, 5), (9000, 37, 0, 6)]
    for n_db, nq, row_base, seed in cases:
        rows, q = synth.reloc_workload(n_db, nq, seed, n_shards=2)
        want = _existing(rows, q, row_base)
        d_rows = torch.from_numpy(np.ascontiguousarray(rows)).to("cuda:0")
        d_q = torch.from_numpy(np.ascontiguousarray(q)).to("cuda:0")
        got = _noalloc(L, d_rows, n_db, row_base, d_q, nq)
        assert np.array_equal(got["best1"], want["best1"])
        assert np.array_equal(got["best2"], want["best2"])
        assert np.array_equal(got["idx"], want["idx"])
    # the same call again, on a caller's stream: identical
    s = torch.cuda.Stream(device=0)
    torch.cuda.synchronize(0)
    again = _noalloc(L, d_rows, n_db, row_base, d_q, nq, stream=s.cuda_stream)
    assert np.array_equal(again, got)


def test_no_alloc_search_empty_database_and_no_queries():
    import torch
    L = _lib.load()
    _, q = synth.reloc_workload(64, 5, 9, n_shards=2)
    d_q = torch.from_numpy(np.ascontiguousarray(q)).to("cuda:0")
    got = _noalloc(L, d_q, 0, 7, d_q, 5)  # rows pointer unused when n_rows == 0
    assert np.all(got["best1"] == 256) and np.all(got["best2"] == 256) and np.all(got["idx"] == -1)
    _lib.check(L.orbm_top2_search_noalloc(0, None, 0, 0, None, 0, None, None))


def test_no_alloc_search_rejects_sizes_beyond_the_static_scratch():
    import torch
    L = _lib.load()
    nq = 304 * 1024 + 1
    d_q = torch.zeros(nq * 32, dtype=torch.uint8, device="cuda:0")
    d_out = torch.empty(nq * _lib.TOP2_DTYPE.itemsize, dtype=torch.uint8, device="cuda:0")
    rc = L.orbm_top2_search_noalloc(0, C.c_void_p(d_q.data_ptr()), 1000, 0, C.c_void_p(d_q.data_ptr()), nq,
                                    C.c_void_p(d_out.data_ptr()), None)
    assert rc != 0
    with pytest.raises(Exception, match="static scratch"):
        _lib.check(rc)
    # misaligned rows
    rc = L.orbm_top2_search_noalloc(0, C.c_void_p(d_q.data_ptr() + 4), 10, 0, C.c_void_p(d_q.data_ptr()), 4,
                                    C.c_void_p(d_out.data_ptr()), None)
    assert rc != 0
