"""N>1 path on CPU: world_size-2 (and 3, with an empty shard) gloo groups run the product's
exchange + exact merge; every rank must end with the result of ONE sequential scan."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from oracle_lib import oracle
from orb_slam2_e_b200 import reloc, synth

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_ranges_tile_the_database():
    for n, w in ((10_000_000, 8), (7, 3), (2, 4), (0, 2)):
        r = [reloc.shard_range(n, k, w) for k in range(w)]
        assert r[0][0] == 0 and r[-1][1] == n
        assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))


@pytest.mark.parametrize("world,n_rows", [(2, 6000), (3, 2)])
def test_gloo_exchange_and_merge(tmp_path, world, n_rows):
    nq = 48
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(HERE, "dist_worker.py"), str(tmp_path), str(n_rows), str(nq)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    db, q = synth.reloc_workload(n_rows, nq, 17, n_shards=world)
    L = oracle()
    d1 = np.zeros(nq, np.int32); i1 = np.zeros(nq, np.int64); d2 = np.zeros(nq, np.int32)
    L.orc_top2(q, nq, db, len(db), d1, i1, d2)
    want_acc = (d1 <= 45) & (d1.astype(np.float32) < np.float32(0.75) * d2.astype(np.float32))
    for rank in range(world):
        m = np.load(tmp_path / f"merged_{rank}.npy")
        assert np.array_equal(m["best1"], d1) and np.array_equal(m["best2"], d2) and np.array_equal(m["idx"], i1)
        assert np.array_equal(np.load(tmp_path / f"acc_{rank}.npy"), want_acc)
