"""Worker for tests/test_dist_cpu.py: one rank of a gloo group.  Scans its shard with the CPU
oracle (the checker stands in for the GPU scan, which needs a device), then runs the PRODUCT's
exchange + merge (orb_slam2_e_b200.reloc) and writes the merged result."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def main():
    out_dir, n_rows, nq = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    import torch.distributed as dist
    from orb_slam2_e_b200 import reloc, synth, _lib
    import oracle_lib
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    db, q = synth.reloc_workload(n_rows, nq, 17, n_shards=world)
    lo, hi = reloc.shard_range(n_rows, rank, world)
    L = oracle_lib.oracle()
    d1 = np.zeros(nq, np.int32); i1 = np.zeros(nq, np.int64); d2 = np.zeros(nq, np.int32)
    if hi > lo:
        L.orc_top2(q, nq, np.ascontiguousarray(db[lo:hi]), hi - lo, d1, i1, d2)
        i1 += lo
    else:
        d1[:] = 256; d2[:] = 256; i1[:] = -1
    local = np.zeros(nq, _lib.TOP2_DTYPE)
    local["best1"], local["best2"], local["idx"] = d1, d2, i1
    gathered = reloc.exchange_top2(local)
    assert gathered.shape == (world, nq)
    merged, acc = reloc.merge_shards(gathered, 0.75, 45)
    np.save(os.path.join(out_dir, f"merged_{rank}.npy"), merged)
    np.save(os.path.join(out_dir, f"acc_{rank}.npy"), acc)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
