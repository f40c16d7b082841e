"""Matching leg of bench.py: brute-force Hamming top-2 of 1000 queries against a 10 M-row
descriptor database (BASELINE.json config 4).  The database is sharded by contiguous row ranges
over the ranks (strong scaling of a fixed database); per-shard (best1, best2, idx) results are
exchanged with ONE NCCL all-gather of nq x 16 B per rank and merged exactly."""
from __future__ import annotations

import time

import numpy as np


def bench_match(args, B, rank, world, local, dist, with_cpu):
    import torch
    from orb_slam2_e_b200 import matcher, reloc, synth, _lib
    from orb_slam2_e_b200.bench_gpu import _barrier, _max_over_ranks
    L = _lib.load()
    D, Q = args.db_rows, args.queries
    lo, hi = reloc.shard_range(D, rank, world)
    rows = hi - lo
    g = torch.Generator(device=f"cuda:{local}")
    g.manual_seed(4000 + rank)
    shard_t = torch.randint(0, 256, (rows, 32), dtype=torch.uint8, device=f"cuda:{local}", generator=g)
    q_host = synth.random_descriptors(Q, 4242)
    # planted neighbours so that acceptance (ratio 0.75, TH_LOW 45) is exercised
    rng = np.random.default_rng(99 + rank)
    plant_rows = rng.integers(0, rows, size=Q // 2)
    plant = np.stack([synth.flip_bits(q_host[i], int(rng.integers(5, 61)), rng) for i in range(Q // 2)])
    shard_t[torch.from_numpy(plant_rows).to(shard_t.device)] = torch.from_numpy(plant).to(shard_t.device)
    db = matcher.DescriptorDB(local, row_base=lo)
    db.append_device(shard_t.data_ptr(), rows)
    q_dev = torch.from_numpy(q_host).cuda(local)
    out_dev = torch.zeros((Q, 2), dtype=torch.int64, device=f"cuda:{local}")  # orbm_top2 = 16 bytes
    gathered = torch.zeros((world, Q, 2), dtype=torch.int64, device=f"cuda:{local}")

    def step():
        db.top2_device(q_dev.data_ptr(), Q, out_dev.data_ptr())
        db.sync()
        if dist is not None:
            dist.all_gather_into_tensor(gathered.view(-1), out_dev.view(-1))
            res = gathered
        else:
            res = out_dev.unsqueeze(0)
        shards = res.cpu().numpy().view(_lib.TOP2_DTYPE).reshape(world, Q)
        return reloc.merge_shards(shards, 0.75, 45)

    for _ in range(max(args.warmup, 1)):
        merged, acc = step()
    _barrier(dist)
    l0 = L.orb_launch_count()
    t0 = time.perf_counter()
    kern_ms = 0.0
    for _ in range(args.steps):
        merged, acc = step()
        kern_ms += db.last_ms()[0]
    torch.cuda.synchronize()
    dt = _max_over_ranks(time.perf_counter() - t0, dist, local)
    kern_ms = _max_over_ranks(kern_ms, dist, local) / args.steps
    launches = int(L.orb_launch_count() - l0)
    _barrier(dist)
    gmatch = Q * D * args.steps / dt / 1e9
    popc_peak = matcher.popc_peak(local)
    achieved = 8.0 * Q * rows / (kern_ms * 1e-3)
    res = {
        "metric": "Hamming Gmatch/s (top-2, ratio 0.75)", "value": gmatch, "unit": "Gmatch/s", "scaling": "strong",
        "kernel_only_gmatch_s_per_gpu": Q * rows / (kern_ms * 1e-3) / 1e9, "ms_per_step": dt / args.steps * 1e3,
        "kernel_ms": kern_ms, "gpu_launches": launches, "accepted": int(acc.sum()),
        "config": {"workload": "%d queries x %d-row database, %d shard(s), NCCL all-gather merge" % (Q, D, world)},
        "roofline": {"bound": "popc", "achieved": achieved / 1e12, "peak": popc_peak / 1e12, "unit": "Tpopc/s",
                     "frac": achieved / popc_peak, "peak_source": "measured in this run (orbm_popc_peak)",
                     "ops_per_match": 8},
    }
    if with_cpu and rank == 0:
        # oracle check + CPU baseline on a bounded slice of this rank's shard
        n_slice = min(rows, args.cpu_db_rows)
        slice_host = shard_t[:n_slice].cpu().numpy()
        threads = __import__("os").cpu_count() or 1
        rate, (d1, d2, i1) = B.cpu_match_rate(slice_host, q_host, threads)
        chk = matcher.DescriptorDB(local, row_base=lo)
        chk.append(slice_host)
        got = chk.top2(q_host)
        ok = bool(np.array_equal(got["best1"], d1) and np.array_equal(got["best2"], d2) and np.array_equal(got["idx"], i1 + lo))
        res["parity_vs_oracle_on_slice"] = ok
        res["cpu_baseline"] = {"value": rate, "unit": "Gmatch/s", "cores": threads, "kind": "port",
                               "sample": "%d queries x %d rows (slice of the shard), SWAR DescriptorDistance + top-2 rule" % (Q, n_slice)}
        if not ok:
            raise SystemExit("bench: GPU top-2 disagrees with the oracle on the checked slice")
    return res
