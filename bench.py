#!/usr/bin/env python
"""bench.py -- ORB front end on B200: frames/s of ORB extraction (640x480, 1000 features, TUM
settings) and Gmatch/s of the brute-force Hamming top-2 search (1000 queries vs 10 M rows).

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun)
    python bench.py --impl reference ...                     (the reference's CPU path)

One JSON line on stdout (rank 0).  A "step" = one pass of the extractor over one batch of
synthetic frames per GPU (frames are independent: sharded over ranks with no collective, weak
scaling), followed -- outside the extraction timing -- by the matching measurement, whose
database IS sharded across the ranks and merged with one NCCL all-gather (strong scaling).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

TUM = dict(nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7)
W, H = 640, 480
SEED = 1000  # seed = 1000*config + index


def headline_config():
    """The `config` object of the JSON line: identical in the GPU arm and in --impl reference, so that the
    driver's same_config check compares like with like; what differs per arm (frames per step, sharding,
    L2 note) goes under `run`."""
    return {"workload": "ORBextractor on synthetic textured 640x480 frames, nFeatures=1000, 8 levels, scale 1.2, FAST 20/7 (TUM mono settings)",
            "frame": "%dx%d u8" % (W, H), "settings": dict(TUM), "frame_seed": "synth.bench_frames('textured', n, 480, 640, %d + 7*rank, ...)" % SEED}


def cpu_cv2_baseline(frames, cfg, seconds=4.0):
    """Second CPU anchor: the extractor's image stages through Python cv2 (SIMD / IPP
    OpenCV 4.13): resize + copyMakeBorder per level, FastFeatureDetector per 30-px cell with the iniTh -> minTh
    fallback, GaussianBlur per level.  No quadtree / orientation / descriptors (cv2 has no equivalent entry
    points), so this is a LOWER bound of a real-OpenCV ORB-SLAM2 frame time; the per-cell Python loop is in it."""
    try:
        import cv2
    except Exception as e:  # pragma: no cover
        return {"unavailable": "cv2 import failed: %s" % e}
    sizes = level_sizes(frames.shape[2], frames.shape[1], cfg["scaleFactor"], cfg["nlevels"])
    det_ini = cv2.FastFeatureDetector_create(cfg["iniThFAST"], True)
    det_min = cv2.FastFeatureDetector_create(cfg["minThFAST"], True)

    def one(img):
        n = 0
        lvl = img
        for l, (w, h) in enumerate(sizes):
            if l:
                lvl = cv2.resize(lvl, (w, h), interpolation=cv2.INTER_LINEAR)
            cv2.copyMakeBorder(lvl, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
            min_b, max_x, max_y = 16, w - 16, h - 16
            n_cols, n_rows = (max_x - min_b) // 30, (max_y - min_b) // 30
            w_cell, h_cell = -(-(max_x - min_b) // n_cols), -(-(max_y - min_b) // n_rows)
            for i in range(n_rows):
                y0 = min_b + i * h_cell
                y1 = min(y0 + h_cell + 6, max_y)
                if y0 >= max_y - 3:
                    continue
                for j in range(n_cols):
                    x0 = min_b + j * w_cell
                    x1 = min(x0 + w_cell + 6, max_x)
                    if x0 >= max_x - 6:
                        continue
                    roi = lvl[y0:y1, x0:x1]
                    k = det_ini.detect(roi)
                    if not k:
                        k = det_min.detect(roi)
                    n += len(k)
            cv2.GaussianBlur(lvl, (7, 7), 2, 2, cv2.BORDER_REFLECT_101)
        return n

    out = {"kind": "cv2-%s image stages (pyramid, per-cell FAST, blur); no quadtree/orientation/descriptors" % cv2.__version__}
    default_threads = cv2.getNumThreads()
    for label, nt in (("threads_1", 1), ("threads_default", default_threads)):
        cv2.setNumThreads(nt)
        one(frames[0])
        t0, done = time.time(), 0
        while time.time() - t0 < seconds / 2 and done < len(frames):
            one(frames[done])
            done += 1
        out[label] = {"value": done / (time.time() - t0), "unit": "frames/s", "cv2_threads": nt, "frames": done}
    cv2.setNumThreads(default_threads)
    return out


def level_sizes(w, h, scale, nlevels):
    s = np.float32(1.0)
    out = []
    for l in range(nlevels):
        if l:
            s = np.float32(np.float64(s) * np.float64(np.float32(scale)))
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
    return out


def algorithmic_bytes_per_frame(w, h, cfg, n_kp):
    """every level read once + every derived level written once + 60 B per keypoint."""
    px = [a * b for a, b in level_sizes(w, h, cfg["scaleFactor"], cfg["nlevels"])]
    return sum(px) + sum(px[1:]) + 60 * n_kp, sum(px)


class ClockSampler:
    """nvidia-smi SM clock / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def mark(self):
        return time.time()

    def summary(self, t0, t1):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        rows = [r for t, r in self.rows if t0 - 0.05 <= t <= t1 + 0.15 and len(r) >= 7] or [r for _, r in self.rows[-3:] if len(r) >= 7]
        try:
            sm = sorted(float(r[0]) for r in rows)
            reasons = [n for i, n in ((3, "hw_slowdown"), (4, "hw_thermal_slowdown"), (5, "sw_thermal_slowdown"), (6, "sw_power_cap"))
                       if any(r[i].lower().startswith("active") for r in rows)]
            return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons, "samples": len(rows)}
        except Exception as e:  # pragma: no cover
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["parse error: %s" % e]}

    def stop(self):
        if self.p is not None:
            self.p.terminate()


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------ CPU reference arm ------
def _cpu_libs():
    """(kind, make_extractor): the reference's own ORBextractor.cc build when oracle/_ref
    travelled with the snapshot, else the restatement ("port").  bench.py only TIMES these."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    if oracle_lib.ref_lib(fast=True) is not None:
        return "reference", (lambda c: oracle_lib.RefExtractor(c["nfeatures"], c["scaleFactor"], c["nlevels"], c["iniThFAST"],
                                                               c["minThFAST"], fast=True)), oracle_lib
    return "port", (lambda c: oracle_lib.PortExtractor(c["nfeatures"], c["scaleFactor"], c["nlevels"], c["iniThFAST"],
                                                       c["minThFAST"], fast=True)), oracle_lib


def cpu_extract_fps(frames, cfg, threads, budget_s=20.0):
    """Frames/s of the CPU path on `threads` host threads (one extractor object per thread, the
    way Frame.cc:78-81 uses two).  ctypes releases the GIL during each call."""
    kind, mk, _ = _cpu_libs()
    exts = [mk(cfg) for _ in range(threads)]
    exts[0].extract(frames[0])  # warm
    done = [0] * threads
    t_end = time.time() + budget_s
    n = len(frames)

    def work(t):
        i = t
        while i < n and time.time() < t_end:
            exts[t].extract(frames[i])
            done[t] += 1
            i += threads

    t0 = time.time()
    th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    [x.start() for x in th]
    [x.join() for x in th]
    dt = time.time() - t0
    return sum(done) / dt, sum(done), kind


def cpu_match_rate(db, q, threads):
    """Gmatch/s of the brute-force scan on host threads.  kind "reference": the reference's own
    ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1848-1864, compiled unmodified into oracle/_ref/libmatchref_fast.so)
    called on one-row cv::Mat views with the update rule of :417-426; kind "port": the restatement, when that library did
    not travel with the snapshot."""
    import ctypes as C
    _, _, ol = _cpu_libs()
    d1 = np.zeros(len(q), np.int32); i1 = np.zeros(len(q), np.int64); d2 = np.zeros(len(q), np.int32)
    path = os.path.join(ROOT, "oracle", "_ref", "libmatchref_fast.so")
    kind = "port"
    fn = None
    if os.path.exists(path):
        try:
            R = C.CDLL(path)
            R.refm_top2_scan_mt.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
            R.refm_top2_scan_mt.restype = None
            fn = lambda: R.refm_top2_scan_mt(q.ctypes.data, len(q), db.ctypes.data, len(db), d1.ctypes.data, i1.ctypes.data, d2.ctypes.data, threads)
            kind = "reference"
        except (OSError, AttributeError):
            fn = None
    if fn is None:
        L = ol.oracle(fast=True)
        fn = lambda: L.orc_top2_mt(q, len(q), db, len(db), d1, i1, d2, threads)
    t0 = time.time()
    fn()
    dt = time.time() - t0
    return len(q) * len(db) / dt / 1e9, (d1, d2, i1), kind


def run_reference_arm(args):
    """--impl reference: the reference's CPU extractor on all host cores, same metric/config."""
    from orb_slam2_e_b200 import synth
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    per_step = 16 * threads  # bounded sample of the workload per step (under a second of wall time)
    frames = synth.bench_frames("textured", per_step, H, W, SEED, min(per_step, 64), workers=threads)
    kind, mk, _ = _cpu_libs()
    for _ in range(args.warmup):
        cpu_extract_fps(frames[:threads], TUM, threads, budget_s=60)
    t0 = time.time()
    total = 0
    for _ in range(args.steps):
        _, n, kind = cpu_extract_fps(frames, TUM, threads, budget_s=600)
        total += n
    dt = time.time() - t0
    fps = total / dt
    line = {
        "impl": "reference", "metric": "frames/s ORB extract (640x480, 1k feats)", "value": fps, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": headline_config(),
        "run": {"frames_per_step": per_step, "arm": "reference's own ORBextractor.cc over the OpenCV-exact scalar shim, -O3 -march=x86-64-v3"},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind,
                         "sample": "%d frames per step, one extractor object per host thread" % per_step},
        "cpu_baseline_cv2": cpu_cv2_baseline(frames[:8], TUM),
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# ------------------------------------------------------------------ GPU arm ----------------
def run_gpu_arm(args):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    B = sys.modules[__name__]
    from orb_slam2_e_b200 import bench_gpu, bench_match
    rank, world, local, dist = bench_gpu._dist_setup()
    ex = bench_gpu.bench_extract(args, B, rank, world, local, dist)
    cpu = None
    if rank == 0 and not args.no_cpu:
        threads = os.cpu_count() or 1
        sample = ex["host"][: min(len(ex["host"]), 40 * threads)]  # ~25 s of CPU work across the host threads
        fps, n_done, kind = cpu_extract_fps(sample, TUM, threads, budget_s=args.cpu_seconds)
        cpu = {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind,
               "sample": "%d of the step's frames, one extractor object per host thread (%s ORBextractor over the OpenCV-exact shim)"
                         % (n_done, "reference's own" if kind == "reference" else "restated")}
    cpu_cv2 = cpu_cv2_baseline(ex["host"][:8], TUM) if (rank == 0 and not args.no_cpu) else None
    ex.pop("host")
    ext = ex.pop("ext")
    ext.close()
    torch.cuda.empty_cache()
    configs = None
    if not args.no_configs:
        from orb_slam2_e_b200 import bench_configs as BC
        with_cpu = not args.no_cpu
        configs = {}
        configs["c3"] = BC.bench_batch_config(B, "c3", BC.HAMLYN, "endoscopy", W, H, args.frames, args.distinct, args, rank, world, local,
                                              dist, with_cpu, 3000)
        configs["c5"] = BC.bench_batch_config(B, "c5", BC.BIG, "textured", 3840, 2160, args.c5_frames, min(args.c5_frames, 4), args, rank,
                                              world, local, dist, with_cpu, 5000)
        configs["c2"] = BC.bench_stereo(B, args, rank, world, local, dist, with_cpu, args.c2_pairs)
        if rank == 0:
            try:
                configs["c1_single"] = BC.bench_single(B, args, local)
            except Exception as e:  # the C++ harness is test infrastructure; its absence must not lose the line
                configs["c1_single"] = {"unavailable": str(e)}
    match = None
    if not args.no_match:
        match = bench_match.bench_match(args, B, rank, world, local, dist, with_cpu=not args.no_cpu)
    if dist is not None:
        dist.barrier()
    if rank == 0:
        line = {
            "metric": "frames/s ORB extract (640x480, 1k feats)", "value": ex["fps"], "unit": "frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ex["ms"] / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": headline_config(),
            "run": {"frames_per_gpu": args.frames, "frames_generated": min(args.distinct, args.frames),
                    "sharding": "frames block-partitioned over ranks, no collective",
                    "l2": "inputs larger than L2 (%d MB of frames per step per GPU), no flush" % (args.frames * W * H >> 20),
                    "keypoints_per_frame": ex["n_kp"], "fast_candidates_per_frame": ex["n_cand"]},
            "e2e": {"value": ex["e2e"], "unit": "frames/s", "h2d_bytes_per_step": ex["h2d"], "d2h_bytes_per_step": ex["d2h"],
                    "h2d_ceiling_gbs": ex["h2d_gbs"], "h2d_ceiling_frames_s": ex["h2d_ceiling_fps"],
                    "frac_of_h2d_ceiling": ex["e2e"] / ex["h2d_ceiling_fps"],
                    "duplex_ceiling_frames_s": ex["duplex_ceiling_fps"], "frac_of_duplex_ceiling": ex["e2e"] / ex["duplex_ceiling_fps"],
                    "duplex_ceiling_note": "the same probe with the step's result bytes copied device-to-host at the same time: what the host link "
                                           "carries when both directions of the e2e call run with no kernels",
                    "h2d_ceiling_note": "copy-only probe in this run: the step's frames from the same pinned buffer, no kernels, "
                                        "all ranks at once, max over ranks (aggregate GB/s)"},
            "gpu_launches": ex["launches"] + (match["gpu_launches"] if match else 0),
            "clocks": ex["clocks"], "roofline": ex["roofline"], "roofline_issue": ex["roofline_issue"], "roofline_pipeline": ex["roofline_pipeline"],
            "cpu_baseline": cpu, "cpu_baseline_cv2": cpu_cv2, "matching": match, "configs": configs,
        }
        _emit(line)
    if dist is not None:
        dist.destroy_process_group()


_REAL_STDOUT = None


def _emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=1024, help="frames per GPU per step")
    ap.add_argument("--distinct", type=int, default=128, help="frames produced by the generator; the rest of --frames are cyclic shifts of them (all different)")
    ap.add_argument("--c5-frames", type=int, default=16, help="3840x2160 frames per GPU per step of the configs.c5 leg")
    ap.add_argument("--c2-pairs", type=int, default=64, help="stereo pairs per GPU per step of the configs.c2 leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-config legs (c1_single, c2, c3, c5)")
    ap.add_argument("--db-rows", type=int, default=10_000_000)
    ap.add_argument("--queries", type=int, default=1000)
    ap.add_argument("--cpu-db-rows", type=int, default=200_000)
    ap.add_argument("--cpu-seconds", type=float, default=20.0)
    ap.add_argument("--chunk-frames", type=int, default=0, help="frames per upload|extract|download chunk of the e2e path")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-match", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    # stdout carries exactly ONE JSON line: native libraries (NCCL prints its version banner on fd 1
    # whenever the box sets NCCL_DEBUG) write to stderr for the whole run, and _emit() writes the line
    # to the saved descriptor
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        if args.warmup < 3:
            args.warmup = 3  # timing rule: at least 3 warm-up steps
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
