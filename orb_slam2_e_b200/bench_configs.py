"""Per-config legs of bench.py: every configuration BASELINE.json names is measured, not only the
headline (round-1 review, row 8d-2).  Each leg returns one object for the `configs` key of the JSON line:

  c1_single  one 640x480 frame through the C++ ORBextractor::operator() (tests/cpp/libdropin.so), microseconds per
             call with and without the mvImagePyramid download -- how Tracking drives the class (Frame.cc:262-268)
  c2         KITTI-shape stereo pairs (1241x376, 2000 features): both images extracted as one batch + the whole
             Frame::ComputeStereoMatches on the device (orbx_stereo_matches), pairs/s
  c3         low-texture endoscopy frames, Hamlyn settings (1200 / 1.1 / 6 / 24 / 7), batch per GPU, frames/s
  c5         3840x2160 frames, 8000 features, 12 levels, frames/s

Timing: CUDA events on the extractor's stream (device-resident `value`), wall clock around the host-buffer
call (`e2e`), >= 3 warm-ups, inputs larger than L2 or stated otherwise; a parity spot-check of 4 frames of the
batch against the CPU oracle and the CPU reference on a sample of the same frames run outside the timed region.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

HAMLYN = dict(nfeatures=1200, scaleFactor=1.1, nlevels=6, iniThFAST=24, minThFAST=7)   # roslaunch/sHamlyn05.yaml:71-84
KITTI = dict(nfeatures=2000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7)    # Examples/Stereo/KITTI00-02.yaml:18-51
BIG = dict(nfeatures=8000, scaleFactor=1.2, nlevels=12, iniThFAST=20, minThFAST=7)     # BASELINE.json config 5
KITTI_BF, KITTI_FX = 386.1448, 718.856                                                 # KITTI00-02.yaml: Camera.bf, Camera.fx


def _mk(cfg, device, w, h, n):
    from orb_slam2_e_b200.extractor import ORBextractor
    return ORBextractor(cfg["nfeatures"], cfg["scaleFactor"], cfg["nlevels"], cfg["iniThFAST"], cfg["minThFAST"],
                        device=device, max_width=w, max_height=h, max_batch=n)


def parity_spot_check(B, cfg, frames, results, n_check=4):
    """results[i] = (keypoints, descriptors) of frames[i] from the CUDA path; compared with the CPU oracle on
    n_check frames spread over the batch: keypoints exact, angles <= 1e-3 rad, descriptor bits <= 1e-4."""
    # the PARITY oracle: the reference's ORBextractor.cc built with the monotonic arena (oracle/_ref/liborbref.so) when it
    # travelled with the snapshot, else the restatement.  (The -O3 "fast" build that bench.py TIMES uses glibc malloc, whose
    # pointer order makes DistributeOctTree's tie-break at ORBextractor.cc:684 vary: not a parity reference.)
    _, _, ol = B._cpu_libs()
    a = (cfg["nfeatures"], cfg["scaleFactor"], cfg["nlevels"], cfg["iniThFAST"], cfg["minThFAST"])
    ext = ol.RefExtractor(*a) if ol.ref_lib() is not None else ol.PortExtractor(*a)
    idx = sorted(set(int(i) for i in np.linspace(0, len(frames) - 1, n_check)))
    worst_angle, bit_bad, bit_all = 0.0, 0, 0
    for i in idx:
        pk, pd = ext.extract(frames[i])
        kps, desc = results[i]
        if len(kps) != len(pk):
            return {"frames_checked": len(idx), "ok": False, "why": "frame %d: %d keypoints, oracle %d" % (i, len(kps), len(pk))}
        for name in ("x", "y", "octave", "response", "size"):
            if not np.array_equal(kps[name], pk[name]):
                return {"frames_checked": len(idx), "ok": False, "why": "frame %d: field %s differs" % (i, name)}
        if len(kps):
            worst_angle = max(worst_angle, float(np.max(np.abs(kps["angle"] - pk["angle"]))))
            bit_bad += int(np.unpackbits(desc ^ pd).sum())
            bit_all += desc.size * 8
    ok = worst_angle < 0.0573 and bit_bad <= 1e-4 * max(bit_all, 1)
    return {"frames_checked": len(idx), "ok": bool(ok), "max_angle_diff_deg": worst_angle, "descriptor_bit_mismatch_rate": bit_bad / max(bit_all, 1)}


def bench_batch_config(B, name, cfg, kind, w, h, n, generated, args, rank, world, local, dist, with_cpu, seed):
    """Batched extraction leg (c3, c5): n frames per GPU per step."""
    import torch
    from orb_slam2_e_b200 import synth, _lib
    from orb_slam2_e_b200.bench_gpu import _barrier, _max_over_ranks
    L = _lib.load()
    workers = max(1, (os.cpu_count() or 1) // max(world, 1))
    host = synth.bench_frames(kind, n, h, w, seed + 7 * rank, generated, workers)
    dev = torch.from_numpy(host).cuda(local)
    ext = _mk(cfg, local, w, h, n)
    stream = torch.cuda.ExternalStream(ext.cuda_stream(), device=local)

    def step():
        ext.extract_batch_device(dev.data_ptr(), n, h, w, w, w * h)

    for _ in range(3):
        step()
    ext.sync()
    counts = ext.counts()
    n_kp = float(counts.mean())
    ext.set_stage_timing(True)
    stages = None
    for _ in range(2):
        step()
        ext.sync()
        st = ext.last_stage_ms()
        if stages is None or sum(st.values()) < sum(stages.values()):
            stages = st
    ext.set_stage_timing(False)
    _barrier(dist)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = L.orb_launch_count()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    e1.synchronize()
    _barrier(dist)
    launches = int(L.orb_launch_count() - l0)
    ms = _max_over_ranks(e0.elapsed_time(e1), dist, local)
    fps = world * n * args.steps / (ms * 1e-3)
    # end to end: pinned host frames in, keypoints + descriptors out
    pin_in = torch.from_numpy(host).pin_memory()
    cap = ext.frame_capacity(w, h)
    pin_kp = torch.empty(n * cap * 28, dtype=torch.uint8).pin_memory()
    pin_desc = torch.empty(n * cap * 32, dtype=torch.uint8).pin_memory()
    cnt = np.zeros(n, np.int32)

    def step_e2e():
        ext.extract_batch_host_ptr_into(pin_in.data_ptr(), n, h, w, w, w * h, pin_kp.data_ptr(), pin_desc.data_ptr(), cap, cnt,
                                        args.chunk_frames)

    step_e2e()
    _barrier(dist)
    tt = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    dt = _max_over_ranks(time.perf_counter() - tt, dist, local)
    _barrier(dist)
    e2e_fps = world * n * args.steps / dt
    same = int(cnt.sum()) == int(counts.sum())
    peak, peak_src = B.measured_peaks()
    b_alg, px = B.algorithmic_bytes_per_frame(w, h, cfg, n_kp)
    pipe = b_alg * n * args.steps / (ms * 1e-3) / 1e9
    out = {
        "workload": "%s: %d %s %dx%d frames per GPU per step, nFeatures=%d, %d levels, scale %.2f, FAST %d/%d" % (
            name, n, kind, w, h, cfg["nfeatures"], cfg["nlevels"], cfg["scaleFactor"], cfg["iniThFAST"], cfg["minThFAST"]),
        "value": fps, "unit": "frames/s", "n_gpus": world, "ms_per_step": ms / args.steps, "steps": args.steps,
        "frames_per_gpu": n, "frames_generated": min(generated, n), "keypoints_per_frame": n_kp, "stage_ms": stages,
        "l2": "%d MB of frames per step per GPU (L2 is 126 MB)" % (host.nbytes >> 20),
        "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": int(host.nbytes), "d2h_bytes_per_step": int(n * cap * 60 + 4 * n),
                "same_keypoint_count_as_resident_path": bool(same)},
        "roofline_pipeline": {"bound": "hbm", "achieved": pipe, "peak": peak, "unit": "GB/s", "frac": pipe / peak,
                              "algorithmic_bytes_per_frame": b_alg, "peak_source": peak_src},
        "gpu_launches": launches,
    }
    if rank == 0:
        k = 4
        idx = sorted(set(int(i) for i in np.linspace(0, n - 1, k)))
        res = {i: None for i in idx}
        step()
        ext.sync()
        for i in idx:
            kp = np.zeros(cap, _lib.KEYPOINT_DTYPE)
            de = np.zeros((cap, 32), np.uint8)
            m = __import__("ctypes").c_int(0)
            _lib.check(L.orbx_batch_read(ext._h, i, _lib.ptr(kp), _lib.ptr(de), cap, __import__("ctypes").byref(m)))
            res[i] = (kp[:m.value].copy(), de[:m.value].copy())
        sub = [host[i] for i in idx]
        out["parity"] = parity_spot_check(B, cfg, sub, [res[i] for i in idx], n_check=len(idx))
        if with_cpu:
            threads = os.cpu_count() or 1
            sample = host[: min(n, max(threads, 4 * threads if w * h < 10 ** 6 else threads))]
            cfps, done, kind_cpu = B.cpu_extract_fps(sample, cfg, threads, budget_s=max(4.0, args.cpu_seconds / 3))
            out["cpu_reference"] = {"value": cfps, "unit": "frames/s", "cores": threads, "kind": kind_cpu,
                                    "sample": "%d of the step's frames, one extractor object per host thread" % done}
    ext.close()
    del dev
    torch.cuda.empty_cache()
    return out


def bench_stereo(B, args, rank, world, local, dist, with_cpu, n_pairs):
    """c2: KITTI-shape pairs.  Both images of every pair are extracted as ONE batch (left = even frames, right = odd
    frames) and every pair then runs Frame::ComputeStereoMatches on the device (row search, 11x11 SAD refinement on
    the resident pyramids, parabola fit, median rejection).  Timed end to end from resident images to (mvuRight,
    mvDepth) on the host: the keypoints/descriptors of every frame are read back because the reference's Frame keeps
    them on the host, and orbx_stereo_matches takes them as host arrays."""
    import torch
    from orb_slam2_e_b200 import synth, _lib, extractor
    from orb_slam2_e_b200.bench_gpu import _barrier, _max_over_ranks
    L = _lib.load()
    w, h, cfg = 1241, 376, KITTI
    generated = min(n_pairs, 32)
    pairs = [synth.stereo_pair(h, w, 2000 + 131 * rank + i) for i in range(generated)]
    host = np.empty((2 * n_pairs, h, w), np.uint8)
    for i in range(n_pairs):
        k, b = divmod(i, generated)
        sh = ((k * 37) % h, (k * 101) % w)
        host[2 * i] = pairs[b][0] if k == 0 else np.roll(pairs[b][0], sh, axis=(0, 1))
        host[2 * i + 1] = pairs[b][1] if k == 0 else np.roll(pairs[b][1], sh, axis=(0, 1))
    # rows of 1241 bytes are not 16-byte multiples: keep the frames in a pitched device buffer like a caller would
    pitch = (w + 63) // 64 * 64
    dev = torch.zeros((2 * n_pairs, h, pitch), dtype=torch.uint8, device=f"cuda:{local}")
    dev[:, :, :w] = torch.from_numpy(host).cuda(local)
    n = 2 * n_pairs
    ext = _mk(cfg, local, w, h, n)
    mb = KITTI_BF / KITTI_FX
    cap = ext.frame_capacity(w, h)
    pin_kp = torch.empty(n * cap * 28, dtype=torch.uint8).pin_memory()
    pin_desc = torch.empty(n * cap * 32, dtype=torch.uint8).pin_memory()
    cnt = np.zeros(n, np.int32)
    kp_all = pin_kp.numpy().view(_lib.KEYPOINT_DTYPE).reshape(n, cap)
    de_all = pin_desc.numpy().reshape(n, cap, 32)

    ur_all = np.zeros((n_pairs, cap), np.float32)
    z_all = np.zeros((n_pairs, cap), np.float32)

    pin_cnt = torch.zeros(n, dtype=torch.int32).pin_memory()
    cnt = pin_cnt.numpy()

    def step():
        ext.extract_batch_device(dev.data_ptr(), n, h, w, pitch, pitch * h)
        # the Frame keeps keypoints and descriptors on the host: their download runs beside the stereo kernels
        ext.read_batch_begin(pin_kp.data_ptr(), pin_desc.data_ptr(), cap, pin_cnt.data_ptr())
        # mvuRight / mvDepth of every pair: three launches over the device-resident keypoints, one transfer
        _lib.check(L.orbx_stereo_matches_batch(ext._h, n_pairs, mb, KITTI_BF, _lib.ptr(ur_all), _lib.ptr(z_all), cap))
        ext.read_batch_end()
        return int((z_all > 0).sum())

    def step_per_pair():  # round-1 path: one orbx_stereo_matches call per pair on host keypoints
        ext.extract_batch_device(dev.data_ptr(), n, h, w, pitch, pitch * h)
        ext.read_batch_into(pin_kp.data_ptr(), pin_desc.data_ptr(), cap, cnt)
        matched = 0
        for i in range(n_pairs):
            nl, nr = int(cnt[2 * i]), int(cnt[2 * i + 1])
            ur, z = extractor.stereo_matches(ext, ext, kp_all[2 * i, :nl], de_all[2 * i, :nl], kp_all[2 * i + 1, :nr], de_all[2 * i + 1, :nr],
                                             mb, KITTI_BF, frame_left=2 * i, frame_right=2 * i + 1)
            matched += int((z > 0).sum())
        return matched

    matched_pp = step_per_pair()
    tpp = time.perf_counter()
    step_per_pair()
    pps_pp = world * n_pairs / (time.perf_counter() - tpp)
    for _ in range(3):
        matched = step()
    _barrier(dist)
    l0 = L.orb_launch_count()
    tt = time.perf_counter()
    for _ in range(args.steps):
        matched = step()
    torch.cuda.synchronize()
    dt = _max_over_ranks(time.perf_counter() - tt, dist, local)
    launches = int(L.orb_launch_count() - l0)
    _barrier(dist)
    # extraction alone, device resident (CUDA events)
    stream = torch.cuda.ExternalStream(ext.cuda_stream(), device=local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        ext.extract_batch_device(dev.data_ptr(), n, h, w, pitch, pitch * h)
    e1.record(stream)
    e1.synchronize()
    ms_ext = _max_over_ranks(e0.elapsed_time(e1), dist, local) / args.steps
    pps = world * n_pairs * args.steps / dt
    n_kp = float(cnt.mean())
    peak, peak_src = B.measured_peaks()
    b_frame, _ = B.algorithmic_bytes_per_frame(w, h, cfg, n_kp)
    b_pair = 2 * b_frame  # 2 x 2,541,578 + 4000 x 60 at 2000 keypoints per image
    out = {
        "workload": "c2: %d KITTI-shape stereo pairs (1241x376, nFeatures=2000, 8 levels, 1.2, FAST 20/7) per GPU per step: "
                    "extraction of both images + ComputeStereoMatches" % n_pairs,
        "value": pps, "unit": "pairs/s", "n_gpus": world, "ms_per_step": dt / args.steps * 1e3, "steps": args.steps,
        "pairs_per_gpu": n_pairs, "pairs_generated": generated, "keypoints_per_image": n_kp,
        "stereo_matches_per_pair": matched / n_pairs, "extraction_only_ms_per_step": ms_ext,
        "extraction_only_pairs_s": world * n_pairs / (ms_ext * 1e-3),
        "timing": "wall clock around extraction + orbx_stereo_matches_batch (device-resident keypoints) + D2H of keypoints/descriptors/mvuRight/mvDepth",
        "same_matches_as_per_pair_calls": bool(matched == matched_pp), "per_pair_calls_pairs_s": pps_pp,
        "roofline_pipeline": {"bound": "hbm", "achieved": b_pair * pps / world / 1e9, "peak": peak, "unit": "GB/s",
                              "frac": b_pair * pps / world / 1e9 / peak, "algorithmic_bytes_per_pair": b_pair, "peak_source": peak_src},
        "gpu_launches": launches,
    }
    if rank == 0:
        idx = sorted(set(int(i) for i in np.linspace(0, n - 1, 4)))
        res = [(kp_all[i, :cnt[i]].copy(), de_all[i, :cnt[i]].copy()) for i in idx]
        out["parity"] = parity_spot_check(B, cfg, [host[i] for i in idx], res, n_check=len(idx))
        if with_cpu:
            threads = os.cpu_count() or 1
            sample = host[: min(n, 2 * threads)]
            cfps, done, kind_cpu = B.cpu_extract_fps(sample, cfg, threads, budget_s=max(4.0, args.cpu_seconds / 3))
            out["cpu_reference"] = {"value": cfps / 2, "unit": "pairs/s (extraction of both images only)", "cores": threads, "kind": kind_cpu,
                                    "sample": "%d images of the step's pairs, one extractor object per host thread" % done}
    ext.close()
    del dev
    torch.cuda.empty_cache()
    return out


def bench_single(B, args, local):
    """c1_single: rank 0, one textured 640x480 frame through the C++ class."""
    sys.path.insert(0, os.path.join(B.ROOT, "tests"))
    import dropin_lib
    from orb_slam2_e_b200 import synth
    D = dropin_lib.load()
    os.environ.setdefault("ORBFRONT_DEVICE", str(local))
    cfg = B.TUM
    img = np.ascontiguousarray(synth.textured_frame(B.H, B.W, B.SEED * 1000))
    h = D.dropin_create(cfg["nfeatures"], cfg["scaleFactor"], cfg["nlevels"], cfg["iniThFAST"], cfg["minThFAST"])
    if not h:
        return {"error": D.dropin_error().decode()}
    try:
        with_pyr = D.dropin_time_extract(h, img.ctypes.data, B.W, B.H, img.strides[0], 200, 1)
        without = D.dropin_time_extract(h, img.ctypes.data, B.W, B.H, img.strides[0], 200, 0)
    finally:
        D.dropin_destroy(h)
    return {"workload": "c1_single: one textured 640x480 frame through the C++ ORBextractor::operator() (pageable host image in, "
                        "std::vector<cv::KeyPoint> + cv::Mat out), nFeatures=1000, 8 levels, 1.2, FAST 20/7",
            "us_per_call_with_pyramid_download": with_pyr, "us_per_call_without_pyramid_download": without,
            "frames_s_with_pyramid_download": 1e6 / with_pyr if with_pyr > 0 else None, "calls": 200, "statistic": "median"}
