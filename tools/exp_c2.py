"""Experiment: where the time of the c2 leg (KITTI pairs: extraction + stereo matching + read-back) goes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from orb_slam2_e_b200 import synth, _lib
from orb_slam2_e_b200.extractor import ORBextractor
L = _lib.load()
w, h, n_pairs = 1241, 376, 64
pairs = [synth.stereo_pair(h, w, 2000 + i) for i in range(16)]
host = np.empty((2 * n_pairs, h, w), np.uint8)
for i in range(n_pairs):
    host[2 * i], host[2 * i + 1] = pairs[i % 16]
pitch = (w + 63) // 64 * 64
dev = torch.zeros((2 * n_pairs, h, pitch), dtype=torch.uint8, device="cuda")
dev[:, :, :w] = torch.from_numpy(host).cuda()
n = 2 * n_pairs
ext = ORBextractor(2000, 1.2, 8, 20, 7, device=0, max_width=w, max_height=h, max_batch=n)
cap = ext.frame_capacity(w, h)
pin_kp = torch.empty(n * cap * 28, dtype=torch.uint8).pin_memory()
pin_desc = torch.empty(n * cap * 32, dtype=torch.uint8).pin_memory()
cnt = np.zeros(n, np.int32)
ur = np.zeros((n_pairs, cap), np.float32); z = np.zeros((n_pairs, cap), np.float32)
mb, bf = 386.1448 / 718.856, 386.1448
def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3
def ex(): ext.extract_batch_device(dev.data_ptr(), n, h, w, pitch, pitch * h); ext.sync()
def st(): _lib.check(L.orbx_stereo_matches_batch(ext._h, n_pairs, mb, bf, _lib.ptr(ur), _lib.ptr(z), cap))
def rd(): ext.read_batch_into(pin_kp.data_ptr(), pin_desc.data_ptr(), cap, cnt)
def all3():
    ext.extract_batch_device(dev.data_ptr(), n, h, w, pitch, pitch * h); st(); rd()
print("extract %.3f ms, stereo batch %.3f ms, read-back %.3f ms, all three %.3f ms (%d pairs, cap %d)" % (t(ex), t(st), t(rd), t(all3), n_pairs, cap))
ext.set_stage_timing(True); ex(); print(ext.last_stage_ms())
