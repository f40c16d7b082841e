"""Experiment: where does the end-to-end (host buffers) time go?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from orb_slam2_e_b200 import synth
from orb_slam2_e_b200.extractor import ORBextractor
n, h, w = 1024, 480, 640
host = synth.frame_batch("textured", n, h, w, 1000, distinct=32)
pin_in = torch.from_numpy(host).pin_memory()
dev = torch.empty_like(pin_in, device="cuda")
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3
print("H2D 315MB ms", t(lambda: dev.copy_(pin_in, non_blocking=True)))
ext = ORBextractor(1000, 1.2, 8, 20, 7, device=0, max_width=w, max_height=h, max_batch=n)
cap = ext.frame_capacity(w, h)
pin_kp = torch.empty(n * cap * 28, dtype=torch.uint8).pin_memory()
pin_desc = torch.empty(n * cap * 32, dtype=torch.uint8).pin_memory()
dk = torch.empty(n * cap * 60, dtype=torch.uint8, device="cuda")
pk = torch.empty(n * cap * 60, dtype=torch.uint8).pin_memory()
print("D2H %dMB ms" % (n*cap*60>>20), t(lambda: pk.copy_(dk, non_blocking=True)))
def dev_step():
    ext.extract_batch_device(dev.data_ptr(), n, h, w, w, w * h); ext.sync()
print("compute ms", t(dev_step))
cnt = np.zeros(n, np.int32)
for c in (0, 128, 256):
    print("pipelined chunk", c, "ms", t(lambda: ext.extract_batch_host_ptr_into(pin_in.data_ptr(), n, h, w, w, w * h, pin_kp.data_ptr(), pin_desc.data_ptr(), cap, cnt, c)))
def two_call():
    ext.extract_batch_host_ptr(pin_in.data_ptr(), n, h, w, w, w * h)
    ext.read_batch_into(pin_kp.data_ptr(), pin_desc.data_ptr(), ext.out_cap if False else cap, cnt)
print("two-call ms", t(two_call))
for nn in (64, 128, 256):
    def small():
        ext.extract_batch_device(dev.data_ptr(), nn, h, w, w, w * h); ext.sync()
    print("compute only n=%d ms per 1024: %.2f" % (nn, t(small) * 1024 / nn))
