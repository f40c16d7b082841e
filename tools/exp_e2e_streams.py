"""Experiment: end-to-end (host buffers) step time of orbx_extract_batch_host against the number of
compute streams (ORBX_COMPUTE_STREAMS, read once per process -> child processes) and the chunk size."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import os, sys, time
sys.path.insert(0, %r)
import numpy as np, torch
from orb_slam2_e_b200 import synth
from orb_slam2_e_b200.extractor import ORBextractor
n, h, w = 1024, 480, 640
host = synth.frame_batch("textured", n, h, w, 1000, distinct=32)
pin_in = torch.from_numpy(host).pin_memory()
ext = ORBextractor(1000, 1.2, 8, 20, 7, device=0, max_width=w, max_height=h, max_batch=n)
cap = ext.frame_capacity(w, h)
pin_kp = torch.empty(n * cap * 28, dtype=torch.uint8).pin_memory()
pin_desc = torch.empty(n * cap * 32, dtype=torch.uint8).pin_memory()
cnt = np.zeros(n, np.int32)
def t(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3
out = []
for c in (32, 64, 96, 128, 192):
    ms = t(lambda: ext.extract_batch_host_ptr_into(pin_in.data_ptr(), n, h, w, w, w * h, pin_kp.data_ptr(), pin_desc.data_ptr(), cap, cnt, c))
    out.append("chunk %%d: %%.3f ms (%%.0f k/s)" %% (c, ms, n / ms))
print("; ".join(out), " keypoints", int(cnt.sum()))
""" % ROOT
for k in (1, 2, 3, 4, 6):
    e = dict(os.environ); e["ORBX_COMPUTE_STREAMS"] = str(k)
    out = subprocess.run([sys.executable, "-c", CHILD], env=e, capture_output=True, text=True)
    print("streams", k, "->", out.stdout.strip() or out.stderr[-400:], flush=True)
