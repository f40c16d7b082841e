"""Experiment: where the end-to-end (host buffers) step time of orbx_extract_batch_host goes.
Child processes set the developer knobs ORBX_E2E_SKIP (1 no upload, 2 no download, 4 no kernels),
ORBX_CHUNK_SCHEDULE ("16,48,64": chunk sizes, last repeats) and ORBX_COMPUTE_STREAMS."""
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
fn = lambda: ext.extract_batch_host_ptr_into(pin_in.data_ptr(), n, h, w, w, w * h, pin_kp.data_ptr(), pin_desc.data_ptr(), cap, cnt, 0)
for _ in range(4): fn()
torch.cuda.synchronize()
ts = []
for _ in range(12):
    t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
print("min %%.3f med %%.3f ms" %% (min(ts), sorted(ts)[len(ts) // 2]))
""" % ROOT
cases = [("default", {})]
for skip, name in ((6, "upload only"), (5, "download only"), (3, "kernels only (chunked, resident)"), (2, "upload + kernels"), (1, "kernels + download"), (4, "upload + download")):
    cases.append((name, {"ORBX_E2E_SKIP": str(skip)}))
for sched in ("16,48,64", "32,64", "16,32,64", "64,128", "32,96,128", "128", "16,48,96"):
    cases.append(("schedule " + sched, {"ORBX_CHUNK_SCHEDULE": sched}))
for k in (2, 3, 6, 8):
    cases.append(("streams %d" % k, {"ORBX_COMPUTE_STREAMS": str(k)}))
for sched, k in (("16,48,64", 6), ("32,96,128", 6), ("32,96,128", 8), ("64,128", 8)):
    cases.append(("schedule %s streams %d" % (sched, k), {"ORBX_CHUNK_SCHEDULE": sched, "ORBX_COMPUTE_STREAMS": str(k)}))
if len(sys.argv) > 2 and sys.argv[1] == "--sched":  # explicit schedules, "64*14,32*2,16*4" style
    def expand(spec):
        out = []
        for part in spec.split(","):
            size, _, rep = part.partition("*")
            out += [size] * int(rep or 1)
        return ",".join(out)
    cases = [("schedule " + a, {"ORBX_CHUNK_SCHEDULE": expand(a)}) for a in sys.argv[2:]]
elif len(sys.argv) > 1:
    cases = [c for c in cases if any(a in c[0] for a in sys.argv[1:])]
for name, env in cases:
    e = dict(os.environ); e.update(env)
    out = subprocess.run([sys.executable, "-c", CHILD], env=e, capture_output=True, text=True)
    print("%-40s %s" % (name, out.stdout.strip() or out.stderr[-300:]), flush=True)
    if "ORBX_E2E_TRACE" in e and out.stdout.strip():
        print(out.stderr, flush=True)
