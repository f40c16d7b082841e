"""Experiment: cycles per quadtree phase summed over every CTA of a 1024-frame batch, by level.
Needs a developer build: ORBFRONT_NVCC_FLAGS=-DORBX_QT_PROFILE2 python -m orb_slam2_e_b200.build --force"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from orb_slam2_e_b200 import synth, _lib
from orb_slam2_e_b200.extractor import ORBextractor
lib = C.CDLL(os.path.join(os.path.dirname(_lib.__file__), "liborbfront.so"))
names = ["attach", "radix", "bounds", "replay", "ordered", "leaf heads", "leaf walk", "leaf sort", "emit", "ctas", "keys"]
if os.environ.get("QP_NAMES"):  # the sort-free kernel's stamps
    names = ["init", "count+best", "pyramid", "replay", "ordered", "leaf heads", "leaf build", "leaf sort", "emit", "ctas", "keys"]
n, h, w = 1024, 480, 640
host = synth.frame_batch("textured", n, h, w, 1000, distinct=32)
dev = torch.from_numpy(host).cuda()
ext = ORBextractor(1000, 1.2, 8, 20, 7, device=0, max_width=w, max_height=h, max_batch=n)
out = (C.c_ulonglong * (16 * 12))()
for _ in range(2):
    ext.extract_batch_device(dev.data_ptr(), n, h, w, w, w * h); ext.sync()
lib.orbx_debug_qt_acc(out)
ext.extract_batch_device(dev.data_ptr(), n, h, w, w, w * h); ext.sync()
lib.orbx_debug_qt_acc(out)
a = np.array(list(out), dtype=np.float64).reshape(16, 12)
tot = a[:, :9].sum()
print("level  " + "  ".join("%9s" % x for x in names) + "   share")
for l in range(8):
    print("%5d  " % l + "  ".join("%9.0f" % (a[l, k] / max(a[l, 9], 1)) for k in range(9)) + "  %9.0f  %9.0f   %.1f%%" % (a[l, 9], a[l, 10] / max(a[l, 9], 1), 100 * a[l, :9].sum() / tot))
print("phase share: " + ", ".join("%s %.1f%%" % (names[k], 100 * a[:, k].sum() / tot) for k in range(9)))
if a[15, :4].sum() > 0: print("radix sub-phases (all CTAs, all passes): zero %.1f%%, histogram %.1f%%, scan %.1f%%, scatter %.1f%%" % tuple(100 * a[15, k] / a[15, :4].sum() for k in range(4)))
