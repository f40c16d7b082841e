"""The keypoint selection has two kernels (quadtree.cu): the sort-free one, and the sorting one that redoes the (frame, level)
pairs the first hands back -- immediately (sparse levels) or midway (the reference's loop would need a deeper grid).
Both must give the reference's kept set in the reference's ORDER.  The library reads its knobs once per process, so every
configuration runs the fuzz cases in a process of its own:
  ORBX_QP_RATIO4=0     the sort-free kernel tries every level, so sparse and clustered levels take the midway hand-backs
  ORBX_QT_NO_PYRAMID=1 the sorting kernel alone
(the default configuration is what every other extractor test runs)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
SCRIPT = r"""
import sys
sys.path.insert(0, %r)
import numpy as np
import test_gpu_fuzz as F
import test_gpu_extract as E
from orb_slam2_e_b200 import synth
n = 0
for seed in range(0, 36, 2):
    F.test_random_case(seed)
    n += 1
for seed in (3, 4):
    E.compare(synth.endoscopy_frame(480, 640, 700 + seed), E.HAMLYN)   # sparse levels: 1.5-1.8 candidates per wanted keypoint
    E.compare(synth.textured_frame(480, 640, 800 + seed), E.TUM)
    n += 2
# clustered corners: a few dense blobs on a flat background push the tree deep
rng = np.random.default_rng(77)
img = np.full((360, 480), 90, np.uint8)
for _ in range(12):
    y, x = int(rng.integers(40, 300)), int(rng.integers(40, 420))
    img[y:y + 24, x:x + 24] = rng.integers(0, 256, (24, 24), dtype=np.uint8)
F.check(img, (600, 1.2, 4, 20, 7))
print("ok", n + 1)
""" % HERE


@pytest.mark.parametrize("env", [{"ORBX_QP_RATIO4": "0"}, {"ORBX_QT_NO_PYRAMID": "1"}], ids=["try_every_level", "sorting_kernel_only"])
def test_selection_paths_agree_with_the_reference(env):
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-c", SCRIPT], env=e, capture_output=True, text=True, timeout=900, cwd=os.path.dirname(HERE))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "ok" in r.stdout
