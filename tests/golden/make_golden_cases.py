"""Case table shared by make_golden.py (generator) and tests/test_golden.py."""
from orb_slam2_e_b200 import synth

EXTRACT_CASES = {  # name -> (generator, (h, w), seed, (nfeatures, scale, levels, ini, min))
    "tum_small": ("textured", (240, 320), 77, (500, 1.2, 4, 20, 7)),
    "tum_full": ("textured", (480, 640), 1, (1000, 1.2, 8, 20, 7)),
    "hamlyn_full": ("endoscopy", (480, 640), 11, (1200, 1.1, 6, 24, 7)),
    "kitti_left": ("textured", (376, 1241), 21, (2000, 1.2, 8, 20, 7)),
    "odd_small": ("textured", (123, 157), 5, (150, 1.3, 3, 15, 5)),
}


def frame(kind, shape, seed):
    return synth.textured_frame(*shape, seed) if kind == "textured" else synth.endoscopy_frame(*shape, seed)
