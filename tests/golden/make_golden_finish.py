"""Generates tests/golden/frame_finish.npz IN THE BUILD CONTAINER:
  * cv2/<camera>      cv2.undistortPoints (OpenCV 4.13 wheel) on seeded points, the pin of
                      oracle/cvprims.hpp undistort_points;
  * ref/<case>/...    the reference's own Frame::UndistortKeyPoints / ComputeImageBounds /
                      ComputeStereoFromRGBD / AssignFeaturesToGrid (src/Frame.cc compiled unmodified into
                      oracle/_ref/libframeref.so; needs /root/reference) on seeded keypoints.
The inputs are regenerated from seeds by the tests.

    python tests/golden/make_golden_finish.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle_lib  # noqa: E402
from make_golden_finish_cases import CAMERAS, FINISH_CASES, depth_image, keypoints, random_points  # noqa: E402


def main():
    import cv2
    FR = oracle_lib.frame_ref_lib()
    assert FR is not None, "oracle/_ref/libframeref.so must be built (needs /root/reference)"
    out = {"cv2_version": np.frombuffer(cv2.__version__.encode(), np.uint8)}
    for name, (w, h, cam, dist, bf) in CAMERAS.items():
        if dist[0] == 0.0:
            continue
        K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
        D = np.array(dist, np.float32).reshape(-1, 1)
        p = random_points(name, 4000, 5)
        out["cv2/" + name] = cv2.undistortPoints(p.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    for case, (camname, n, seed, with_depth) in FINISH_CASES.items():
        x, y = keypoints(camname, n, seed)
        depth = depth_image(camname, seed) if with_depth else None
        r = oracle_lib.ref_frame_finish(FR, x, y, CAMERAS[camname], depth)
        for k, v in r.items():
            out["ref/%s/%s" % (case, k)] = v
        print(case, n, "keypoints,", r["cell_off"][-1], "in grid, bounds", r["bounds"][:4])
    path = os.path.join(HERE, "frame_finish.npz")
    np.savez_compressed(path, **out)
    print("frame_finish.npz", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
