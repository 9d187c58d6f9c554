"""Generates tests/golden/*.npz: outputs of the reference's own cv2 calls (oracle/cv_literal.py = the call
expressions of mono_slam.py:106-112 and stereo_slam.py:85,234,264-298, issued with the reference's arguments)
on small seeded inputs.  cv2 = opencv-python-headless 4.13.0 in this image.  Run: python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cv2  # noqa: E402
import util  # noqa: E402
from oracle import cv_literal as lit  # noqa: E402


def main():
    out = {}
    # A2 Hamming crossCheck
    for i, (n0, n1) in enumerate([(300, 280), (64, 500), (257, 255)]):
        d0, d1 = util.binary_descriptors(n0, n1, 100 + i)
        q, t, d = lit.bf_hamming_match(d0, d1)
        out[f"ham{i}_d0"], out[f"ham{i}_d1"] = d0, d1
        out[f"ham{i}_q"], out[f"ham{i}_t"], out[f"ham{i}_dist"] = q, t, d
        out[f"ham{i}_ratio07"] = lit.bf_hamming_ratio(d0, d1, 0.7)
    # A4 / A5 mono
    for i, (n, o) in enumerate([(120, 0.3), (400, 0.5), (33, 0.1), (5, 0.0)]):
        p0, p1, _, _, _ = util.two_view(n, 200 + i, outliers=o)
        E, mask = lit.find_essential(p0, p1)
        E3 = E[:3]
        ng, R, t, pm = lit.recover_pose(E3, p0, p1)
        out[f"ess{i}_p0"], out[f"ess{i}_p1"] = p0, p1
        out[f"ess{i}_E"], out[f"ess{i}_mask"] = E, mask
        out[f"ess{i}_R"], out[f"ess{i}_t"], out[f"ess{i}_pmask"], out[f"ess{i}_ngood"] = R, t, pm, np.int32(ng)
    # A6 / A7 stereo
    for i, (n, o) in enumerate([(150, 0.3), (500, 0.45), (12, 0.0)]):
        P3, p2, _, _ = util.pnp_scene(n, 300 + i, outliers=o)
        ok, rvec, tvec, inl = lit.solve_pnp_ransac(P3, p2)
        out[f"pnp{i}_P3"], out[f"pnp{i}_p2"] = P3, p2
        out[f"pnp{i}_ok"], out[f"pnp{i}_rvec"], out[f"pnp{i}_tvec"] = np.bool_(ok), rvec, tvec
        out[f"pnp{i}_inliers"] = inl if inl is not None else np.zeros((0, 1), np.int32)
    rng = np.random.default_rng(400)
    kp = np.stack([rng.uniform(4, 955, 200), rng.uniform(4, 595, 200)], 1).astype(np.float32)
    disp = rng.uniform(-1, 90, 200).astype(np.float32)
    disp[::17] = 0.0
    disp[5::23] = -1.0
    disp[7::29] = 1e-4  # beyond the 1000 m gate
    p3, valid = lit.backproject(kp, disp)
    out["bp_kp"], out["bp_disp"], out["bp_p3"], out["bp_valid"] = kp, disp, p3, valid
    out["cv2_version"] = np.array(cv2.__version__)
    np.savez_compressed(os.path.join(HERE, "cv2_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "cv2_golden.npz"), len(out), "arrays")


if __name__ == "__main__":
    main()
