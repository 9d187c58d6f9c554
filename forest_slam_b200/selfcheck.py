"""smoke(): one small invocation of the hot path on cuda:0, checked against the oracle.

This is the ONLY module of the package that touches ``oracle/`` (as the checker, never as the thing run):
``__graft_entry__.smoke()`` calls it.  The product path (api.py / sequence.py) never imports it.
"""
from __future__ import annotations

import os
import sys

import numpy as np


def smoke(verbose: bool = True) -> None:
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("smoke() needs cuda:0 — forest_slam_b200 has no CPU path")
    torch.cuda.set_device(0)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    from oracle import cv_literal as lit  # checker only
    from . import api, synth

    world = synth.make_world(n_trunks=60, n_ground=1500, n_canopy=1500)
    frames = synth.make_sequence(2, 512, world=world)
    f0, f1 = frames

    # ---- mono: float match -> gather -> E-RANSAC -> recoverPose  (mono_slam.py:100-112)
    got = api.match_mutual_nn(f0.desc_f32, f1.desc_f32, threshold=0.2)
    ref = lit.float_mutual_nn(f0.desc_f32, f1.desc_f32, threshold=0.2)
    keep = ~ref["ambiguous0"]
    assert np.array_equal(got["matches0"][keep], ref["matches0"][keep]), "float match parity"
    mk0, mk1 = lit.gather_matches(f0.kpts, f1.kpts, got["matches0"])
    E, mask = api.findEssentialMat(mk0, mk1, focal=lit.K0[0, 0], pp=(lit.K0[0, 2], lit.K0[1, 2]), method=api.RANSAC,
                                   prob=0.999, threshold=1.0)
    E_cv, mask_cv = lit.find_essential(mk0, mk1)
    assert np.array_equal(mask, mask_cv), "findEssentialMat inlier mask parity"
    n, R, t, pm = api.recoverPose(E, mk0, mk1, focal=lit.K0[0, 0], pp=(lit.K0[0, 2], lit.K0[1, 2]))
    n_cv, R_cv, t_cv, pm_cv = lit.recover_pose(E_cv, mk0, mk1)
    assert n == n_cv and np.array_equal(pm, pm_cv), "recoverPose mask parity"
    assert lit.angle_error_mat(R, R_cv) < 1e-4 and lit.angle_error_vec(t.ravel(), t_cv.ravel()) < 1e-4

    # ---- stereo: Hamming match -> back-projection -> PnP RANSAC  (stereo_slam.py:234-246, :264-298)
    ms = api.BFMatcher(api.NORM_HAMMING, crossCheck=True).match(f0.desc_bin, f1.desc_bin)
    q, tr, d = lit.bf_hamming_match(f0.desc_bin, f1.desc_bin)
    assert [m.queryIdx for m in ms] == q.tolist() and [m.trainIdx for m in ms] == tr.tolist(), "Hamming parity"
    qi = np.array([m.queryIdx for m in ms])
    ti = np.array([m.trainIdx for m in ms])
    p3, valid = api.backproject_stereo(f0.kpts[qi], f0.disparity[qi])
    p3_ref, valid_ref = lit.backproject(f0.kpts[qi], f0.disparity[qi])
    assert np.array_equal(valid.cpu().numpy().astype(bool), valid_ref), "depth gate parity"
    assert np.allclose(p3.cpu().numpy(), p3_ref, rtol=1e-3), "back-projection parity"
    P3 = p3_ref[valid_ref]
    p2 = f1.kpts[ti][valid_ref]
    ok, rvec, tvec, inl = api.solvePnPRansac(P3, p2, lit.K0, lit.DIST_L, reprojectionError=1.0, confidence=0.99,
                                             iterationsCount=1000, flags=api.SOLVEPNP_ITERATIVE)
    ok_cv, rvec_cv, tvec_cv, inl_cv = lit.solve_pnp_ransac(P3, p2)
    assert ok == ok_cv and np.array_equal(inl, inl_cv), "solvePnPRansac inlier parity"
    assert lit.angle_error_mat(api.Rodrigues(rvec)[0], api.Rodrigues(rvec_cv)[0]) < 1e-4
    assert np.linalg.norm(tvec - tvec_cv) < 1e-4 * max(1.0, np.linalg.norm(tvec_cv))
    if verbose:
        print(f"smoke ok: {len(mk0)} float matches / {int(mask.sum())} E-inliers / {n} cheirality-good; "
              f"{len(ms)} Hamming matches / {len(inl)} PnP inliers")
