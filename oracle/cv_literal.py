"""ORACLE (test infrastructure only — never imported by the product path).

O-lit: the reference's own call expressions, issued literally with the reference's arguments, on NumPy inputs.
The reference scripts cannot be imported (ROS calls at import time, mono_slam.py:26, stereo_slam.py:26;
SuperGlue weights missing), so each function below is the line of the script it cites, nothing more.
cv2 here is opencv-python-headless 4.13.0 (the reference's Dockerfile:62-67 builds an unpinned master);
the reference holds no golden vectors for this path, so parity is pinned to THIS binary's outputs
("parity unpinned" w.r.t. the reference's own tests — SURVEY.md §8c).  This module is also the timed CPU
baseline of bench.py (`cpu_baseline` / `--impl reference`).
"""
from __future__ import annotations

import numpy as np
import cv2

# camera constants, stereo_slam.py:45-64 / mono_slam.py:40-50
K0 = np.array([[642.9165664800531, 0., 460.1840658156501],
               [0., 641.9171825800378, 308.5846449100310],
               [0., 0., 1.]])
DIST_L = np.array([-0.060164620903866, 0.094005180631043, 0.0, 0.0, 0])
BASELINE = 0.253736175410149  # np.linalg.norm(T_rgb0_rgb1[:3, 3]) for the 1x16 array of stereo_slam.py:61-64,270


# ---------------------------------------------------------------------------------------------- A2
def bf_hamming_match(descs0: np.ndarray, descs1: np.ndarray, cross_check: bool = True):
    """stereo_slam.py:85 + :234: cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(d0, d1).
    → (queryIdx[], trainIdx[], distance[]) sorted by query index."""
    bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=cross_check)
    if descs0.shape[0] == 0 or descs1.shape[0] == 0:
        return np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.float32)
    ms = bf.match(np.ascontiguousarray(descs0), np.ascontiguousarray(descs1))
    q = np.array([m.queryIdx for m in ms], dtype=np.int32)
    t = np.array([m.trainIdx for m in ms], dtype=np.int32)
    d = np.array([m.distance for m in ms], dtype=np.float32)
    o = np.argsort(q, kind="stable")
    return q[o], t[o], d[o]


def bf_hamming_ratio(descs0: np.ndarray, descs1: np.ndarray, ratio: float):
    """ORB_SLAM3/src/Frame.cc:1144-1151: knnMatch(k=2) then keep if d0 < d1 * ratio."""
    bf = cv2.BFMatcher(cv2.NORM_HAMMING)
    ms = bf.knnMatch(np.ascontiguousarray(descs0), np.ascontiguousarray(descs1), k=2)
    out = np.full(descs0.shape[0], -1, dtype=np.int32)
    r = float(np.float32(ratio))
    for pair in ms:
        if len(pair) >= 2 and float(pair[0].distance) < float(pair[1].distance) * r:
            out[pair[0].queryIdx] = pair[0].trainIdx
    return out


_POP8 = np.array([bin(i).count("1") for i in range(256)], dtype=np.int32)


def hamming_matrix(descs0: np.ndarray, descs1: np.ndarray) -> np.ndarray:
    """O-ham: popcount(d0[q] ^ d1[t]) over 256 bits (ORB_SLAM3/src/ORBmatcher.cc:2058-2074), NumPy only."""
    x = descs0[:, None, :] ^ descs1[None, :, :]
    return _POP8[x].sum(axis=2)


def hamming_mutual_numpy(descs0, descs1, cross_check=True, ratio=None):
    """Independent restatement of BFMatcher semantics (lowest index wins ties in both directions)."""
    n0, n1 = descs0.shape[0], descs1.shape[0]
    m0 = np.full(n0, -1, dtype=np.int32)
    if n0 == 0 or n1 == 0:
        return m0, np.zeros(n0, np.int32), np.zeros(n0, np.int32)
    D = hamming_matrix(descs0, descs1)
    best_t = D.argmin(axis=1)
    best_q = D.argmin(axis=0)
    d_best = D[np.arange(n0), best_t]
    Ds = np.sort(D, axis=1)
    d_second = Ds[:, 1] if n1 > 1 else np.full(n0, np.iinfo(np.int32).max)
    ok = np.ones(n0, bool)
    if cross_check:
        ok &= best_q[best_t] == np.arange(n0)
    if ratio:
        r = float(np.float32(ratio))
        ok &= (n1 > 1) & (d_best.astype(np.float64) < d_second.astype(np.float64) * r)
    m0[ok] = best_t[ok]
    return m0, d_best.astype(np.int32), d_second.astype(np.int32)


# ---------------------------------------------------------------------------------------------- A1
def float_mutual_nn(desc0: np.ndarray, desc1: np.ndarray, scale: float = 1.0, threshold=None, ratio=None,
                    margin: float = 1e-5):
    """superglue.py:259-260 (einsum 'bdn,bdm->bnm' * scale) + :268-278 (max over rows / columns, mutual check,
    threshold, -1 fill), evaluated in float64, plus the north star's Lowe ratio on the L2 distance.
    desc [256, N].  Also returns `ambiguous0`: rows whose decision is within `margin` of flipping (top-1 vs
    top-2 score gap, threshold gap, ratio gap, or the partner column's gap) — the float32 GPU result is only
    required to match on the other rows."""
    d0 = desc0.astype(np.float64)
    d1 = desc1.astype(np.float64)
    n0, n1 = d0.shape[1], d1.shape[1]
    m0 = np.full(n0, -1, dtype=np.int64)
    m1 = np.full(n1, -1, dtype=np.int64)
    if n0 == 0 or n1 == 0:
        return {"matches0": m0, "matches1": m1, "score0": np.zeros(n0), "ambiguous0": np.zeros(n0, bool)}
    S = d0.T @ d1
    i0 = S.argmax(axis=1)
    i1 = S.argmax(axis=0)
    b0 = S[np.arange(n0), i0]
    Ssort = np.sort(S, axis=1)
    s0 = Ssort[:, -2] if n1 > 1 else np.full(n0, -np.inf)
    Scs = np.sort(S, axis=0)
    colgap = (Scs[-1] - Scs[-2]) if n0 > 1 else np.full(n1, np.inf)
    mutual = i1[i0] == np.arange(n0)
    ok = mutual.copy()
    amb = (b0 - s0) < margin
    amb |= colgap[i0] < margin
    if threshold is not None:
        ok &= b0 * scale > threshold
        amb |= np.abs(b0 * scale - threshold) < margin
    if ratio:
        db = np.sqrt(np.maximum(0.0, 2.0 - 2.0 * b0))
        ds = np.sqrt(np.maximum(0.0, 2.0 - 2.0 * s0))
        r = float(np.float32(ratio))
        ok &= (n1 > 1) & (db < r * ds)
        amb |= np.abs(db - r * ds) < 10 * margin
    m0[ok] = i0[ok]
    m1[i0[ok]] = np.nonzero(ok)[0]
    return {"matches0": m0, "matches1": m1, "score0": b0 * scale, "second0": s0 * scale, "ambiguous0": amb}


# ---------------------------------------------------------------------------------------------- A3
def gather_matches(kpts0: np.ndarray, kpts1: np.ndarray, matches: np.ndarray):
    """mono_slam.py:106-108."""
    valid = matches > -1
    mkpts0 = kpts0[valid]
    mkpts1 = kpts1[matches[valid]]
    return mkpts0, mkpts1


# ---------------------------------------------------------------------------------------------- A4 / A5
def find_essential(mkpts0: np.ndarray, mkpts1: np.ndarray):
    """mono_slam.py:111."""
    return cv2.findEssentialMat(mkpts0, mkpts1, focal=K0[0, 0], pp=(K0[0, 2], K0[1, 2]), method=cv2.RANSAC,
                                prob=0.999, threshold=1.0)


def recover_pose(E: np.ndarray, mkpts0: np.ndarray, mkpts1: np.ndarray):
    """mono_slam.py:112."""
    return cv2.recoverPose(E, mkpts0, mkpts1, focal=K0[0, 0], pp=(K0[0, 2], K0[1, 2]))


def mono_pair(kpts0, kpts1, matches):
    """mono_slam.py:105-118 for one frame pair → dict(E, mask, R, t, T_rel)."""
    mkpts0, mkpts1 = gather_matches(kpts0, kpts1, matches)
    E, mask = find_essential(mkpts0, mkpts1)
    _, rotation, translation, pmask = recover_pose(E, mkpts0, mkpts1)
    translation = translation.reshape(3)
    T = np.eye(4)
    T[:3, 3] = translation
    T[:3, :3] = rotation
    return {"mkpts0": mkpts0, "mkpts1": mkpts1, "E": E, "mask": mask, "R": rotation, "t": translation,
            "pose_mask": pmask, "T_rel": T}


# ---------------------------------------------------------------------------------------------- A6
def backproject(mkpts0_l: np.ndarray, disparity, K=K0, baseline: float = BASELINE):
    """stereo_slam.py:117-121 + :264-289 with the dtype semantics of the reference container (NumPy 1.x
    value-based casting: float64 *scalars* do not promote the float32 arrays, so every array op is float32).
    `disparity` is either a dense [H, W] float32 map (looked up at int(y), int(x)) or one value per keypoint.
    → (points3D [K', 3] float32, valid_depth [K] bool)."""
    f32 = np.float32
    disp = np.asarray(disparity, dtype=f32).copy()
    disp[disp == 0.0] = 0.1
    disp[disp == -1.0] = 0.1
    cx, cy, fx, fy = f32(K[0, 2]), f32(K[1, 2]), f32(K[0, 0]), f32(K[1, 1])
    depth = f32(K[0, 0] * baseline) / disp
    X = mkpts0_l[:, 0].astype(f32)
    Y = mkpts0_l[:, 1].astype(f32)
    if depth.ndim == 2:
        Z = depth[Y.astype(int), X.astype(int)]
    else:
        Z = depth
    X = ((X - cx) / fx) * Z
    Y = ((Y - cy) / fy) * Z
    points3D = np.column_stack((X, Y, Z)).astype(f32)
    valid_depth = (Z > f32(0.1)) & (Z < f32(1000))
    return points3D, valid_depth


# ---------------------------------------------------------------------------------------------- A7 / A8
def solve_pnp_ransac(points3D: np.ndarray, mkpts1_l: np.ndarray, K=K0, dist=DIST_L):
    """stereo_slam.py:294-295."""
    return cv2.solvePnPRansac(points3D, mkpts1_l, K, dist, reprojectionError=1.0, confidence=0.99,
                              iterationsCount=1000, flags=cv2.SOLVEPNP_ITERATIVE)


def stereo_pair(mkpts0_l, mkpts1_l, disparity):
    """stereo_slam.py:264-306 for one frame pair (matches already gathered)."""
    points3D, valid = backproject(mkpts0_l, disparity)
    points3D = points3D[valid]
    mk1 = mkpts1_l[valid]
    out = {"points3D": points3D, "mkpts1": mk1, "ok": False}
    if len(points3D) >= 6:  # stereo_slam.py:292
        ok, rvec, tvec, inliers = solve_pnp_ransac(points3D, mk1)
        R, _ = cv2.Rodrigues(rvec)
        T = np.eye(4)
        T[:3, :3] = R
        T[:3, 3] = tvec.T
        out.update({"ok": bool(ok), "rvec": rvec, "tvec": tvec, "inliers": inliers, "R": R, "T_rel": T})
    return out


# ---------------------------------------------------------------------------------------------- error metrics
def angle_error_mat(R1, R2):
    """SuperGluePretrainedNetwork/models/utils.py:377-380, in radians."""
    cos = (np.trace(np.dot(R1.T, R2)) - 1) / 2
    cos = np.clip(cos, -1., 1.)
    return float(np.abs(np.arccos(cos)))


def angle_error_vec(v1, v2):
    """SuperGluePretrainedNetwork/models/utils.py:383-385, in radians."""
    n = np.linalg.norm(v1) * np.linalg.norm(v2)
    return float(np.arccos(np.clip(np.dot(v1, v2) / n, -1.0, 1.0)))


# ---------------------------------------------------------------------------------------------- timed CPU baseline
def superglue_match_tail(desc0: np.ndarray, desc1: np.ndarray, match_threshold: float = 0.2):
    """SuperGluePretrainedNetwork/models/superglue.py:259-260 and :268-278 issued literally with torch on the CPU
    (the neural steps between them — GNN, Sinkhorn — are out of scope and skipped): dense score contraction,
    row/column max, mutual check, exp(score) > threshold, -1 fill.  desc [256, N] float32 → matches0 [N0] int64."""
    import torch
    mdesc0 = torch.from_numpy(desc0)[None]
    mdesc1 = torch.from_numpy(desc1)[None]
    scores = torch.einsum('bdn,bdm->bnm', mdesc0, mdesc1)
    scores = scores / 256 ** .5
    max0, max1 = scores.max(2), scores.max(1)
    indices0, indices1 = max0.indices, max1.indices
    arange = torch.arange(indices0.shape[1])[None]
    mutual0 = arange == indices1.gather(1, indices0)
    zero = scores.new_tensor(0)
    mscores0 = torch.where(mutual0, max0.values.exp(), zero)
    valid0 = mutual0 & (mscores0 > match_threshold)
    indices0 = torch.where(valid0, indices0, indices0.new_tensor(-1))
    return indices0[0].numpy()


def mono_frame_pair(kpts0, desc0, kpts1, desc1):
    """One iteration of the mono_slam.py loop body restricted to the hot path (mono_slam.py:100-118)."""
    matches = superglue_match_tail(desc0, desc1)
    return mono_pair(kpts0, kpts1, matches)


def stereo_frame_pair(kpts0, descb0, disp0, kpts1, descb1):
    """One iteration of the stereo_slam.py loop body restricted to the hot path, ORB/BF variant
    (stereo_slam.py:232-246, :264-306); the per-keypoint disparity stands in for the SGBM map lookup."""
    q, t, _ = bf_hamming_match(descb0, descb1)
    return stereo_pair(kpts0[q], kpts1[t], disp0[q])
