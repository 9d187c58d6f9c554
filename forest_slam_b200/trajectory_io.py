"""Trajectory artefacts of the reference: accumulation, TUM files, APE / RPE.

mono_slam.py:115-139 / stereo_slam.py:301-338 accumulate T_cum = T_cum @ T_rel, convert the rotation to a quaternion
(tf.transformations.quaternion_from_matrix) and append `t x y z qx qy qz qw`; mono_slam.py:196 writes the list with
np.savetxt(..., fmt='%f').  The shipped files under ros_ws/src/pose_estimation_results/*/*.txt have that format.
The pose prefix product itself runs on the device (api.compose_trajectory, fs_compose_trajectory).
"""
from __future__ import annotations

import numpy as np


def quaternion_from_matrix(R: np.ndarray) -> np.ndarray:
    """(qx, qy, qz, qw) of a rotation matrix exactly as ROS tf.transformations.quaternion_from_matrix computes it (the
    call at mono_slam.py:127 / stereo_slam.py:327): homogeneous 4x4 with M[3,3] = 1, branch on trace(M) > M[3,3], else
    the largest diagonal element; every component scaled by 0.5 / sqrt(t * M[3,3]); NO sign normalisation — in the
    second branch qw can come out negative, and the reference writes it that way."""
    M = np.eye(4)
    M[:3, :3] = np.asarray(R, dtype=np.float64)[:3, :3]
    q = np.empty(4)
    t = np.trace(M)
    if t > M[3, 3]:
        q[3] = t
        q[2] = M[1, 0] - M[0, 1]
        q[1] = M[0, 2] - M[2, 0]
        q[0] = M[2, 1] - M[1, 2]
    else:
        i, j, k = 0, 1, 2
        if M[1, 1] > M[0, 0]:
            i, j, k = 1, 2, 0
        if M[2, 2] > M[i, i]:
            i, j, k = 2, 0, 1
        t = M[i, i] - (M[j, j] + M[k, k]) + M[3, 3]
        q[i] = t
        q[j] = M[i, j] + M[j, i]
        q[k] = M[k, i] + M[i, k]
        q[3] = M[k, j] - M[j, k]
    q *= 0.5 / np.sqrt(t * M[3, 3])
    return q


def matrix_from_quaternion(q: np.ndarray) -> np.ndarray:
    x, y, z, w = np.asarray(q, dtype=np.float64) / np.linalg.norm(q)
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def tum_rows(T_cum: np.ndarray, stamps: np.ndarray) -> np.ndarray:
    """[n, 4, 4] cumulative poses + [n] time stamps -> [n, 8] rows `t x y z qx qy qz qw` (mono_slam.py:127-139)."""
    T_cum = np.asarray(T_cum, dtype=np.float64)
    rows = np.empty((T_cum.shape[0], 8))
    rows[:, 0] = stamps
    rows[:, 1:4] = T_cum[:, :3, 3]
    for i in range(T_cum.shape[0]):
        rows[i, 4:8] = quaternion_from_matrix(T_cum[i])
    return rows


def write_tum(path: str, T_cum: np.ndarray, stamps: np.ndarray) -> None:
    np.savetxt(path, tum_rows(T_cum, stamps), fmt="%f")  # mono_slam.py:196


def read_tum(path: str):
    """-> stamps [n], T [n, 4, 4] of a TUM trajectory file (the reference's pose_estimation_results format)."""
    a = np.loadtxt(path, ndmin=2)
    T = np.tile(np.eye(4), (a.shape[0], 1, 1))
    T[:, :3, 3] = a[:, 1:4]
    for i in range(a.shape[0]):
        T[i, :3, :3] = matrix_from_quaternion(a[i, 4:8])
    return a[:, 0], T


def umeyama(src: np.ndarray, dst: np.ndarray, with_scale: bool = False):
    """Least-squares similarity (R, t, s) with dst ~ s R src + t (Umeyama 1991) — the alignment evo applies for APE."""
    mu_s, mu_d = src.mean(0), dst.mean(0)
    xs, xd = src - mu_s, dst - mu_d
    U, D, Vt = np.linalg.svd(xd.T @ xs / src.shape[0])
    S = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        S[2, 2] = -1
    R = U @ S @ Vt
    s = float(np.trace(np.diag(D) @ S) / (xs ** 2).sum() * src.shape[0]) if with_scale else 1.0
    return R, mu_d - s * R @ mu_s, s


def ape_translation(T_est: np.ndarray, T_ref: np.ndarray, align: bool = True, with_scale: bool = False) -> dict:
    """Absolute pose error of the translation part (evo_ape semantics): rmse / mean / median / max in metres."""
    p, q = T_est[:, :3, 3], T_ref[:, :3, 3]
    if align:
        R, t, s = umeyama(p, q, with_scale)
        p = (s * (R @ p.T)).T + t
    e = np.linalg.norm(p - q, axis=1)
    return {"rmse": float(np.sqrt((e ** 2).mean())), "mean": float(e.mean()), "median": float(np.median(e)),
            "max": float(e.max())}


def rpe_translation(T_est: np.ndarray, T_ref: np.ndarray, delta: int = 1) -> dict:
    """Relative pose error over `delta` frames (evo_rpe, translation part)."""
    e = []
    for i in range(T_est.shape[0] - delta):
        d_est = np.linalg.inv(T_est[i]) @ T_est[i + delta]
        d_ref = np.linalg.inv(T_ref[i]) @ T_ref[i + delta]
        err = np.linalg.inv(d_ref) @ d_est
        e.append(np.linalg.norm(err[:3, 3]))
    e = np.asarray(e)
    return {"rmse": float(np.sqrt((e ** 2).mean())), "mean": float(e.mean()), "max": float(e.max())}
