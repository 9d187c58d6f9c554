"""Procedural forest + stereo rig: the synthetic stand-in for the BotanicGarden rosbags.

The reference reads 960x600 stereo images from a rosbag (mono_slam.py:37, stereo_slam.py:35) that is not
available offline; the hot path starts at keypoints + descriptors, so this module produces exactly those:
per frame a set of keypoints (pixel coordinates, float32), a 256-bit ORB-like and a 256-d SuperPoint-like
descriptor per keypoint, and the stereo disparity of every keypoint.  Everything is seeded NumPy on the host;
it is workload generation, not part of the timed path.
"""
from __future__ import annotations

import dataclasses
from typing import List, Optional

import numpy as np

from .api import ForestCamera

SEED0 = 20240


@dataclasses.dataclass
class Frame:
    kpts: np.ndarray        # [N, 2] float32 pixels (x, y)
    desc_bin: np.ndarray    # [N, 32] uint8
    desc_f32: np.ndarray    # [256, N] float32, unit L2 columns (SuperPoint layout)
    landmark: np.ndarray    # [N] int64, -1 for spurious keypoints
    disparity: np.ndarray   # [N] float32, uL - uR of the keypoint (0 where unknown)
    T_wc: np.ndarray        # [4, 4] camera-to-world


@dataclasses.dataclass
class ForestWorld:
    points: np.ndarray      # [L, 3] world coordinates (x right, y down, z forward)
    desc_bin: np.ndarray    # [L, 32] uint8
    desc_f32: np.ndarray    # [L, 256] float32 unit rows
    priority: np.ndarray    # [L] fixed random priority (keeps the keypoint set stable between frames)
    trunks: Optional[np.ndarray] = None   # [n_trunks, 4] float64: axis x, axis z, radius, height (csrc/render.cu)
    ground_y: float = 1.2                 # world y of the ground plane (y points down; the rig is at y = 0)


def make_world(seed: int = SEED0, n_trunks: int = 400, pts_per_trunk: int = 60, n_ground: int = 8000,
               n_canopy: int = 8000, extent: float = 80.0, cam_height: float = 1.2) -> ForestWorld:
    rng = np.random.default_rng(seed)
    half = extent / 2
    # trunk positions: dart-throwing Poisson disc (min spacing 2 m)
    pos = []
    while len(pos) < n_trunks:
        c = np.array([rng.uniform(-half, half), rng.uniform(0.0, extent)])
        if all((c[0] - q[0]) ** 2 + (c[1] - q[1]) ** 2 > 4.0 for q in pos):
            pos.append(c)
    pos = np.array(pos)
    radius = rng.uniform(0.1, 0.4, n_trunks)
    height = rng.uniform(4.0, 15.0, n_trunks)
    ang = rng.uniform(0, 2 * np.pi, (n_trunks, pts_per_trunk))
    h = rng.uniform(0, 1, (n_trunks, pts_per_trunk)) * height[:, None]
    tx = pos[:, 0:1] + radius[:, None] * np.cos(ang)
    tz = pos[:, 1:2] + radius[:, None] * np.sin(ang)
    ty = cam_height - h  # y points down, ground is at +cam_height below the camera
    trunks = np.stack([tx, ty, tz], axis=-1).reshape(-1, 3)
    ground = np.stack([rng.uniform(-half, half, n_ground), cam_height + rng.normal(0, 0.05, n_ground),
                       rng.uniform(0.0, extent, n_ground)], axis=-1)
    canopy = np.stack([rng.uniform(-half, half, n_canopy), cam_height - rng.uniform(5.0, 15.0, n_canopy),
                       rng.uniform(0.0, extent, n_canopy)], axis=-1)
    pts = np.concatenate([trunks, ground, canopy], axis=0)
    L = pts.shape[0]
    dbin = rng.integers(0, 256, (L, 32), dtype=np.uint8)
    df = rng.standard_normal((L, 256)).astype(np.float32)
    df /= np.linalg.norm(df, axis=1, keepdims=True)
    return ForestWorld(points=pts, desc_bin=dbin, desc_f32=df, priority=rng.random(L),
                       trunks=np.ascontiguousarray(np.stack([pos[:, 0], pos[:, 1], radius, height], axis=1)),
                       ground_y=float(cam_height))


def _rot_y(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])


def _rot_x(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]])


def _rot_z(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])


def make_trajectory(n_frames: int, seed: int = SEED0, step: float = 0.13, z0: float = 2.0) -> List[np.ndarray]:
    """Camera-to-world poses: forward 0.13 m/frame, yaw 0.01 sin(i/15) rad/frame, pitch/roll noise 0.002 rad."""
    rng = np.random.default_rng(seed + 7)
    poses = []
    yaw = 0.0
    p = np.array([0.0, 0.0, z0])
    for i in range(n_frames):
        R = _rot_y(yaw) @ _rot_x(rng.normal(0, 0.002)) @ _rot_z(rng.normal(0, 0.002))
        T = np.eye(4)
        T[:3, :3] = R
        T[:3, 3] = p
        poses.append(T)
        yaw += 0.01 * np.sin(i / 15.0)
        p = p + _rot_y(yaw) @ np.array([0.0, 0.0, step])
    return poses


def observe(world: ForestWorld, T_wc: np.ndarray, n_kpts: int, frame_seed: int,
            cam: ForestCamera = ForestCamera(), spurious: float = 0.25, pix_sigma: float = 0.5,
            bit_flip: float = 0.06, f32_sigma: float = 0.05, zmin: float = 1.0, zmax: float = 60.0,
            border: int = 4, pos_outlier: float = 0.0) -> Frame:
    rng = np.random.default_rng(frame_seed)
    T_cw = np.linalg.inv(T_wc)
    P = world.points @ T_cw[:3, :3].T + T_cw[:3, 3]
    z = P[:, 2]
    ok = (z > zmin) & (z < zmax)
    u = np.where(ok, cam.fx * P[:, 0] / np.where(ok, z, 1.0) + cam.cx, -1.0)
    v = np.where(ok, cam.fy * P[:, 1] / np.where(ok, z, 1.0) + cam.cy, -1.0)
    ok &= (u >= border) & (u < cam.width - border) & (v >= border) & (v < cam.height - border)
    vis = np.nonzero(ok)[0]
    n_true = min(int(round(n_kpts / (1.0 + spurious))), vis.size)
    sel = vis[np.argsort(world.priority[vis], kind="stable")[:n_true]]
    n_spur = n_kpts - n_true
    kp = np.stack([u[sel], v[sel]], axis=-1) + rng.normal(0, pix_sigma, (n_true, 2))
    if pos_outlier > 0:
        # gross mismatches: the descriptor still matches but the keypoint sits somewhere else (~20 % of the
        # matches of the RANSAC benchmarks are corrupted so that the outlier ratio is ~0.3)
        bad = rng.random(n_true) < pos_outlier
        kp[bad, 0] = rng.uniform(border, cam.width - border, int(bad.sum()))
        kp[bad, 1] = rng.uniform(border, cam.height - border, int(bad.sum()))
    kp[:, 0] = np.clip(kp[:, 0], 0.0, cam.width - 1.001)
    kp[:, 1] = np.clip(kp[:, 1], 0.0, cam.height - 1.001)
    disp = cam.fx * cam.baseline / z[sel] + rng.normal(0, 0.1, n_true)
    # descriptors of the observations
    flips = (rng.random((n_true, 256)) < bit_flip)
    dbin = world.desc_bin[sel] ^ np.packbits(flips, axis=1)
    df = world.desc_f32[sel] + rng.normal(0, f32_sigma, (n_true, 256)).astype(np.float32)
    df /= np.linalg.norm(df, axis=1, keepdims=True)
    # spurious keypoints: uniform position, fresh random descriptors
    kp_s = np.stack([rng.uniform(border, cam.width - border, n_spur), rng.uniform(border, cam.height - border, n_spur)],
                    axis=-1)
    dbin_s = rng.integers(0, 256, (n_spur, 32), dtype=np.uint8)
    df_s = rng.standard_normal((n_spur, 256)).astype(np.float32)
    df_s /= np.linalg.norm(df_s, axis=1, keepdims=True)
    disp_s = rng.uniform(2.0, 90.0, n_spur)
    kpts = np.concatenate([kp, kp_s]).astype(np.float32)
    desc_bin = np.concatenate([dbin, dbin_s])
    desc_f32 = np.concatenate([df, df_s]).astype(np.float32)
    lm = np.concatenate([sel, np.full(n_spur, -1, dtype=np.int64)])
    disparity = np.concatenate([disp, disp_s]).astype(np.float32)
    perm = rng.permutation(n_kpts)
    return Frame(kpts=np.ascontiguousarray(kpts[perm]), desc_bin=np.ascontiguousarray(desc_bin[perm]),
                 desc_f32=np.ascontiguousarray(desc_f32[perm].T), landmark=lm[perm],
                 disparity=np.ascontiguousarray(disparity[perm]), T_wc=T_wc)


def make_sequence(n_frames: int, n_kpts: int, seed: int = SEED0, world: Optional[ForestWorld] = None,
                  **kw) -> List[Frame]:
    world = world or make_world(seed)
    poses = make_trajectory(n_frames, seed)
    return [observe(world, poses[i], n_kpts, seed + i, **kw) for i in range(n_frames)]


def make_loop_trajectory(n_frames: int, seed: int = SEED0, laps: float = 2.0, radius: float = 25.0,
                         centre=(0.0, 40.0)) -> List[np.ndarray]:
    """Keyframe poses of a loop-closure job (BASELINE config 5): `laps` laps of a circle inside the forest, heading
    along the tangent, pitch/roll noise 0.002 rad — every place is revisited one lap later."""
    rng = np.random.default_rng(seed + 11)
    poses = []
    for i in range(n_frames):
        a = 2 * np.pi * laps * i / n_frames
        p = np.array([centre[0] + radius * np.sin(a), 0.0, centre[1] - radius * np.cos(a)])
        yaw = a + np.pi / 2      # tangent: d/da (sin a, -cos a) = (cos a, sin a) = rot_y(yaw) e_z with yaw = pi/2 - ... see below
        # rot_y(yaw) e_z = (sin yaw, 0, cos yaw); the tangent is (cos a, 0, sin a) = (sin(pi/2 - a), 0, cos(pi/2 - a))
        yaw = np.pi / 2 - a
        R = _rot_y(yaw) @ _rot_x(rng.normal(0, 0.002)) @ _rot_z(rng.normal(0, 0.002))
        T = np.eye(4)
        T[:3, :3] = R
        T[:3, 3] = p
        poses.append(T)
    return poses


def make_loop_sequence(n_frames: int, n_kpts: int, seed: int = SEED0, world: Optional[ForestWorld] = None,
                       **kw) -> List[Frame]:
    """`n_frames` keyframes along make_loop_trajectory (binary + float descriptors, as make_sequence)."""
    world = world or make_world(seed)
    poses = make_loop_trajectory(n_frames, seed)
    return [observe(world, poses[i], n_kpts, seed + 50000 + i, **kw) for i in range(n_frames)]


def right_pose(T_wc: np.ndarray, baseline: float = ForestCamera().baseline) -> np.ndarray:
    """Pose of the right camera of the rectified rig: the left camera shifted by the baseline along its own +x."""
    T = np.array(T_wc, dtype=np.float64)
    T[:3, 3] = T[:3, 3] + T[:3, :3] @ np.array([baseline, 0.0, 0.0])
    return T


def keypoint_angles(world_size: int, landmark: np.ndarray, frame_seed: int, sigma: float = 2.0) -> np.ndarray:
    """ORB-style keypoint orientation in degrees [0, 360): a fixed angle per landmark plus observation noise; spurious
    keypoints get a random one (the synthetic frames have no image patch to take an orientation from)."""
    base = np.random.default_rng(4242).uniform(0, 360, world_size)
    rng = np.random.default_rng(frame_seed + 99)
    a = np.where(landmark >= 0, base[np.maximum(landmark, 0)] + rng.normal(0, sigma, landmark.shape[0]),
                 rng.uniform(0, 360, landmark.shape[0]))
    return np.mod(a, 360.0).astype(np.float32)


def relative_pose(f0: Frame, f1: Frame):
    """(R, t) with X_cam1 = R X_cam0 + t — what recoverPose / solvePnP estimate for the pair (f0, f1)."""
    T = np.linalg.inv(f1.T_wc) @ f0.T_wc
    return T[:3, :3], T[:3, 3]


def corrupt_matches(mk1: np.ndarray, frac: float, seed: int, cam: ForestCamera = ForestCamera()) -> np.ndarray:
    """Replace a fraction of the partner points by uniform pixels (RANSAC-only benchmarks, configs 2 and 4)."""
    rng = np.random.default_rng(seed)
    out = mk1.copy()
    n = out.shape[0]
    k = int(frac * n)
    idx = rng.choice(n, k, replace=False)
    out[idx, 0] = rng.uniform(0, cam.width, k)
    out[idx, 1] = rng.uniform(0, cam.height, k)
    return out.astype(np.float32)


def stack_frames(frames: List[Frame]):
    """Frame bank arrays: kpts [F, N, 2], desc_bin [F, N, 32], desc_f32 [F, 256, N], disparity [F, N], counts [F]."""
    return (np.stack([f.kpts for f in frames]), np.stack([f.desc_bin for f in frames]),
            np.stack([f.desc_f32 for f in frames]), np.stack([f.disparity for f in frames]),
            np.array([f.kpts.shape[0] for f in frames], dtype=np.int32))
