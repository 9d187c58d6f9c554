"""Shared helpers of the test-suite: seeded synthetic correspondences and the host-harness binding."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

FOCAL = 642.9165664800531
FY = 641.9171825800378
PP = (460.1840658156501, 308.5846449100310)


def rodrigues(r):
    r = np.asarray(r, dtype=np.float64)
    th = np.linalg.norm(r)
    if th < 1e-15:
        return np.eye(3)
    k = r / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.cos(th) * np.eye(3) + (1 - np.cos(th)) * np.outer(k, k) + np.sin(th) * Kx


def two_view(n, seed, outliers=0.3, noise=0.5, rvec=(0.01, 0.05, -0.02), t=(0.1, -0.02, 0.5)):
    """n pixel correspondences of a random scene seen from two poses (+ noise, + uniform outliers) →
    p0, p1 float32 [n, 2], R, t, X (scene points in camera 0)."""
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(-8, 8, n), rng.uniform(-3, 3, n), rng.uniform(4, 40, n)], 1)
    R = rodrigues(rvec)
    t = np.asarray(t, dtype=np.float64)
    p0 = np.stack([FOCAL * X[:, 0] / X[:, 2] + PP[0], FY * X[:, 1] / X[:, 2] + PP[1]], 1)
    Y = X @ R.T + t
    p1 = np.stack([FOCAL * Y[:, 0] / Y[:, 2] + PP[0], FY * Y[:, 1] / Y[:, 2] + PP[1]], 1)
    p0 += rng.normal(0, noise, p0.shape)
    p1 += rng.normal(0, noise, p1.shape)
    k = int(outliers * n)
    if k:
        idx = rng.choice(n, k, replace=False)
        p1[idx] = np.stack([rng.uniform(0, 960, k), rng.uniform(0, 600, k)], 1)
    return p0.astype(np.float32), p1.astype(np.float32), R, t, X


def pnp_scene(n, seed, outliers=0.3, noise=0.5, rvec=(0.01, 0.03, -0.01), t=(0.02, -0.01, -0.13)):
    """3D points (float32, camera-0 frame) and their noisy pixels in camera 1 (+ uniform outliers)."""
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(-8, 8, n), rng.uniform(-3, 3, n), rng.uniform(4, 40, n)], 1).astype(np.float32)
    R = rodrigues(rvec)
    t = np.asarray(t, dtype=np.float64)
    Y = X.astype(np.float64) @ R.T + t
    p = np.stack([FOCAL * Y[:, 0] / Y[:, 2] + PP[0], FY * Y[:, 1] / Y[:, 2] + PP[1]], 1)
    p += rng.normal(0, noise, p.shape)
    k = int(outliers * n)
    if k:
        idx = rng.choice(n, k, replace=False)
        p[idx] = np.stack([rng.uniform(0, 960, k), rng.uniform(0, 600, k)], 1)
    return X, p.astype(np.float32), R, t


def normalise(p, focal=FOCAL, pp=PP):
    q = np.empty(p.shape, dtype=np.float64)
    q[:, 0] = (p[:, 0].astype(np.float64) - pp[0]) / focal
    q[:, 1] = (p[:, 1].astype(np.float64) - pp[1]) / focal
    return np.ascontiguousarray(q)


def binary_descriptors(n0, n1, seed, shared=0.6, flip=0.06):
    """Two sets of 256-bit descriptors sharing a fraction of (bit-flipped) landmarks, shuffled."""
    rng = np.random.default_rng(seed)
    ns = int(min(n0, n1) * shared)
    base = rng.integers(0, 256, (ns, 32), dtype=np.uint8)

    def obs(n):
        d = np.empty((n, 32), np.uint8)
        flips = np.packbits(rng.random((ns, 256)) < flip, axis=1)
        d[:ns] = base ^ flips
        d[ns:] = rng.integers(0, 256, (n - ns, 32), dtype=np.uint8)
        return np.ascontiguousarray(d[rng.permutation(n)])
    return obs(n0), obs(n1)


def float_descriptors(n0, n1, seed, shared=0.6, sigma=0.05):
    """Two sets of unit 256-d descriptors, SuperPoint layout [256, N]."""
    rng = np.random.default_rng(seed)
    ns = int(min(n0, n1) * shared)
    base = rng.standard_normal((ns, 256))

    def obs(n):
        d = np.empty((n, 256))
        d[:ns] = base / np.linalg.norm(base, axis=1, keepdims=True) + rng.normal(0, sigma, (ns, 256))
        d[ns:] = rng.standard_normal((n - ns, 256))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        return np.ascontiguousarray(d[rng.permutation(n)].T.astype(np.float32))
    return obs(n0), obs(n1)


def sinkhorn_scores(m, n, seed, gain=20.0, shared=0.6):
    """A SuperGlue-like score matrix [m, n] f32: gain * cosine of noisy shared descriptors (the reference's scores are
    dot products of unnormalised 256-d projections divided by 16, superglue.py:259-260: O(10) for matches, O(1) else)."""
    d0, d1 = float_descriptors(m, n, seed, shared=shared)
    return np.ascontiguousarray((gain * (d0.T.astype(np.float64) @ d1.astype(np.float64))).astype(np.float32))


_hh = None


def host_harness():
    """TEST-ONLY host build of csrc/*.cuh (see tests/host_harness/host_harness.cu)."""
    global _hh
    if _hh is None:
        import __graft_entry__ as ge
        _hh = C.CDLL(ge.build_host_harness())
        _hh.hh_five_point.restype = C.c_int
        _hh.hh_update_num_iters.restype = C.c_int
        _hh.hh_update_num_iters.argtypes = [C.c_double, C.c_double, C.c_int, C.c_int]
    return _hh


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


_orb = None


def orb_oracle():
    """TEST-ONLY: oracle/orb_guided.c (the reference's ORB-SLAM3 guided matchers restated in C) as a ctypes library."""
    global _orb
    if _orb is None:
        import __graft_entry__ as ge
        ge._build_oracle()
        _orb = C.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_build",
                                   "liboracle.so"))
        _orb.orb_search_for_initialization.restype = C.c_int
        _orb.orb_search_for_initialization.argtypes = [C.c_int] + [C.c_void_p] * 4 + [C.c_int] + [C.c_void_p] * 4 + [
            C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p]
        _orb.orb_compute_stereo_matches.restype = None
        _orb.orb_compute_stereo_matches.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                                    C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        _orb.orb_descriptor_distance.restype = C.c_int
        _orb.orb_descriptor_distance.argtypes = [C.c_void_p, C.c_void_p]
    return _orb


def orb_search_init_oracle(kp1, ang1, oct1, d1, kp2, ang2, oct2, d2, prev, window=100, ratio=0.9, check=True,
                           bounds=(0.0, 960.0, 0.0, 600.0)):
    """-> (matches12 [n1] int32, n_matches, prev_matched updated copy)"""
    lib = orb_oracle()
    c = lambda a, dt: np.ascontiguousarray(a, dtype=dt)
    kp1, kp2, prev = c(kp1, np.float32), c(kp2, np.float32), c(prev, np.float32).copy()
    ang1, ang2 = c(ang1, np.float32), c(ang2, np.float32)
    oct1, oct2, d1, d2 = c(oct1, np.int32), c(oct2, np.int32), c(d1, np.uint8), c(d2, np.uint8)
    m12 = np.empty(kp1.shape[0], np.int32)
    n = lib.orb_search_for_initialization(kp1.shape[0], vp(kp1), vp(ang1), vp(oct1), vp(d1), kp2.shape[0], vp(kp2),
                                          vp(ang2), vp(oct2), vp(d2), vp(prev), int(window), float(ratio), int(check),
                                          bounds[0], bounds[1], bounds[2], bounds[3], vp(m12))
    return m12, n, prev


def orb_stereo_oracle(kpL, octL, dL, kpR, octR, dR, imL, imR, mb, mbf):
    """-> (u_right, depth, sad)"""
    lib = orb_oracle()
    c = lambda a, dt: np.ascontiguousarray(a, dtype=dt)
    kpL, kpR = c(kpL, np.float32), c(kpR, np.float32)
    octL, octR, dL, dR = c(octL, np.int32), c(octR, np.int32), c(dL, np.uint8), c(dR, np.uint8)
    imL, imR = c(imL, np.uint8), c(imR, np.uint8)
    n = kpL.shape[0]
    ur, dp, sad = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.int32)
    lib.orb_compute_stereo_matches(n, vp(kpL), vp(octL), vp(dL), kpR.shape[0], vp(kpR), vp(octR), vp(dR), vp(imL),
                                   vp(imR), imL.shape[0], imL.shape[1], float(mb), float(mbf), vp(ur), vp(dp), vp(sad))
    return ur, dp, sad


def stereo_scene(n_kpts=1200, seed=0, n_trunks=60, border=16, pose_index=3):
    """A rendered stereo frame for the ORB stereo tests: world with trunk geometry, left/right keypoints + descriptors
    observed from the rectified rig, and the two poses.  (Images come from the renderer under test.)"""
    from forest_slam_b200 import synth
    world = synth.make_world(synth.SEED0 + seed, n_trunks=n_trunks, n_ground=6000, n_canopy=500)
    T_l = synth.make_trajectory(pose_index + 1, synth.SEED0 + seed)[pose_index]
    T_r = synth.right_pose(T_l)
    fl = synth.observe(world, T_l, n_kpts, 7000 + seed, border=border)
    fr = synth.observe(world, T_r, n_kpts, 8000 + seed, border=border)
    return world, T_l, T_r, fl, fr
