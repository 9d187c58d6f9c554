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
