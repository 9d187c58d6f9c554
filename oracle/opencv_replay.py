"""ORACLE (test infrastructure only — never imported by the product path).

O-replay: NumPy / pure-Python restatement of the OpenCV algorithms behind the reference's calls
    cv2.findEssentialMat(..., RANSAC)      ros_ws/src/mono_slam.py:111
    cv2.recoverPose                        ros_ws/src/mono_slam.py:112
    cv2.solvePnPRansac(..., ITERATIVE)     ros_ws/src/stereo_slam.py:294-295
The arithmetic lives in OpenCV (calib3d five-point.cpp / ptsetreg.cpp / solvepnp.cpp / epnp.cpp /
triangulate.cpp; core rand.cpp / lapack.cpp / mathfuncs_core), which is NOT vendored under /root/reference
(Dockerfile:62-67 clones an unpinned master).  This file restates the published algorithms; it is pinned
against the cv2 binary of this image (opencv-python-headless 4.13.0) by tests/test_oracle_cpu.py, stage
by stage (sampler, null-space basis, candidate order, masks).  The reference repo itself holds no test or
golden vector for this path: "parity unpinned" w.r.t. the reference's own tests.

The restatement exposes the intermediate stages (subset indices, candidate models, float32 residuals,
counts, adaptive iteration cap) so that every CUDA stage can be checked in isolation.
"""
from __future__ import annotations

import math

import numpy as np

# ------------------------------------------------------------------------------------------ cv::RNG
CV_RNG_COEFF = 4164903690
MASK64 = (1 << 64) - 1


class CvRNG:
    """cv::RNG — multiply-with-carry (core/operations.hpp RNG::next)."""

    def __init__(self, state: int = MASK64):
        self.state = state if state else 0xFFFFFFFF

    def next(self) -> int:
        self.state = ((self.state & 0xFFFFFFFF) * CV_RNG_COEFF + (self.state >> 32)) & MASK64
        return self.state & 0xFFFFFFFF

    def uniform(self, a: int, b: int) -> int:
        return a + self.next() % (b - a) if a != b else a


def ransac_subsets(count: int, n_iters: int, model_points: int = 5) -> np.ndarray:
    """RANSACPointSetRegistrator::getSubset (ptsetreg.cpp): RNG((uint64)-1) per run(); for each iteration
    draw uniform(0, count) until `model_points` distinct indices are collected (draw order kept)."""
    rng = CvRNG(MASK64)
    out = np.zeros((n_iters, model_points), dtype=np.int32)
    for it in range(n_iters):
        idx = []
        for i in range(model_points):
            v = rng.uniform(0, count)
            while v in idx:
                v = rng.uniform(0, count)
            idx.append(v)
        out[it] = idx
    return out


def ransac_update_num_iters(p: float, ep: float, model_points: int, max_iters: int) -> int:
    """RANSACUpdateNumIters (ptsetreg.cpp)."""
    p = max(p, 0.0)
    p = min(p, 1.0)
    ep = max(ep, 0.0)
    ep = min(ep, 1.0)
    num = max(1.0 - p, np.finfo(np.float64).tiny)
    denom = 1.0 - math.pow(1.0 - ep, model_points)
    if denom < np.finfo(np.float64).tiny:
        return 0
    num = math.log(num)
    denom = math.log(denom)
    return max_iters if (denom >= 0 or -num >= max_iters * (-denom)) else int(round(num / denom))


# ------------------------------------------------------------------------------------------ small linear algebra
_SVD_SIGNS = None


def _svd_fill_signs(m: int, count: int) -> np.ndarray:
    """The fixed pseudo-random +-1/m vectors JacobiSVDImpl_ (core/lapack.cpp) uses to complete a basis:
    RNG(0x12345678), sign = (next() & 256) ? + : -."""
    rng = CvRNG(0x12345678)
    out = np.zeros((count, m))
    for i in range(count):
        for k in range(m):
            out[i, k] = (1.0 / m) if (rng.next() & 256) else (-1.0 / m)
    return out


def nullspace_opencv(Q: np.ndarray) -> np.ndarray:
    """Rows 5..8 of Vt from SVD::compute(Q[5x9], FULL_UV): OpenCV completes the right singular basis by
    projecting fixed sign vectors onto the orthogonal complement of what it already has (two
    Gram-Schmidt passes) — so the null-space basis is a deterministic function of the row space."""
    n, m = Q.shape  # 5 x 9
    # any orthonormal basis of the row space gives the same projections (up to rounding)
    _, _, vt = np.linalg.svd(Q, full_matrices=False)
    rows = [vt[i].copy() for i in range(n)]
    signs = _svd_fill_signs(m, m - n)
    eps = np.finfo(np.float64).eps * 10
    for i in range(n, m):
        v = signs[i - n].copy()
        for _ in range(2):
            for j in range(i):
                sd = float(np.dot(v, rows[j]))
                v = v - sd * rows[j]
                asum = float(np.abs(v).sum())
                v = v * ((1.0 / asum) if asum > eps * 100 else 0.0)
        v = v / math.sqrt(float((v * v).sum()))
        rows.append(v)
    return np.array(rows[n:])


# monomial table of the 5-point constraint matrix (Nister ordering used by OpenCV's five-point.cpp):
# index of variable: 0=x 1=y 2=z 3=1
_MONO = [(0, 0, 0), (1, 1, 1), (0, 0, 1), (0, 1, 1), (0, 0, 2), (0, 0, 3), (1, 1, 2), (1, 1, 3), (0, 1, 2),
         (0, 1, 3), (0, 2, 2), (0, 2, 3), (0, 3, 3), (1, 2, 2), (1, 2, 3), (1, 3, 3), (2, 2, 2), (2, 2, 3),
         (2, 3, 3), (3, 3, 3)]
_MONO_INDEX = {m: i for i, m in enumerate(_MONO)}


def five_point_constraints(X: np.ndarray) -> np.ndarray:
    """10 x 20 coefficient matrix of det(E) = 0 and 2 E E^T E - tr(E E^T) E = 0 for E = x X0 + y X1 + z X2 + X3."""
    A = np.zeros((10, 20))
    for a in range(4):
        for b in range(4):
            XaXbT = X[a] @ X[b].T
            tr = np.trace(XaXbT)
            for c in range(4):
                col = _MONO_INDEX[tuple(sorted((a, b, c)))]
                M = 2.0 * (XaXbT @ X[c]) - tr * X[c]
                A[0:9, col] += M.reshape(9)
                A[9, col] += float(np.dot(X[a][0], np.cross(X[b][1], X[c][2])))
    return A


def solve_poly_dk(coeffs: np.ndarray, max_iters: int = 300) -> np.ndarray:
    """cv::solvePoly (core/mathfuncs.cpp): Durand-Kerner with Gauss-Seidel updates, start (1+i)^k,
    stop when the largest update is exactly 0 or after max_iters sweeps.  coeffs[k] multiplies z^k."""
    n = len(coeffs) - 1
    while n > 1 and abs(coeffs[n]) <= np.finfo(np.float64).eps:
        n -= 1
    c = [complex(coeffs[i], 0.0) for i in range(n + 1)]
    roots = []
    p = complex(1, 0)
    r = complex(1, 1)
    for i in range(n):
        roots.append(p)
        p = p * r
    for _ in range(max_iters):
        max_diff = 0.0
        for i in range(n):
            p = roots[i]
            num = c[n]
            denom = c[n]
            num_same = 1
            for j in range(n):
                num = num * p + c[n - j - 1]
                if j != i:
                    d = p - roots[j]
                    if d.real != 0 or d.imag != 0:
                        denom = denom * d
                    else:
                        num_same += 1
            num = num / denom
            if num_same > 1:  # multiple-root branch of OpenCV: not reached for generic data
                raise NotImplementedError("coincident Durand-Kerner iterates")
            roots[i] = p - num
            max_diff = max(max_diff, abs(num))
        if max_diff <= 0:
            break
    out = np.array(roots, dtype=np.complex128)
    out.imag[np.abs(out.imag) < 1e-100] = 0.0
    return out


def _poly_det3(B: np.ndarray) -> np.ndarray:
    """det of the 3x3 polynomial matrix whose row i is B[i] = [cubic(4) | cubic(4) | quartic(5)], highest
    power first inside every block; returns coefficients c[k] of z^k, k = 0..10."""
    def blk(i, j):
        s = B[i]
        return s[0:4] if j == 0 else (s[4:8] if j == 1 else s[8:13])
    pm = np.polymul
    det = (pm(blk(0, 0), np.polysub(pm(blk(1, 1), blk(2, 2)), pm(blk(1, 2), blk(2, 1))))
           - pm(blk(0, 1), np.polysub(pm(blk(1, 0), blk(2, 2)), pm(blk(1, 2), blk(2, 0))))
           + pm(blk(0, 2), np.polysub(pm(blk(1, 0), blk(2, 1)), pm(blk(1, 1), blk(2, 0)))))
    det = np.concatenate([np.zeros(11 - len(det)), det])
    return det[::-1].copy()


def null_vector_3x3(M: np.ndarray) -> np.ndarray:
    """SVD::solveZ: right singular vector of the smallest singular value."""
    _, _, vt = np.linalg.svd(M)
    return vt[-1]


def five_point(q1: np.ndarray, q2: np.ndarray):
    """EMEstimatorCallback::runKernel (calib3d/five-point.cpp) on 5 normalised correspondences.
    Returns the candidate essential matrices in solver (root) order, each 3x3, unit Frobenius norm."""
    n = q1.shape[0]
    # x2^T E x1 = 0 with e = vec(E) row-major: e[3i+j] multiplies x2_i * x1_j.  (Pinned against cv2 4.13:
    # this column order reproduces both the candidate order and the sign of every candidate.)
    Q = np.zeros((n, 9))
    Q[:, 0] = q2[:, 0] * q1[:, 0]
    Q[:, 1] = q2[:, 0] * q1[:, 1]
    Q[:, 2] = q2[:, 0]
    Q[:, 3] = q2[:, 1] * q1[:, 0]
    Q[:, 4] = q2[:, 1] * q1[:, 1]
    Q[:, 5] = q2[:, 1]
    Q[:, 6] = q1[:, 0]
    Q[:, 7] = q1[:, 1]
    Q[:, 8] = 1.0
    EE = nullspace_opencv(Q)  # 4 x 9
    X = EE.reshape(4, 3, 3)
    A = five_point_constraints(X)
    G = np.linalg.solve(A[:, :10], A[:, 10:])
    B = np.zeros((3, 13))
    for i in range(3):
        r1 = G[4 + 2 * i]
        r2 = G[5 + 2 * i]
        row1 = np.zeros(13)
        row2 = np.zeros(13)
        row1[1:4] = r1[0:3]
        row1[5:8] = r1[3:6]
        row1[9:13] = r1[6:10]
        row2[0:3] = r2[0:3]
        row2[4:7] = r2[3:6]
        row2[8:12] = r2[6:10]
        B[i] = row1 - row2
    c = _poly_det3(B)
    roots = solve_poly_dk(c)
    models = []
    for rt in roots:
        if abs(rt.imag) > 1e-10:
            continue
        z1 = rt.real
        z2 = z1 * z1
        z3 = z2 * z1
        z4 = z3 * z1
        bz = np.zeros((3, 3))
        for j in range(3):
            br = B[j]
            bz[j, 0] = br[0] * z3 + br[1] * z2 + br[2] * z1 + br[3]
            bz[j, 1] = br[4] * z3 + br[5] * z2 + br[6] * z1 + br[7]
            bz[j, 2] = br[8] * z4 + br[9] * z3 + br[10] * z2 + br[11] * z1 + br[12]
        xy1 = null_vector_3x3(bz)
        if abs(xy1[2]) < 1e-10:
            continue
        x = xy1[0] / xy1[2]
        y = xy1[1] / xy1[2]
        e = EE[0] * x + EE[1] * y + EE[2] * z1 + EE[3]
        e = e / math.sqrt(float((e * e).sum()))
        models.append(e.reshape(3, 3).copy())
    return models, {"EE": EE, "A": A, "B": B, "coeffs": c, "roots": roots}


def sampson_error_f32(E: np.ndarray, q1: np.ndarray, q2: np.ndarray) -> np.ndarray:
    """EMEstimatorCallback::computeError: double arithmetic, result stored as float32."""
    x1 = np.concatenate([q1, np.ones((q1.shape[0], 1))], axis=1)
    x2 = np.concatenate([q2, np.ones((q2.shape[0], 1))], axis=1)
    Ex1 = x1 @ E.T
    Etx2 = x2 @ E
    x2tEx1 = (x2 * Ex1).sum(axis=1)
    den = Ex1[:, 0] ** 2 + Ex1[:, 1] ** 2 + Etx2[:, 0] ** 2 + Etx2[:, 1] ** 2
    return (x2tEx1 * x2tEx1 / den).astype(np.float32)


def cv_normalise(x: np.ndarray, c: float, f: float) -> np.ndarray:
    """cv::findEssentialMat's `points.col(k) = (points.col(k) - c) / f` as OpenCV evaluates it: the MatExpr folds the
    division into scale = 1/f and shift = -c * (1/f), and convertTo's double kernel is a fused multiply-add.  Exact
    rational arithmetic + one rounding = fma (pinned by tests/test_oracle_cpu.py::test_normalisation_is_fma)."""
    from fractions import Fraction
    a = 1.0 / f
    b = -c * a
    fa, fb = Fraction(a), Fraction(b)
    return np.array([float(Fraction(float(v)) * fa + fb) for v in np.asarray(x, dtype=np.float32)], dtype=np.float64)


def find_essential_ransac(p0: np.ndarray, p1: np.ndarray, focal: float, pp, prob: float = 0.999,
                          threshold: float = 1.0, max_iters: int = 1000, solver=None):
    """cv::findEssentialMat(RANSAC) → dict(E, mask, iters, n_inliers, trace).  `solver` may replace the
    minimal solver (e.g. with cv2's own, to isolate the loop logic)."""
    solver = solver or (lambda a, b: five_point(a, b)[0])
    n = p0.shape[0]
    if n < 5:
        return None
    q1 = np.empty((n, 2))
    q2 = np.empty((n, 2))
    q1[:, 0] = cv_normalise(p0[:, 0], pp[0], focal)
    q1[:, 1] = cv_normalise(p0[:, 1], pp[1], focal)
    q2[:, 0] = cv_normalise(p1[:, 0], pp[0], focal)
    q2[:, 1] = cv_normalise(p1[:, 1], pp[1], focal)
    thr = threshold / ((focal + focal) / 2)
    thr2 = np.float32(thr * thr)
    subsets = ransac_subsets(n, max_iters) if n > 5 else np.arange(5, dtype=np.int32)[None]
    niters = max(max_iters, 1)
    best_count = 0
    best_E = None
    best_mask = None
    it = 0
    trace = []
    while it < niters:
        idx = subsets[it]
        models = solver(q1[idx], q2[idx])
        for k, E in enumerate(models):
            err = sampson_error_f32(E, q1, q2)
            mask = err <= thr2
            good = int(mask.sum())
            trace.append((it, k, good))
            if good > max(best_count, 4):
                best_count = good
                best_E = E.copy()
                best_mask = mask.copy()
                niters = ransac_update_num_iters(prob, (n - good) / n, 5, niters)
        it += 1
    if best_E is None:
        return {"E": None, "mask": None, "iters": it, "n_inliers": 0, "trace": trace}
    return {"E": best_E, "mask": best_mask.astype(np.uint8), "iters": it, "n_inliers": best_count,
            "trace": trace, "q1": q1, "q2": q2}
