"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement of the SuperGlue score head that follows the contraction:
  * sinkhorn()               <- log_sinkhorn_iterations   SuperGluePretrainedNetwork/models/superglue.py:143-150
  * optimal_transport()      <- log_optimal_transport     superglue.py:153-172
  * match_head()             <- the tail of SuperGlue.forward, superglue.py:268-285
Hyper-parameters in the reference: sinkhorn_iterations 20, match_threshold 0.2 (mono_slam.py:59-63), bin_score
initial value 1.0 (superglue.py:220).

Pinning: unlike the cv2 rows, this part of the reference IS importable in the build container (pure torch, no
weights needed), so tests/golden/make_sinkhorn_golden.py runs the reference's own log_optimal_transport + tail on
seeded inputs and commits the outputs (tests/golden/sinkhorn_golden.npz); tests/test_oracle_cpu.py checks this
restatement against them.  NumPy, float32 or float64 selectable: float32 mirrors the reference's arithmetic type,
float64 is used by the GPU tests to decide which rows are numerically ambiguous.
"""
from __future__ import annotations

import numpy as np


def _lse(a: np.ndarray, axis: int) -> np.ndarray:
    """log-sum-exp the way torch.logsumexp evaluates it: subtract the maximum, sum, log, add back."""
    mx = np.max(a, axis=axis, keepdims=True)
    mx = np.where(np.isfinite(mx), mx, 0)
    return (np.log(np.sum(np.exp(a - mx), axis=axis, keepdims=True)) + mx).squeeze(axis)


def sinkhorn(couplings: np.ndarray, log_mu: np.ndarray, log_nu: np.ndarray, iters: int):
    """superglue.py:143-150 for ONE pair: returns (Z + u + v, u, v)."""
    u = np.zeros_like(log_mu)
    v = np.zeros_like(log_nu)
    for _ in range(iters):
        u = log_mu - _lse(couplings + v[None, :], axis=1)
        v = log_nu - _lse(couplings + u[:, None], axis=0)
    return (couplings + u[:, None]) + v[None, :], u, v


def optimal_transport(scores: np.ndarray, alpha: float, iters: int, dtype=np.float32):
    """superglue.py:153-172 for ONE [m, n] score matrix: returns (Z [m+1, n+1], u [m+1], v [n+1])."""
    scores = np.asarray(scores, dtype=dtype)
    m, n = scores.shape
    a = dtype(alpha)
    aug = np.full((m + 1, n + 1), a, dtype=dtype)     # dust-bin row / column / corner all hold alpha
    aug[:m, :n] = scores
    norm = -np.log(dtype(m) + dtype(n))
    log_mu = np.full((m + 1,), norm, dtype=dtype)
    log_nu = np.full((n + 1,), norm, dtype=dtype)
    log_mu[m] = np.log(dtype(n)) + norm
    log_nu[n] = np.log(dtype(m)) + norm
    Z, u, v = sinkhorn(aug, log_mu, log_nu, iters)
    return Z - norm, u, v


def match_head(Z: np.ndarray, match_threshold: float):
    """superglue.py:268-285 on one transported matrix Z [m+1, n+1] → the forward()'s dictionary (NumPy)."""
    core = Z[:-1, :-1]
    m, n = core.shape
    i0 = np.argmax(core, axis=1)     # first maximum, like torch.max
    i1 = np.argmax(core, axis=0)
    mutual0 = i1[i0] == np.arange(m)
    mutual1 = i0[i1] == np.arange(n)
    ms0 = np.where(mutual0, np.exp(core[np.arange(m), i0]), 0).astype(Z.dtype)
    ms1 = np.where(mutual1, ms0[i1], 0).astype(Z.dtype)
    valid0 = mutual0 & (ms0 > Z.dtype.type(match_threshold))
    valid1 = mutual1 & valid0[i1]
    return {"matches0": np.where(valid0, i0, -1).astype(np.int64),
            "matches1": np.where(valid1, i1, -1).astype(np.int64),
            "matching_scores0": ms0, "matching_scores1": ms1}


def superglue_head(scores: np.ndarray, alpha: float = 1.0, iters: int = 20, match_threshold: float = 0.2,
                   dtype=np.float32):
    """optimal_transport + match_head; empty inputs follow superglue.py:235-242."""
    scores = np.asarray(scores)
    m, n = scores.shape
    if m == 0 or n == 0:
        return {"matches0": np.full((m,), -1, np.int64), "matches1": np.full((n,), -1, np.int64),
                "matching_scores0": np.zeros((m,), dtype), "matching_scores1": np.zeros((n,), dtype)}
    Z, u, v = optimal_transport(scores, alpha, iters, dtype)
    out = match_head(Z, match_threshold)
    out["Z"], out["u"], out["v"] = Z, u, v
    return out


def ambiguous_rows(Z64: np.ndarray, margin: float = 1e-4):
    """Rows / columns of the float64 transported matrix whose top two entries are closer than `margin`: a float32
    evaluation may legitimately pick either (used to mask index comparisons, like the float matcher's tie flags)."""
    core = Z64[:-1, :-1]
    m, n = core.shape
    def gap(a, axis):
        if a.shape[axis] < 2:
            return np.full(a.shape[1 - axis], np.inf)
        s = np.sort(a, axis=axis)
        return (s.take(-1, axis) - s.take(-2, axis))
    return gap(core, 1) < margin, gap(core, 0) < margin
