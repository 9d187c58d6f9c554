"""CPU: trajectory artefacts and the loop-closure pair enumeration (config 5).

tests/golden/tum_1018_13_*_head.txt are the first 24 lines of the reference's own shipped trajectories
(ros_ws/src/pose_estimation_results/1018_13/SuperPoint_SuperGlue_Stereo.txt and 1018_13_Ground_Truth.txt, copied with
`head -24`): they pin the on-disk format (`t x y z qx qy qz qw`, '%f')."""
import itertools
import os

import numpy as np
import torch

from forest_slam_b200 import trajectory_io as tio
from tests import util


def test_tum_roundtrip_on_reference_files(tmp_path):
    for name in ("tum_1018_13_stereo_head.txt", "tum_1018_13_gt_head.txt"):
        path = os.path.join(util.GOLDEN, name)
        stamps, T = tio.read_tum(path)
        assert T.shape == (24, 4, 4)
        out = tmp_path / name
        tio.write_tum(str(out), T, stamps)
        a, b = np.loadtxt(path), np.loadtxt(out)
        # a file only holds 6 decimals, so the matrix read back is not the matrix that was written: values agree to the
        # last printed digit; the SIGN is tf's (no normalisation): rows the reference wrote through tf (the estimate)
        # come back with the sign they had, a ground-truth row with qw < 0 (not produced by tf) comes back negated
        assert np.allclose(a[:, :4], b[:, :4], atol=1e-6)
        sgn = np.sign((a[:, 4:8] * b[:, 4:8]).sum(1, keepdims=True))
        assert np.allclose(a[:, 4:8], sgn * b[:, 4:8], atol=2e-6)
        if "stereo" in name:
            assert np.all(sgn > 0)
        # format: 8 columns of '%f'
        line = open(out).readline().split()
        assert len(line) == 8 and all(len(x.split(".")[1]) == 6 for x in line)


def test_quaternion_conventions():
    rng = np.random.default_rng(3)
    for _ in range(50):
        R = util.rodrigues(rng.normal(0, 1.5, 3))
        q = tio.quaternion_from_matrix(R)
        assert abs(np.linalg.norm(q) - 1) < 1e-12
        assert np.abs(tio.matrix_from_quaternion(q) - R).max() < 1e-12
    assert np.allclose(tio.quaternion_from_matrix(np.eye(3)), [0, 0, 0, 1])
    # tf's algorithm, literally (ROS tf.transformations): trace branch / largest-diagonal branch, no sign flip
    def tf_ref(R):
        import math
        M = np.eye(4); M[:3, :3] = R
        q = np.empty(4); t = np.trace(M)
        if t > M[3, 3]:
            q[3] = t; q[2] = M[1, 0] - M[0, 1]; q[1] = M[0, 2] - M[2, 0]; q[0] = M[2, 1] - M[1, 2]
        else:
            i, j, k = 0, 1, 2
            if M[1, 1] > M[0, 0]: i, j, k = 1, 2, 0
            if M[2, 2] > M[i, i]: i, j, k = 2, 0, 1
            t = M[i, i] - (M[j, j] + M[k, k]) + M[3, 3]
            q[i] = t; q[j] = M[i, j] + M[j, i]; q[k] = M[k, i] + M[i, k]; q[3] = M[k, j] - M[j, k]
        return q * (0.5 / math.sqrt(t * M[3, 3]))
    neg = 0
    for _ in range(200):
        R = util.rodrigues(rng.normal(0, 2.5, 3))
        q = tio.quaternion_from_matrix(R)
        assert np.array_equal(q, tf_ref(R))
        neg += q[3] < 0
    assert neg > 0   # the second branch does return qw < 0, and it is kept


def test_ape_rpe():
    rng = np.random.default_rng(5)
    T = np.tile(np.eye(4), (40, 1, 1))
    for i in range(1, 40):
        d = np.eye(4)
        d[:3, :3] = util.rodrigues(rng.normal(0, 0.02, 3))
        d[:3, 3] = [0.01, 0.0, 0.13]
        T[i] = T[i - 1] @ d
    G = np.eye(4)
    G[:3, :3] = util.rodrigues([0.3, -0.2, 0.9])
    G[:3, 3] = [4, -2, 7]
    moved = np.einsum("ij,njk->nik", G, T)
    assert tio.ape_translation(moved, T)["rmse"] < 1e-9           # a rigid motion is aligned away
    assert tio.rpe_translation(moved, T)["rmse"] < 1e-9           # and does not enter relative errors
    assert tio.ape_translation(moved, T, align=False)["rmse"] > 1.0
    noisy = T.copy()
    noisy[:, :3, 3] += rng.normal(0, 0.05, (40, 3))
    r = tio.ape_translation(noisy, T)
    assert 0.04 < r["rmse"] < 0.12 and r["max"] >= r["median"]
    # the shipped estimate against its ground truth: finite and in metres-scale (format sanity, stamps differ)
    _, Te = tio.read_tum(os.path.join(util.GOLDEN, "tum_1018_13_stereo_head.txt"))
    _, Tg = tio.read_tum(os.path.join(util.GOLDEN, "tum_1018_13_gt_head.txt"))
    assert np.isfinite(tio.ape_translation(Te, Tg)["rmse"])


def test_triangular_enumeration_and_sharding():
    from forest_slam_b200 import sequence as seq
    for n in (2, 3, 7, 50, 2000):
        total = n * (n - 1) // 2
        lin = torch.arange(total) if n <= 50 else torch.tensor([0, 1, 1998, 1999, 3997, total // 2, total - 2, total - 1])
        a, b = seq.triangular_index(n, lin)
        if n <= 50:
            ref = np.array(list(itertools.combinations(range(n), 2)))
            assert np.array_equal(a.numpy(), ref[:, 0]) and np.array_equal(b.numpy(), ref[:, 1])
        else:
            start = a * (2 * n - a - 1) // 2
            assert torch.all(a < b) and torch.all(b < n) and torch.equal(start + (b - a - 1), lin)
    for world in (1, 2, 4, 8):
        parts = [seq.shard_all_pairs(100, r, world, block=64) for r in range(world)]
        assert torch.equal(torch.cat(parts).sort().values, torch.arange(4950))
        # block-cyclic: every rank touches (almost) every keyframe
        for p in parts:
            a, b = seq.triangular_index(100, p)
            assert torch.unique(torch.cat([a, b])).numel() >= 90
