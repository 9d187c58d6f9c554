"""GPU parity: the SuperGlue score head (fs_sinkhorn_match through the C ABI) against the golden outputs of the
reference's own log_optimal_transport + match tail and against oracle/superglue_tail.py.

Bar: the head is float32 arithmetic in the reference; the dual potentials u, v and the matching scores must agree
within 2e-4 absolute / 1e-3 relative (CUDA's exp / log differ from torch's in the last ulps and the sums are
associated differently); match index sets must be identical wherever the float64 oracle does not flag the decision
as a numerical tie (top-2 gap < 1e-4) or the score as sitting on the threshold (|score - thr| < 1e-4)."""
import os

import numpy as np
import pytest

from tests import util
from oracle import superglue_tail as st

pytestmark = pytest.mark.gpu
SG = np.load(os.path.join(util.GOLDEN, "sinkhorn_golden.npz"))
U_TOL = 2e-4


@pytest.fixture(scope="module")
def fsb():
    import torch
    assert torch.cuda.is_available()
    from forest_slam_b200 import api
    return api


def _check_indices(got0, got1, ref0, ref1, r64, thr):
    amb_r, amb_c = st.ambiguous_rows(r64["Z"])
    edge0 = np.abs(r64["matching_scores0"] - thr) < 1e-4
    m, n = len(ref0), len(ref1)
    # a row's outcome depends on its own arg-max and on the arg-max of the column it points at
    for k in range(m):
        if got0[k] == ref0[k]:
            continue
        js = [j for j in (got0[k], ref0[k]) if j >= 0]
        assert amb_r[k] or edge0[k] or any(amb_c[j] for j in js), (k, got0[k], ref0[k])
    for j in range(n):
        if got1[j] == ref1[j]:
            continue
        ks = [k for k in (got1[j], ref1[j]) if k >= 0]
        assert amb_c[j] or any(amb_r[k] or edge0[k] for k in ks), (j, got1[j], ref1[j])


@pytest.mark.parametrize("i", range(int(SG["n_cases"])))
def test_sinkhorn_reference_golden(fsb, i):
    m, n, _, _, alpha, iters, thr = SG[f"c{i}_params"]
    sc = SG[f"c{i}_scores"]
    got = fsb.superglue_match_head(sc, bin_score=float(alpha), iters=int(iters), match_threshold=float(thr))
    r64 = st.superglue_head(sc, float(alpha), int(iters), float(thr), np.float64)
    _check_indices(got["matches0"], got["matches1"], SG[f"c{i}_matches0"], SG[f"c{i}_matches1"], r64, float(thr))
    same0 = got["matches0"] == SG[f"c{i}_matches0"]
    assert same0.mean() > 0.98
    mut = (got["matching_scores0"] > 0) & (SG[f"c{i}_mscores0"] > 0)
    assert np.allclose(got["matching_scores0"][mut], SG[f"c{i}_mscores0"][mut], rtol=1e-3, atol=1e-6)
    mut1 = (got["matching_scores1"] > 0) & (SG[f"c{i}_mscores1"] > 0)
    assert np.allclose(got["matching_scores1"][mut1], SG[f"c{i}_mscores1"][mut1], rtol=1e-3, atol=1e-6)
    if i == 3:  # exact duplicate rows / columns tie exactly: lowest index wins, like torch.max
        assert np.array_equal(got["matches0"], SG["c3_matches0"]) and np.array_equal(got["matches1"], SG["c3_matches1"])


@pytest.mark.parametrize("m_max,n_max,P,seed", [(256, 256, 5, 1), (1024, 1024, 3, 2), (300, 2048, 2, 3),
                                                (1500, 132, 2, 4), (64, 4096, 1, 5)])
def test_sinkhorn_batch_vs_oracle(fsb, m_max, n_max, P, seed):
    """Ragged batch (per-pair m, n below the padded sizes), potentials and matches against the float32 / float64 oracle."""
    import torch
    rng = np.random.default_rng(seed)
    ms = [m_max] + [int(rng.integers(1, m_max + 1)) for _ in range(P - 1)]
    ns = [n_max] + [int(rng.integers(1, n_max + 1)) for _ in range(P - 1)]
    buf = np.full((P, m_max, n_max), np.nan, np.float32)   # whatever lies outside a pair's block must never be read
    for p in range(P):
        buf[p, :ms[p], :ns[p]] = util.sinkhorn_scores(ms[p], ns[p], 1000 * seed + p)
    dev = torch.device("cuda:0")
    r = fsb.sinkhorn_match_batch(torch.from_numpy(buf).to(dev), torch.tensor(ms, dtype=torch.int32),
                                 torch.tensor(ns, dtype=torch.int32), bin_score=1.0, iters=20, match_threshold=0.2,
                                 want_potentials=True)
    torch.cuda.synchronize()
    for p in range(P):
        m, n = ms[p], ns[p]
        sc = buf[p, :m, :n]
        r64 = st.superglue_head(sc, 1.0, 20, 0.2, np.float64)
        r32 = st.superglue_head(sc, 1.0, 20, 0.2, np.float32)
        u = r["u"][p, :m].cpu().numpy()
        v = r["v"][p, :n].cpu().numpy()
        assert np.abs(u - r64["u"][:m]).max() < U_TOL, np.abs(u - r64["u"][:m]).max()
        assert np.abs(v - r64["v"][:n]).max() < U_TOL
        g0 = r["matches0"][p].cpu().numpy()
        g1 = r["matches1"][p].cpu().numpy()
        assert (g0[m:] == -1).all() and (g1[n:] == -1).all()
        _check_indices(g0[:m], g1[:n], r32["matches0"], r32["matches1"], r64, 0.2)
        s0 = r["matching_scores0"][p, :m].cpu().numpy()
        mut = (s0 > 0) & (r64["matching_scores0"] > 0)
        assert mut.sum() > 0 or m < 4
        assert np.allclose(s0[mut], r64["matching_scores0"][mut], rtol=1e-3, atol=1e-6)


def test_sinkhorn_degenerate(fsb):
    import torch
    out = fsb.superglue_match_head(np.zeros((0, 7), np.float32))
    assert out["matches0"].shape == (0,) and (out["matches1"] == -1).all() and out["matches1"].shape == (7,)
    # one keypoint on each side: Z = log-OT of a 2x2 matrix, mutual by construction
    sc = np.array([[3.0]], np.float32)
    got = fsb.superglue_match_head(sc)
    ref = st.superglue_head(sc, 1.0, 20, 0.2, np.float32)
    assert np.array_equal(got["matches0"], ref["matches0"]) and np.array_equal(got["matches1"], ref["matches1"])
    assert np.allclose(got["matching_scores0"], ref["matching_scores0"], rtol=1e-3)
    # zero iterations = plain mutual arg-max of the raw scores (u = v = 0)
    sc = util.sinkhorn_scores(50, 60, 77)
    got = fsb.superglue_match_head(sc, iters=0, match_threshold=0.0)
    ref = st.superglue_head(sc, 1.0, 0, 0.0, np.float32)
    assert np.array_equal(got["matches0"], ref["matches0"]) and np.array_equal(got["matches1"], ref["matches1"])
    # a pair with m = 0 inside a batch leaves its outputs at -1 / 0 and does not disturb its neighbour
    buf = torch.zeros((2, 40, 44), dtype=torch.float32, device="cuda:0")
    buf[1] = torch.from_numpy(util.sinkhorn_scores(40, 44, 78)).cuda()
    r = fsb.sinkhorn_match_batch(buf, torch.tensor([0, 40], dtype=torch.int32), torch.tensor([44, 44], dtype=torch.int32))
    assert (r["matches0"][0] == -1).all() and (r["matches1"][0] == -1).all()
    ref = st.superglue_head(buf[1].cpu().numpy(), 1.0, 20, 0.2, np.float32)
    assert np.array_equal(r["matches0"][1].cpu().numpy(), ref["matches0"])


def test_sinkhorn_group_size_invariance(fsb):
    """pairs_per_group only changes launch grouping (L2 residency), never a bit of the result."""
    import torch
    P, m, n = 6, 200, 256
    buf = torch.from_numpy(np.stack([util.sinkhorn_scores(m, n, 900 + p) for p in range(P)])).cuda()
    cm = torch.full((P,), m, dtype=torch.int32)
    cn = torch.full((P,), n, dtype=torch.int32)
    a = fsb.sinkhorn_match_batch(buf, cm, cn, want_potentials=True, pairs_per_group=0)
    b = fsb.sinkhorn_match_batch(buf, cm, cn, want_potentials=True, pairs_per_group=1)
    for k in ("matches0", "matches1", "matching_scores0", "u", "v"):
        assert torch.equal(a[k], b[k]), k


def test_sinkhorn_streaming_variant_agrees(fsb):
    """The per-iteration (L2-streaming) fallback and the shared-memory-resident kernel compute the same thing."""
    import torch
    P, m, n = 3, 333, 512
    buf = torch.from_numpy(np.stack([util.sinkhorn_scores(m, n, 950 + p) for p in range(P)])).cuda()
    cm = torch.tensor([m, 17, 200], dtype=torch.int32)
    cn = torch.tensor([n, 512, 3], dtype=torch.int32)
    a = fsb.sinkhorn_match_batch(buf, cm, cn, want_potentials=True)
    b = fsb.sinkhorn_match_batch(buf, cm, cn, want_potentials=True, pairs_per_group=-1)
    torch.cuda.synchronize()
    assert (a["u"] - b["u"]).abs().max().item() < 1e-4 and (a["v"] - b["v"]).abs().max().item() < 1e-4
    assert (a["matches0"] != b["matches0"]).float().mean().item() < 0.01
    assert torch.allclose(a["matching_scores0"], b["matching_scores0"], rtol=1e-3, atol=1e-6) or \
        ((a["matching_scores0"] > 0) != (b["matching_scores0"] > 0)).float().mean().item() < 0.01


def test_superglue_head_from_descriptors(fsb):
    """Contraction on the tensor cores (fs_float_scores) + Sinkhorn head against the oracle fed with the float64
    score matrix: scores to 2e-6 absolute (fp32-grade), then the usual head bars."""
    import torch
    n_max, F = 512, 4
    counts = [512, 400, 130, 0]
    bank = np.zeros((F, 256, n_max), np.float32)
    for f in range(F - 1):
        d0, _ = util.float_descriptors(counts[f], 8, 40 + f)
        bank[f, :, :counts[f]] = d0
    # frames 0 and 1 observe shared landmarks
    d0, d1 = util.float_descriptors(counts[0], counts[1], 77)
    bank[0, :, :counts[0]], bank[1, :, :counts[1]] = d0, d1
    dev = torch.device("cuda:0")
    gain = 20.0  # scores of matched pairs near 20 * cos, like tests.util.sinkhorn_scores
    r = fsb.superglue_head_batch(torch.from_numpy(bank).to(dev), torch.tensor(counts, dtype=torch.int32),
                                 [0, 1, 2, 0], [1, 2, 0, 3], scale=gain)
    torch.cuda.synchronize()
    for p, (a, b) in enumerate([(0, 1), (1, 2), (2, 0), (0, 3)]):
        m, n = counts[a], counts[b]
        g0 = r["matches0"][p].cpu().numpy()
        g1 = r["matches1"][p].cpu().numpy()
        if m == 0 or n == 0:
            assert (g0 == -1).all() and (g1 == -1).all()
            continue
        ref_sc = (gain * (bank[a, :, :m].T.astype(np.float64) @ bank[b, :, :n].astype(np.float64)))
        got_sc = r["scores"][p, :m, :n].cpu().numpy()
        assert np.abs(got_sc - ref_sc).max() < 2e-6 * gain, np.abs(got_sc - ref_sc).max()
        r64 = st.superglue_head(ref_sc, 1.0, 20, 0.2, np.float64)
        r32 = st.superglue_head(ref_sc.astype(np.float32), 1.0, 20, 0.2, np.float32)
        _check_indices(g0[:m], g1[:n], r32["matches0"], r32["matches1"], r64, 0.2)
        assert (g0[m:] == -1).all() and (g1[n:] == -1).all()
        if (a, b) == (0, 1):
            assert (g0[:m] >= 0).sum() > 100
