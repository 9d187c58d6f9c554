"""Host-side mirror of the call expressions the reference's frame loop makes.

Two layers:

* batched operators (``*_batch``) taking CUDA ``torch.Tensor`` frame banks / correspondence batches and
  returning CUDA tensors — the form the sequence and loop-closure drivers use;
* cv2-shaped single-pair functions (``BFMatcher``, ``findEssentialMat``, ``recoverPose``,
  ``solvePnPRansac``, ``Rodrigues`` …) taking and returning NumPy arrays exactly like the calls in
  ``ros_ws/src/mono_slam.py:106-112`` and ``ros_ws/src/stereo_slam.py:85,234,264-298``, so that the
  reference scripts can swap ``cv2.`` for ``fsb.`` on those lines.

Everything runs in the CUDA extension through the C ABI (include/forestslam_b200.h); PyTorch only owns
device memory and streams.  There is no CPU path in this module.
"""
from __future__ import annotations

import collections
import ctypes as C
import dataclasses
import math
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import FSCamera

NORM_HAMMING = 6  # cv2.NORM_HAMMING
NORM_L2 = 4       # cv2.NORM_L2
RANSAC = 8        # cv2.RANSAC
SOLVEPNP_ITERATIVE = 0

# the tcgen05 matchers are the default: csrc/float_top3.cu (certified single fp16 pass per frame pair; also the Hamming
# path on +-1 operands, csrc/hamming_tc.cu); csrc/float_tc.cu (three-MMA split-fp16) on request or above 4096 columns;
# the exact-fp32 CUDA-core kernel serves descriptor counts that are not a multiple of 128 and `exact_fp32=True`
TENSOR_CORE_MATCHER = True

DMatch = collections.namedtuple("DMatch", ["queryIdx", "trainIdx", "distance"])


# ------------------------------------------------------------------------------------------- constants
@dataclasses.dataclass(frozen=True)
class ForestCamera:
    """Intrinsics / distortion / baseline of the BotanicGarden rig (stereo_slam.py:45-64, mono_slam.py:40-50)."""
    fx: float = 642.9165664800531
    fy: float = 641.9171825800378
    cx: float = 460.1840658156501
    cy: float = 308.5846449100310
    dist: Tuple[float, ...] = (-0.060164620903866, 0.094005180631043, 0.0, 0.0, 0.0)
    baseline: float = 0.253736175410149  # np.linalg.norm(T_rgb0_rgb1[:3, 3]) of the 1x16 array, stereo_slam.py:270
    width: int = 960
    height: int = 600

    @property
    def K(self) -> np.ndarray:
        return np.array([[self.fx, 0.0, self.cx], [0.0, self.fy, self.cy], [0.0, 0.0, 1.0]])

    def as_struct(self) -> FSCamera:
        s = FSCamera()
        s.fx, s.fy, s.cx, s.cy = self.fx, self.fy, self.cx, self.cy
        d = list(self.dist) + [0.0] * (5 - len(self.dist))
        for i in range(5):
            s.dist[i] = float(d[i])
        return s


def camera_from_K(K, dist=None) -> ForestCamera:
    K = np.asarray(K, dtype=np.float64)
    d = np.zeros(5) if dist is None else np.asarray(dist, dtype=np.float64).reshape(-1)
    d = np.concatenate([d, np.zeros(max(0, 5 - d.size))])[:5]
    return ForestCamera(fx=float(K[0, 0]), fy=float(K[1, 1]), cx=float(K[0, 2]), cy=float(K[1, 2]),
                        dist=tuple(float(x) for x in d))


# ------------------------------------------------------------------------------------------- plumbing
def _dev(device=None) -> torch.device:
    if not torch.cuda.is_available():
        raise _lib.ForestSlamError("a CUDA device is required: forest_slam_b200 has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


def _p(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


_workspaces = {}
_ws_scope = ["global"]


class workspace_scope:
    """Private workspace set for everything enqueued inside the `with` block.  Workspaces are cached per
    (device, operator); two launch sequences that may run CONCURRENTLY (two captured pipelines replayed on two
    streams) must not share them, so each captures under its own scope."""

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        _ws_scope.append(self.name)
        return self

    def __exit__(self, *exc):
        _ws_scope.pop()
        return False


def _workspace(nbytes: int, device: torch.device, slot: str = "default") -> torch.Tensor:
    key = (device.index, slot, _ws_scope[-1])
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def _chk(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError(f"{name} must be a CUDA torch.Tensor")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.contiguous()


def _to_cuda(a, dtype, device) -> torch.Tensor:
    if isinstance(a, torch.Tensor):
        return a.to(device=device, dtype=dtype, non_blocking=True).contiguous()
    return torch.from_numpy(np.ascontiguousarray(a)).to(device=device, dtype=dtype, non_blocking=True)


def _pairs(pair_a, pair_b, device):
    a = _to_cuda(pair_a, torch.int32, device)
    b = _to_cuda(pair_b, torch.int32, device)
    if a.shape != b.shape or a.dim() != 1:
        raise ValueError("pair_a / pair_b must be 1-D and of equal length")
    return a, b


# ------------------------------------------------------------------------------------------- A2
HAMMING_TC_MAX_COLS = 4096  # fs_hamming_match_tc packs the column index into 12 key bits


def hamming_prepare(desc: torch.Tensor, counts: torch.Tensor) -> torch.Tensor:
    """[F, n_max, 32] uint8 ORB rows → [F, n_max + 1, 512] fp16 operand bank of the tensor-core Hamming matcher
    (every bit a +-1 operand: dot = 256 - 2 * hamming, exact).  Once per FRAME; any number of pairs reuse it."""
    lib = _lib.load()
    desc = _chk(desc, torch.uint8, "desc")
    if desc.dim() != 3 or desc.shape[2] != 32:
        raise ValueError("desc must be [n_frames, n_max, 32] uint8")
    dev = desc.device
    F, n_max = int(desc.shape[0]), int(desc.shape[1])
    counts = _chk(counts.to(dev), torch.int32, "counts")
    ops = torch.empty((F, n_max + 1, 512), dtype=torch.float16, device=dev)
    assert ops.numel() * 2 == lib.fs_hamming_operand_bytes(F, n_max) or F == 0 or n_max == 0
    _lib.check(lib.fs_hamming_prepare(_p(desc), F, n_max, _p(counts), _p(ops), _stream()), "fs_hamming_prepare")
    return ops


def hamming_match_batch(desc: Optional[torch.Tensor], counts: torch.Tensor, pair_a, pair_b, *,
                        cross_check: bool = True, ratio: Optional[float] = None, want_matches1: bool = True,
                        operands: Optional[torch.Tensor] = None, popc: bool = False):
    """Batched cv2.BFMatcher(NORM_HAMMING, crossCheck).match over a frame bank (stereo_slam.py:85,234,242).

    desc [F, n_max, 32] uint8, counts [F] int32 → dict(matches0, dist0, second0, matches1, n_matches).
    Default: the tensor-core kernel (hamming_tc.cu) on ``operands`` from hamming_prepare (prepared on the fly when
    absent); ``popc=True`` or n_max > 4096 selects the POPC-pipe kernel (hamming.cu).  Both give the same bits."""
    lib = _lib.load()
    if desc is not None:
        desc = _chk(desc, torch.uint8, "desc")
        if desc.dim() != 3 or desc.shape[2] != 32:
            raise ValueError("desc must be [n_frames, n_max, 32] uint8")
        dev = desc.device
        F, n_max = int(desc.shape[0]), int(desc.shape[1])
    else:
        if operands is None:
            raise ValueError("either desc or operands is required")
        dev = operands.device
        F, n_max = int(operands.shape[0]), int(operands.shape[1]) - 1
    counts = _chk(counts.to(dev), torch.int32, "counts")
    a, b = _pairs(pair_a, pair_b, dev)
    P = int(a.numel())
    m0 = torch.empty((P, n_max), dtype=torch.int32, device=dev)
    d0 = torch.empty_like(m0)
    s0 = torch.empty_like(m0)
    m1 = torch.empty_like(m0) if want_matches1 else None
    nm = torch.empty((P,), dtype=torch.int32, device=dev)
    out = {"matches0": m0, "dist0": d0, "second0": s0, "matches1": m1, "n_matches": nm}
    use_tc = TENSOR_CORE_MATCHER and not popc and n_max <= HAMMING_TC_MAX_COLS
    if use_tc:
        if operands is None:
            operands = hamming_prepare(desc, counts)
        operands = _chk(operands, torch.float16, "operands")
        if tuple(operands.shape) != (F, n_max + 1, 512):
            raise ValueError("operands must be the [n_frames, n_max + 1, 512] fp16 bank hamming_prepare returns")
        ws = _workspace(lib.fs_hamming_match_tc_workspace_bytes(P, n_max), dev)
        _lib.check(lib.fs_hamming_match_tc(_p(operands), F, n_max, _p(counts), _p(a), _p(b), P,
                                           int(bool(cross_check)), float(ratio) if ratio else 0.0, _p(m0), _p(d0),
                                           _p(s0), _p(m1), _p(nm), _p(ws), ws.numel(), _stream()),
                   "fs_hamming_match_tc")
        return out
    if desc is None:
        raise ValueError("the POPC kernel needs desc")
    nbytes = lib.fs_hamming_match_workspace_bytes(P, n_max)
    ws = _workspace(nbytes, dev)
    _lib.check(lib.fs_hamming_match(_p(desc), F, n_max, _p(counts), _p(a), _p(b), P, int(bool(cross_check)),
                                    float(ratio) if ratio else 0.0, _p(m0), _p(d0), _p(s0), _p(m1), _p(nm),
                                    _p(ws), ws.numel(), _stream()), "fs_hamming_match")
    return out


# ------------------------------------------------------------------------------------------- A1
def float_prepare(desc: torch.Tensor, counts: torch.Tensor) -> torch.Tensor:
    """[F, 256, n_max] fp32 channel-major (SuperPoint layout) → [F, n_max + 1, 512] fp16 operand bank: rows < n_max
    are the (hi | lo) split of the descriptors, row n_max holds the frame's norm statistics (float_ops.cuh)."""
    lib = _lib.load()
    desc = _chk(desc, torch.float32, "desc")
    if desc.dim() != 3 or desc.shape[1] != 256:
        raise ValueError("desc must be [n_frames, 256, n_max] float32")
    dev = desc.device
    F, n_max = int(desc.shape[0]), int(desc.shape[2])
    counts = _chk(counts.to(dev), torch.int32, "counts")
    ops = torch.empty((F, n_max + 1, 512), dtype=torch.float16, device=dev)
    assert ops.numel() * 2 == lib.fs_float_operand_bytes(F, n_max) or F == 0 or n_max == 0
    _lib.check(lib.fs_float_prepare(_p(desc), F, n_max, _p(counts), _p(ops), _stream()), "fs_float_prepare")
    return ops


FLOAT_TOP3_MAX_COLS = 4096  # fs_float_match packs the column index into 12 key bits


def float_match_batch(desc: torch.Tensor, counts: torch.Tensor, pair_a, pair_b, *, scale: float = 1.0,
                      threshold: Optional[float] = None, ratio: Optional[float] = None,
                      operands: Optional[torch.Tensor] = None, exact_fp32: bool = False, split3: bool = False,
                      want_diag: bool = False, cross_check: bool = True):
    """Batched float-descriptor mutual-NN (+threshold, +ratio): superglue.py:259-260, :268-278.

    desc [F, 256, n_max] fp32.  Default: the certified single-pass tcgen05 kernel on the fp16 operand bank
    (``operands`` from float_prepare, or prepared on the fly); ``split3`` selects the three-MMA split-fp16 tensor-core
    kernel, ``exact_fp32`` the CUDA-core fp32 kernel on ``desc``.  All three give the index sets of the exact score."""
    lib = _lib.load()
    if desc is not None:
        desc = _chk(desc, torch.float32, "desc")
        dev = desc.device
        F, n_max = int(desc.shape[0]), int(desc.shape[2])
    else:
        if operands is None:
            raise ValueError("either desc or operands is required")
        dev = operands.device
        F, n_max = int(operands.shape[0]), int(operands.shape[1]) - 1
    counts = _chk(counts.to(dev), torch.int32, "counts")
    a, b = _pairs(pair_a, pair_b, dev)
    P = int(a.numel())
    m0 = torch.empty((P, n_max), dtype=torch.int32, device=dev)
    m1 = torch.empty_like(m0)
    sc = torch.empty((P, n_max), dtype=torch.float32, device=dev)
    se = torch.empty_like(sc)
    nm = torch.empty((P,), dtype=torch.int32, device=dev)
    thr = -math.inf if threshold is None else float(threshold)
    rat = float(ratio) if ratio else 0.0
    use_tc = TENSOR_CORE_MATCHER and (not exact_fp32) and n_max % 128 == 0
    if use_tc and n_max > FLOAT_TOP3_MAX_COLS:
        split3 = True
    if not cross_check and (not use_tc or split3):
        raise ValueError("cross_check=False runs on the certified tensor-core kernel: n_max % 128 == 0, n_max <= 4096")
    out = {"matches0": m0, "matches1": m1, "score0": sc, "second0": se, "n_matches": nm}
    if use_tc:
        if operands is None:
            operands = float_prepare(desc, counts)
        operands = _chk(operands, torch.float16, "operands")
        if tuple(operands.shape) != (F, n_max + 1, 512):
            raise ValueError("operands must be the [n_frames, n_max + 1, 512] fp16 bank float_prepare returns")
        if split3:
            ws = _workspace(lib.fs_float_match_split3_workspace_bytes(P, n_max), dev)
            _lib.check(lib.fs_float_match_split3(_p(operands), F, n_max, _p(counts), _p(a), _p(b), P, float(scale), thr,
                                                 rat, _p(m0), _p(sc), _p(se), _p(m1), _p(nm), _p(ws), ws.numel(),
                                                 _stream()), "fs_float_match_split3")
        else:
            ws = _workspace(lib.fs_float_match_tc_workspace_bytes(P, n_max), dev)
            diag = torch.empty((4,), dtype=torch.int32, device=dev) if want_diag else None
            if not cross_check:
                _lib.check(lib.fs_float_match_knn(_p(operands), F, n_max, _p(counts), _p(a), _p(b), P, float(scale), thr,
                                                  rat, _p(m0), _p(sc), _p(se), _p(nm), _p(diag), _p(ws), ws.numel(),
                                                  _stream()), "fs_float_match_knn")
                out["matches1"] = None
                if want_diag:
                    out["diag"] = diag
                return out
            _lib.check(lib.fs_float_match(_p(operands), F, n_max, _p(counts), _p(a), _p(b), P, float(scale), thr,
                                          rat, _p(m0), _p(sc), _p(se), _p(m1), _p(nm), _p(diag), _p(ws), ws.numel(),
                                          _stream()), "fs_float_match")
            if want_diag:
                out["diag"] = diag
    else:
        if desc is None:
            raise ValueError("the fp32 kernel needs desc")
        ws = _workspace(lib.fs_float_match_workspace_bytes(P, n_max), dev)
        _lib.check(lib.fs_float_match_fp32(_p(desc), F, n_max, _p(counts), _p(a), _p(b), P, float(scale), thr,
                                           rat, _p(m0), _p(sc), _p(se), _p(m1), _p(nm), _p(ws), ws.numel(),
                                           _stream()), "fs_float_match_fp32")
    return out


# ------------------------------------------------------------------------------------------- A1' (Sinkhorn head)
def sinkhorn_match_batch(scores: torch.Tensor, m_counts: torch.Tensor, n_counts: torch.Tensor, *,
                         bin_score: float = 1.0, iters: int = 20, match_threshold: float = 0.2,
                         pairs_per_group: int = 0, want_potentials: bool = False):
    """Batched SuperGlue score head after the contraction: log_optimal_transport + mutual arg-max + exp +
    match_threshold (superglue.py:143-172, :263-285; hyper-parameters of mono_slam.py:59-63).

    scores [P, m_max, n_max] f32 on the device (n_max a multiple of 4; pair p uses its m_counts[p] x n_counts[p] block)
    → dict(matches0, matching_scores0 [P, m_max]; matches1, matching_scores1 [P, n_max]) (+ u, v if asked)."""
    lib = _lib.load()
    scores = _chk(scores, torch.float32, "scores")
    if scores.dim() != 3:
        raise ValueError("scores must be [n_pairs, m_max, n_max]")
    dev = scores.device
    P, m_max, n_max = (int(x) for x in scores.shape)
    m_counts = _chk(m_counts.to(dev), torch.int32, "m_counts")
    n_counts = _chk(n_counts.to(dev), torch.int32, "n_counts")
    if m_counts.numel() != P or n_counts.numel() != P:
        raise ValueError("one (m, n) count per pair is required")
    m0 = torch.empty((P, m_max), dtype=torch.int32, device=dev)
    m1 = torch.empty((P, n_max), dtype=torch.int32, device=dev)
    s0 = torch.empty((P, m_max), dtype=torch.float32, device=dev)
    s1 = torch.empty((P, n_max), dtype=torch.float32, device=dev)
    u = torch.zeros((P, m_max), dtype=torch.float32, device=dev) if want_potentials else None
    v = torch.zeros((P, n_max), dtype=torch.float32, device=dev) if want_potentials else None
    ws = _workspace(lib.fs_sinkhorn_workspace_bytes(P, m_max, n_max), dev, "sinkhorn")
    _lib.check(lib.fs_sinkhorn_match(_p(scores), P, m_max, n_max, _p(m_counts), _p(n_counts), float(bin_score),
                                     int(iters), float(match_threshold), int(pairs_per_group), _p(m0), _p(m1),
                                     _p(s0), _p(s1), _p(u), _p(v), _p(ws), ws.numel(), _stream()),
               "fs_sinkhorn_match")
    out = {"matches0": m0, "matches1": m1, "matching_scores0": s0, "matching_scores1": s1}
    if want_potentials:
        out["u"], out["v"] = u, v
    return out


def float_scores_batch(counts: torch.Tensor, pair_a, pair_b, *, operands: torch.Tensor, scale: float = 0.0625):
    """scores[p] = scale * desc_a^T desc_b (superglue.py:259-260; scale = 1/sqrt(256)) on the tensor cores →
    [P, n_max, n_max] f32 on the device (rows / columns beyond the frames' counts carry no information)."""
    lib = _lib.load()
    operands = _chk(operands, torch.float16, "operands")
    dev = operands.device
    F, n_max = int(operands.shape[0]), int(operands.shape[1]) - 1
    counts = _chk(counts.to(dev), torch.int32, "counts")
    a, b = _pairs(pair_a, pair_b, dev)
    P = int(a.numel())
    out = torch.empty((P, n_max, n_max), dtype=torch.float32, device=dev)
    _lib.check(lib.fs_float_scores(_p(operands), F, n_max, _p(counts), _p(a), _p(b), P, float(scale), _p(out),
                                   _stream()), "fs_float_scores")
    return out


def superglue_head_batch(desc: Optional[torch.Tensor], counts: torch.Tensor, pair_a, pair_b, *,
                         operands: Optional[torch.Tensor] = None, scale: float = 0.0625, bin_score: float = 1.0,
                         iters: int = 20, match_threshold: float = 0.2):
    """The whole tail of SuperGlue.forward after final_proj (superglue.py:259-285) for a batch of frame pairs:
    contraction / sqrt(d) on the tensor cores → log-Sinkhorn with dust-bins → mutual arg-max, exp, threshold.
    desc [F, 256, n_max] f32 (n_max a multiple of 128) or the prepared operand bank."""
    if operands is None:
        if desc is None:
            raise ValueError("either desc or operands is required")
        operands = float_prepare(_chk(desc, torch.float32, "desc"), counts)
    dev = operands.device
    counts = _chk(counts.to(dev), torch.int32, "counts")
    a, b = _pairs(pair_a, pair_b, dev)
    scores = float_scores_batch(counts, a, b, operands=operands, scale=scale)
    r = sinkhorn_match_batch(scores, counts[a.long()].contiguous(), counts[b.long()].contiguous(),
                             bin_score=bin_score, iters=iters, match_threshold=match_threshold)
    r["scores"] = scores
    return r


def superglue_match_head(scores, *, bin_score: float = 1.0, iters: int = 20, match_threshold: float = 0.2):
    """One pair, host arrays in / host arrays out: the dictionary SuperGlue.forward returns (superglue.py:280-285)
    for a [m, n] score matrix (already scaled by 1/sqrt(descriptor_dim), superglue.py:260)."""
    dev = _dev()
    sc = np.asarray(scores, dtype=np.float32)
    m, n = sc.shape
    if m == 0 or n == 0:  # superglue.py:235-242
        return {"matches0": np.full((m,), -1, np.int64), "matches1": np.full((n,), -1, np.int64),
                "matching_scores0": np.zeros((m,), np.float32), "matching_scores1": np.zeros((n,), np.float32)}
    n_pad = (n + 3) // 4 * 4
    buf = torch.zeros((1, m, n_pad), dtype=torch.float32, device=dev)
    buf[0, :, :n] = _to_cuda(sc, torch.float32, dev)
    cnt_m = torch.tensor([m], dtype=torch.int32, device=dev)
    cnt_n = torch.tensor([n], dtype=torch.int32, device=dev)
    r = sinkhorn_match_batch(buf, cnt_m, cnt_n, bin_score=bin_score, iters=iters, match_threshold=match_threshold)
    return {"matches0": r["matches0"][0, :m].cpu().numpy().astype(np.int64),
            "matches1": r["matches1"][0, :n].cpu().numpy().astype(np.int64),
            "matching_scores0": r["matching_scores0"][0, :m].cpu().numpy(),
            "matching_scores1": r["matching_scores1"][0, :n].cpu().numpy()}


# ------------------------------------------------------------------------------------------- A3
def gather_matches_batch(kpts: torch.Tensor, counts: torch.Tensor, pair_a, pair_b, matches0: torch.Tensor):
    """valid = matches > -1; mkpts0 = kpts0[valid]; mkpts1 = kpts1[matches[valid]] (mono_slam.py:106-108)."""
    lib = _lib.load()
    kpts = _chk(kpts, torch.float32, "kpts")
    dev = kpts.device
    F, n_max = int(kpts.shape[0]), int(kpts.shape[1])
    counts = _chk(counts.to(dev), torch.int32, "counts")
    a, b = _pairs(pair_a, pair_b, dev)
    P = int(a.numel())
    matches0 = _chk(matches0, torch.int32, "matches0")
    mk0 = torch.empty((P, n_max, 2), dtype=torch.float32, device=dev)  # rows >= n_corr[p] are never read
    mk1 = torch.empty_like(mk0)
    nc = torch.empty((P,), dtype=torch.int32, device=dev)
    src = torch.empty((P, n_max), dtype=torch.int32, device=dev)
    _lib.check(lib.fs_gather_matches(_p(kpts), F, n_max, _p(counts), _p(a), _p(b), P, _p(matches0), _p(mk0),
                                     _p(mk1), _p(nc), _p(src), _stream()), "fs_gather_matches")
    return {"mkpts0": mk0, "mkpts1": mk1, "n_corr": nc, "src_idx": src}


# ------------------------------------------------------------------------------------------- A6
def backproject_stereo(pts, disparity, *, camera: ForestCamera = ForestCamera(), baseline: Optional[float] = None,
                       zmin: float = 0.1, zmax: float = 1000.0, map_index=None):
    """depth = fx*B/disp; Z = depth[int(y), int(x)]; X, Y pin-hole; gate (stereo_slam.py:264-289).

    pts [n, 2] f32; disparity [n] (per point) or [H, W] / [M, H, W] dense maps.  → (points3D [n,3], valid [n])."""
    lib = _lib.load()
    dev = pts.device if isinstance(pts, torch.Tensor) and pts.is_cuda else _dev()
    pts_t = _to_cuda(pts, torch.float32, dev).reshape(-1, 2)
    disp = _to_cuda(disparity, torch.float32, dev)
    n = int(pts_t.shape[0])
    if disp.dim() == 1:
        if disp.numel() != n:
            raise ValueError("per-point disparity must have one value per point")
        mh = mw = 0
    else:
        if disp.dim() == 2:
            disp = disp[None]
        mh, mw = int(disp.shape[1]), int(disp.shape[2])
    mi = None if map_index is None else _to_cuda(map_index, torch.int32, dev)
    out = torch.empty((n, 3), dtype=torch.float32, device=dev)
    valid = torch.empty((n,), dtype=torch.uint8, device=dev)
    cam = camera.as_struct()
    B = camera.baseline if baseline is None else float(baseline)
    _lib.check(lib.fs_backproject_stereo(_p(pts_t), _p(disp), _p(mi), n, mh, mw, C.byref(cam), B, zmin, zmax,
                                         _p(out), _p(valid), _stream()), "fs_backproject_stereo")
    return out, valid


def backproject_filter_batch(mkpts0, mkpts1, disparity, n_corr, *, camera: ForestCamera = ForestCamera(),
                             baseline: Optional[float] = None, zmin: float = 0.1, zmax: float = 1000.0):
    """Batched A6 + the joint filter of stereo_slam.py:288-289 → dict(points3d, pts2d, n_out)."""
    lib = _lib.load()
    mkpts0 = _chk(mkpts0, torch.float32, "mkpts0")
    mkpts1 = _chk(mkpts1, torch.float32, "mkpts1")
    disparity = _chk(disparity, torch.float32, "disparity")
    n_corr = _chk(n_corr, torch.int32, "n_corr")
    dev = mkpts0.device
    P, n_max = int(mkpts0.shape[0]), int(mkpts0.shape[1])
    p3 = torch.empty((P, n_max, 3), dtype=torch.float32, device=dev)  # rows >= n_out[p] are never read
    p2 = torch.empty((P, n_max, 2), dtype=torch.float32, device=dev)
    no = torch.empty((P,), dtype=torch.int32, device=dev)
    cam = camera.as_struct()
    B = camera.baseline if baseline is None else float(baseline)
    _lib.check(lib.fs_backproject_filter(_p(mkpts0), _p(mkpts1), _p(disparity), _p(n_corr), P, n_max,
                                         C.byref(cam), B, zmin, zmax, _p(p3), _p(p2), _p(no), _stream()),
               "fs_backproject_filter")
    return {"points3d": p3, "pts2d": p2, "n_out": no}


def backproject_filter_bank(mkpts0, mkpts1, n_corr, disparity, pair_a, src_idx=None, *,
                            camera: ForestCamera = ForestCamera(), baseline: Optional[float] = None,
                            zmin: float = 0.1, zmax: float = 1000.0):
    """A6 over a frame bank with the reference's lookup (stereo_slam.py:274-289): ``disparity`` is the per-keypoint
    bank [F, n_max] (needs ``src_idx`` from gather_matches_batch) or the frames' dense maps [F, H, W] (looked up at
    (int(y), int(x)) of the previous-left keypoint) → dict(points3d, pts2d, n_out)."""
    lib = _lib.load()
    mkpts0 = _chk(mkpts0, torch.float32, "mkpts0")
    mkpts1 = _chk(mkpts1, torch.float32, "mkpts1")
    disparity = _chk(disparity, torch.float32, "disparity")
    n_corr = _chk(n_corr, torch.int32, "n_corr")
    dev = mkpts0.device
    P, n_max = int(mkpts0.shape[0]), int(mkpts0.shape[1])
    pa = _chk(_to_cuda(pair_a, torch.int32, dev), torch.int32, "pair_a")
    dense = disparity.dim() == 3
    if dense:
        F, H, W = (int(x) for x in disparity.shape)
    else:
        if src_idx is None:
            raise ValueError("per-keypoint disparity needs src_idx (gather_matches_batch)")
        if int(disparity.shape[1]) != n_max:
            raise ValueError("disparity must be [n_frames, n_max]")
        src_idx = _chk(src_idx, torch.int32, "src_idx")
        F, H, W = int(disparity.shape[0]), 0, 0
    p3 = torch.empty((P, n_max, 3), dtype=torch.float32, device=dev)  # rows >= n_out[p] are never read
    p2 = torch.empty((P, n_max, 2), dtype=torch.float32, device=dev)
    no = torch.empty((P,), dtype=torch.int32, device=dev)
    cam = camera.as_struct()
    B = camera.baseline if baseline is None else float(baseline)
    _lib.check(lib.fs_backproject_filter_bank(_p(mkpts0), _p(mkpts1), _p(n_corr), P, n_max, _p(disparity), F, H, W,
                                              _p(pa), _p(src_idx), C.byref(cam), B, zmin, zmax, _p(p3), _p(p2),
                                              _p(no), _stream()), "fs_backproject_filter_bank")
    return {"points3d": p3, "pts2d": p2, "n_out": no}


def filter_matches_distance(matches0: torch.Tensor, dist0: torch.Tensor, max_distance: int) -> torch.Tensor:
    """In place: matches0[p, q] = -1 where dist0[p, q] > max_distance (ORBmatcher.cc:704 style gate) → n_matches [P]."""
    lib = _lib.load()
    matches0 = _chk(matches0, torch.int32, "matches0")
    dist0 = _chk(dist0, torch.int32, "dist0")
    P, n_max = int(matches0.shape[0]), int(matches0.shape[1])
    nm = torch.empty((P,), dtype=torch.int32, device=matches0.device)
    _lib.check(lib.fs_filter_matches_distance(_p(matches0), _p(dist0), P, n_max, int(max_distance), _p(nm), _stream()),
               "fs_filter_matches_distance")
    return nm


def pack_records(n_matches, n_inliers, status, R, t, n_good=None, iters=None) -> torch.Tensor:
    """[P, 20] f64 per-pair result records (n_matches, n_inliers, status, n_good, R, t, iters) for the one gather."""
    lib = _lib.load()
    R = _chk(R, torch.float64, "R")
    t = _chk(t, torch.float64, "t")
    P = int(R.shape[0])
    i32 = lambda x, n: None if x is None else _chk(x, torch.int32, n)
    rec = torch.empty((P, 20), dtype=torch.float64, device=R.device)
    _lib.check(lib.fs_pack_records(_p(i32(n_matches, "n_matches")), _p(i32(n_inliers, "n_inliers")),
                                   _p(i32(status, "status")), _p(i32(n_good, "n_good")), _p(R), _p(t),
                                   _p(i32(iters, "iters")), P, _p(rec), _stream()), "fs_pack_records")
    return rec


# ------------------------------------------------------------------------------------------- A4 / A5
def find_essential_batch(mkpts0, mkpts1, n_corr, *, focal: float, pp: Tuple[float, float], prob: float = 0.999,
                         threshold: float = 1.0, max_iters: int = 1000):
    """Batched cv2.findEssentialMat(..., method=RANSAC) (mono_slam.py:111)."""
    lib = _lib.load()
    mkpts0 = _chk(mkpts0, torch.float32, "mkpts0")
    mkpts1 = _chk(mkpts1, torch.float32, "mkpts1")
    n_corr = _chk(n_corr, torch.int32, "n_corr")
    dev = mkpts0.device
    P, n_max = int(mkpts0.shape[0]), int(mkpts0.shape[1])
    E = torch.empty((P, 3, 3), dtype=torch.float64, device=dev)  # the kernel writes every output of every pair
    mask = torch.empty((P, n_max), dtype=torch.uint8, device=dev)
    ninl = torch.empty((P,), dtype=torch.int32, device=dev)
    iters = torch.empty((P,), dtype=torch.int32, device=dev)
    status = torch.empty((P,), dtype=torch.int32, device=dev)
    nbytes = lib.fs_find_essential_workspace_bytes(P, n_max, max_iters)
    ws = _workspace(nbytes, dev, "ransac")
    _lib.check(lib.fs_find_essential_ransac(_p(mkpts0), _p(mkpts1), _p(n_corr), P, n_max, float(focal),
                                            float(pp[0]), float(pp[1]), float(prob), float(threshold),
                                            int(max_iters), _p(E), _p(mask), _p(ninl), _p(iters), _p(status),
                                            _p(ws), ws.numel(), _stream()), "fs_find_essential_ransac")
    return {"E": E, "mask": mask, "n_inliers": ninl, "iters": iters, "status": status}


def recover_pose_batch(E, mkpts0, mkpts1, n_corr, *, focal: float, pp: Tuple[float, float],
                       distance_thresh: float = 50.0, status=None):
    """Batched cv2.recoverPose(E, p0, p1, focal=, pp=) (mono_slam.py:112)."""
    lib = _lib.load()
    E = _chk(E, torch.float64, "E")
    mkpts0 = _chk(mkpts0, torch.float32, "mkpts0")
    mkpts1 = _chk(mkpts1, torch.float32, "mkpts1")
    n_corr = _chk(n_corr, torch.int32, "n_corr")
    dev = mkpts0.device
    P, n_max = int(mkpts0.shape[0]), int(mkpts0.shape[1])
    R = torch.empty((P, 3, 3), dtype=torch.float64, device=dev)
    t = torch.empty((P, 3), dtype=torch.float64, device=dev)
    mask = torch.empty((P, n_max), dtype=torch.uint8, device=dev)
    ng = torch.empty((P,), dtype=torch.int32, device=dev)
    ws = _workspace(lib.fs_recover_pose_workspace_bytes(P), dev, "recover_pose")
    _lib.check(lib.fs_recover_pose(_p(E), _p(mkpts0), _p(mkpts1), _p(n_corr), _p(status), P, n_max,
                                   float(focal), float(pp[0]), float(pp[1]), float(distance_thresh), _p(R),
                                   _p(t), _p(mask), _p(ng), _p(ws), ws.numel(), _stream()), "fs_recover_pose")
    return {"R": R, "t": t, "mask": mask, "n_good": ng}


def ransac_subsets(n_corr, max_iters: int) -> torch.Tensor:
    """OpenCV's RANSAC sampler (RNG(2^64-1), 5 distinct indices per iteration) → [P, max_iters, 5] int32."""
    lib = _lib.load()
    dev = n_corr.device if isinstance(n_corr, torch.Tensor) and n_corr.is_cuda else _dev()
    n_corr = _to_cuda(n_corr, torch.int32, dev)
    P = int(n_corr.numel())
    out = torch.zeros((P, max_iters, 5), dtype=torch.int32, device=dev)
    _lib.check(lib.fs_ransac_subsets(_p(n_corr), P, int(max_iters), _p(out), _stream()), "fs_ransac_subsets")
    return out


def score_essential_batch(mkpts0, mkpts1, n_corr, models_E, *, focal, pp, threshold: float = 1.0):
    """Sampson-score [P, H, 3, 3] hypotheses against every correspondence → counts [P, H] (config 4 sweep)."""
    lib = _lib.load()
    mkpts0 = _chk(mkpts0, torch.float32, "mkpts0")
    mkpts1 = _chk(mkpts1, torch.float32, "mkpts1")
    n_corr = _chk(n_corr, torch.int32, "n_corr")
    models_E = _chk(models_E, torch.float64, "models_E")
    P, n_max = int(mkpts0.shape[0]), int(mkpts0.shape[1])
    H = int(models_E.shape[1])
    counts = torch.zeros((P, H), dtype=torch.int32, device=mkpts0.device)
    _lib.check(lib.fs_score_essential(_p(mkpts0), _p(mkpts1), _p(n_corr), P, n_max, float(focal), float(pp[0]),
                                      float(pp[1]), float(threshold), _p(models_E), H, _p(counts), _stream()),
               "fs_score_essential")
    return counts


def score_pnp_batch(points3d, pts2d, n_corr, models_rt, *, camera: ForestCamera = ForestCamera(),
                    reproj_error: float = 1.0):
    """Reprojection-score [P, H, 6] (rvec|tvec) hypotheses → counts [P, H] (config 4 sweep)."""
    lib = _lib.load()
    points3d = _chk(points3d, torch.float32, "points3d")
    pts2d = _chk(pts2d, torch.float32, "pts2d")
    n_corr = _chk(n_corr, torch.int32, "n_corr")
    models_rt = _chk(models_rt, torch.float64, "models_rt")
    P, n_max = int(points3d.shape[0]), int(points3d.shape[1])
    H = int(models_rt.shape[1])
    counts = torch.zeros((P, H), dtype=torch.int32, device=points3d.device)
    cam = camera.as_struct()
    _lib.check(lib.fs_score_pnp(_p(points3d), _p(pts2d), _p(n_corr), P, n_max, C.byref(cam), float(reproj_error),
                                _p(models_rt), H, _p(counts), _stream()), "fs_score_pnp")
    return counts


# ------------------------------------------------------------------------------------------- A7
def solve_pnp_ransac_batch(points3d, pts2d, n_corr, *, camera: ForestCamera = ForestCamera(),
                           reproj_error: float = 1.0, confidence: float = 0.99, max_iters: int = 1000,
                           min_points: int = 6):
    """Batched cv2.solvePnPRansac(..., SOLVEPNP_ITERATIVE) with the >= 6 gate of stereo_slam.py:292-295."""
    lib = _lib.load()
    points3d = _chk(points3d, torch.float32, "points3d")
    pts2d = _chk(pts2d, torch.float32, "pts2d")
    n_corr = _chk(n_corr, torch.int32, "n_corr")
    dev = points3d.device
    P, n_max = int(points3d.shape[0]), int(points3d.shape[1])
    rvec = torch.empty((P, 3), dtype=torch.float64, device=dev)
    tvec = torch.empty((P, 3), dtype=torch.float64, device=dev)
    mask = torch.empty((P, n_max), dtype=torch.uint8, device=dev)
    idx = torch.empty((P, n_max), dtype=torch.int32, device=dev)   # first n_inliers[p] entries are valid
    ninl = torch.empty((P,), dtype=torch.int32, device=dev)
    iters = torch.empty((P,), dtype=torch.int32, device=dev)
    status = torch.empty((P,), dtype=torch.int32, device=dev)
    cam = camera.as_struct()
    nbytes = lib.fs_solve_pnp_workspace_bytes(P, n_max, max_iters)
    ws = _workspace(nbytes, dev, "ransac")
    _lib.check(lib.fs_solve_pnp_ransac(_p(points3d), _p(pts2d), _p(n_corr), P, n_max, C.byref(cam),
                                       float(reproj_error), float(confidence), int(max_iters), int(min_points),
                                       _p(rvec), _p(tvec), _p(mask), _p(idx), _p(ninl), _p(iters), _p(status),
                                       _p(ws), ws.numel(), _stream()), "fs_solve_pnp_ransac")
    return {"rvec": rvec, "tvec": tvec, "mask": mask, "inlier_idx": idx, "n_inliers": ninl, "iters": iters,
            "status": status}


# ------------------------------------------------------------------------------------------- A8
def rodrigues_batch(rvec: torch.Tensor) -> torch.Tensor:
    lib = _lib.load()
    rvec = _chk(rvec, torch.float64, "rvec").reshape(-1, 3)
    R = torch.empty((rvec.shape[0], 3, 3), dtype=torch.float64, device=rvec.device)
    _lib.check(lib.fs_rodrigues(_p(rvec), int(rvec.shape[0]), _p(R), _stream()), "fs_rodrigues")
    return R


def compose_trajectory(R: torch.Tensor, t: torch.Tensor, status: Optional[torch.Tensor] = None,
                       T_init: Optional[np.ndarray] = None) -> torch.Tensor:
    """T_cum[k] = T_init @ T_0 @ … @ T_k with T_k = [R_k | t_k] (mono_slam.py:115-118, stereo_slam.py:301-306)."""
    lib = _lib.load()
    R = _chk(R, torch.float64, "R")
    t = _chk(t, torch.float64, "t")
    n = int(R.shape[0])
    out = torch.empty((max(n, 1), 4, 4), dtype=torch.float64, device=R.device)
    T0 = None
    if T_init is not None:
        T0 = np.ascontiguousarray(T_init, dtype=np.float64)
        T0p = T0.ctypes.data_as(C.c_void_p)
    else:
        T0p = None
    st = None if status is None else _chk(status, torch.int32, "status")
    _lib.check(lib.fs_compose_trajectory(_p(R), _p(t), _p(st), n, T0p, _p(out), _stream()),
               "fs_compose_trajectory")
    return out[:n]


def transform_points_batch(points3d: torch.Tensor, n_pts: torch.Tensor, T: torch.Tensor) -> torch.Tensor:
    """T_cum @ [P; 1] for every pair (stereo_slam.py:309-311)."""
    lib = _lib.load()
    points3d = _chk(points3d, torch.float32, "points3d")
    n_pts = _chk(n_pts, torch.int32, "n_pts")
    T = _chk(T, torch.float64, "T")
    P, n_max = int(points3d.shape[0]), int(points3d.shape[1])
    out = torch.zeros_like(points3d)
    _lib.check(lib.fs_transform_points(_p(points3d), _p(n_pts), P, n_max, _p(T), _p(out), _stream()),
               "fs_transform_points")
    return out


# =========================================================================================== cv2-shaped
class BFMatcher:
    """Drop-in for ``cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True)`` (stereo_slam.py:85) on uint8 [N, 32]
    descriptors, and for ``NORM_L2`` on float32 [N, 256] unit descriptors."""

    def __init__(self, normType: int = NORM_HAMMING, crossCheck: bool = False):
        if normType not in (NORM_HAMMING, NORM_L2):
            raise ValueError("only NORM_HAMMING and NORM_L2 are on the hot path")
        self.normType = normType
        self.crossCheck = bool(crossCheck)

    def _run(self, d0, d1, ratio=None):
        dev = _dev()
        d0 = np.asarray(d0)
        d1 = np.asarray(d1)
        n0, n1 = d0.shape[0], d1.shape[0]
        if n0 == 0 or n1 == 0:
            return None, n0
        if self.normType == NORM_HAMMING:
            n_max = max(n0, n1)
            bank = torch.zeros((2, n_max, 32), dtype=torch.uint8, device=dev)
            bank[0, :n0] = _to_cuda(d0, torch.uint8, dev)
            bank[1, :n1] = _to_cuda(d1, torch.uint8, dev)
            counts = torch.tensor([n0, n1], dtype=torch.int32, device=dev)
            r = hamming_match_batch(bank, counts, [0], [1], cross_check=self.crossCheck, ratio=ratio)
            return (r["matches0"][0, :n0].cpu().numpy(), r["dist0"][0, :n0].cpu().numpy().astype(np.float32),
                    r["second0"][0, :n0].cpu().numpy().astype(np.float32)), n0
        n_max = (max(n0, n1) + 127) // 128 * 128
        bank = torch.zeros((2, 256, n_max), dtype=torch.float32, device=dev)
        bank[0, :, :n0] = _to_cuda(d0, torch.float32, dev).T
        bank[1, :, :n1] = _to_cuda(d1, torch.float32, dev).T
        counts = torch.tensor([n0, n1], dtype=torch.int32, device=dev)
        r = float_match_batch(bank, counts, [0], [1], ratio=ratio, cross_check=self.crossCheck)
        m0 = r["matches0"][0, :n0].cpu().numpy()
        cos = r["score0"][0, :n0].cpu().numpy()
        cos2 = r["second0"][0, :n0].cpu().numpy()
        d = np.sqrt(np.maximum(0.0, 2.0 - 2.0 * cos)).astype(np.float32)
        d2 = np.sqrt(np.maximum(0.0, 2.0 - 2.0 * cos2)).astype(np.float32)
        return (m0, d, d2), n0

    def match(self, queryDescriptors, trainDescriptors):
        """→ tuple of DMatch sorted by queryIdx, like cv2 (stereo_slam.py:234)."""
        res, n0 = self._run(queryDescriptors, trainDescriptors)
        if res is None:
            return ()
        m0, d, _ = res
        return tuple(DMatch(int(q), int(m0[q]), float(d[q])) for q in range(n0) if m0[q] >= 0)

    def ratioMatch(self, queryDescriptors, trainDescriptors, ratio: float = 0.7):
        """knnMatch(k=2) + Lowe ratio in one call (ORB_SLAM3/src/Frame.cc:1144-1151 semantics)."""
        res, n0 = self._run(queryDescriptors, trainDescriptors, ratio=ratio)
        if res is None:
            return ()
        m0, d, _ = res
        return tuple(DMatch(int(q), int(m0[q]), float(d[q])) for q in range(n0) if m0[q] >= 0)


def match_mutual_nn(descriptors0, descriptors1, *, scale: float = 1.0, threshold: Optional[float] = None,
                    ratio: Optional[float] = None, exact_fp32: bool = False, split3: bool = False,
                    want_diag: bool = False):
    """Tail of SuperGlue.forward on the raw contraction (superglue.py:259-260, :268-278).

    descriptors0 [256, N0], descriptors1 [256, N1] float32 (SuperPoint layout) →
    dict(matches0, matches1, matching_scores0, matching_scores1) with -1 for unmatched, like the reference.
    RAW-SCORE CONVENTION (differs from SuperGlue's output dict): `threshold` is compared with the scaled score itself and
    matching_scores are the scaled scores, not exp(score); a caller replacing superglue.py:273-278 passes
    threshold=log(match_threshold) (bench.py: log 0.2) and applies exp() to the scores it keeps."""
    dev = _dev()
    d0 = np.asarray(descriptors0, dtype=np.float32)
    d1 = np.asarray(descriptors1, dtype=np.float32)
    n0, n1 = d0.shape[1], d1.shape[1]
    if n0 == 0 or n1 == 0:  # superglue.py:235-242
        return {"matches0": np.full((n0,), -1, np.int64), "matches1": np.full((n1,), -1, np.int64),
                "matching_scores0": np.zeros((n0,), np.float32), "matching_scores1": np.zeros((n1,), np.float32)}
    n_max = (max(n0, n1) + 127) // 128 * 128
    bank = torch.zeros((2, 256, n_max), dtype=torch.float32, device=dev)
    bank[0, :, :n0] = _to_cuda(d0, torch.float32, dev)
    bank[1, :, :n1] = _to_cuda(d1, torch.float32, dev)
    counts = torch.tensor([n0, n1], dtype=torch.int32, device=dev)
    r = float_match_batch(bank, counts, [0], [1], scale=scale, threshold=threshold, ratio=ratio,
                          exact_fp32=exact_fp32, split3=split3, want_diag=want_diag)
    m0 = r["matches0"][0, :n0].cpu().numpy().astype(np.int64)
    m1 = r["matches1"][0, :n1].cpu().numpy().astype(np.int64)
    s0 = r["score0"][0, :n0].cpu().numpy()
    ms0 = np.where(m0 >= 0, s0, 0.0).astype(np.float32)
    ms1 = np.zeros((n1,), np.float32)
    ms1[m0[m0 >= 0]] = ms0[m0 >= 0]
    out = {"matches0": m0, "matches1": m1, "matching_scores0": ms0, "matching_scores1": ms1}
    if want_diag and "diag" in r:
        out["diag"] = r["diag"].cpu().numpy()
    return out


def _corr_batch(p0, p1, dev):
    p0 = np.ascontiguousarray(np.asarray(p0, dtype=np.float32).reshape(-1, 2))
    p1 = np.ascontiguousarray(np.asarray(p1, dtype=np.float32).reshape(-1, 2))
    if p0.shape != p1.shape:
        raise ValueError("point arrays must have the same shape")
    n = p0.shape[0]
    t0 = _to_cuda(p0, torch.float32, dev)[None]
    t1 = _to_cuda(p1, torch.float32, dev)[None]
    if n == 0:
        t0 = torch.zeros((1, 1, 2), dtype=torch.float32, device=dev)
        t1 = torch.zeros((1, 1, 2), dtype=torch.float32, device=dev)
    return t0, t1, torch.tensor([n], dtype=torch.int32, device=dev), n


def findEssentialMat(points1, points2, focal: float = 1.0, pp=(0.0, 0.0), method: int = RANSAC,
                     prob: float = 0.999, threshold: float = 1.0, maxIters: int = 1000):
    """cv2.findEssentialMat(p0, p1, focal=, pp=, method=cv2.RANSAC, prob=, threshold=) (mono_slam.py:111).
    → (E [3,3] float64, mask [N,1] uint8 in {0,1}); (None, None) below 5 points, like cv2."""
    if method != RANSAC:
        raise ValueError("only method=cv2.RANSAC is on the hot path")
    dev = _dev()
    t0, t1, nc, n = _corr_batch(points1, points2, dev)
    if n < 5:
        return None, None
    r = find_essential_batch(t0, t1, nc, focal=focal, pp=pp, prob=prob, threshold=threshold, max_iters=maxIters)
    if int(r["status"][0].item()) != 0:
        return None, None
    return r["E"][0].cpu().numpy(), r["mask"][0, :n].cpu().numpy().reshape(-1, 1)


def recoverPose(E, points1, points2, focal: float = 1.0, pp=(0.0, 0.0), distanceThresh: float = 50.0):
    """cv2.recoverPose(E, p0, p1, focal=, pp=) (mono_slam.py:112) → (n_good, R, t [3,1], mask [N,1] in {0,255})."""
    dev = _dev()
    t0, t1, nc, n = _corr_batch(points1, points2, dev)
    Et = _to_cuda(np.asarray(E, dtype=np.float64).reshape(1, 3, 3), torch.float64, dev)
    r = recover_pose_batch(Et, t0, t1, nc, focal=focal, pp=pp, distance_thresh=distanceThresh)
    return (int(r["n_good"][0].item()), r["R"][0].cpu().numpy(), r["t"][0].cpu().numpy().reshape(3, 1),
            r["mask"][0, :n].cpu().numpy().reshape(-1, 1))


def solvePnPRansac(objectPoints, imagePoints, cameraMatrix, distCoeffs, reprojectionError: float = 8.0,
                   confidence: float = 0.99, iterationsCount: int = 100, flags: int = SOLVEPNP_ITERATIVE):
    """cv2.solvePnPRansac(P3, p2, K, dist, reprojectionError=, confidence=, iterationsCount=, flags=ITERATIVE)
    (stereo_slam.py:294-295) → (ok, rvec [3,1], tvec [3,1], inliers [I,1] int32)."""
    if flags != SOLVEPNP_ITERATIVE:
        raise ValueError("only SOLVEPNP_ITERATIVE is on the hot path")
    dev = _dev()
    P3 = np.ascontiguousarray(np.asarray(objectPoints, dtype=np.float32).reshape(-1, 3))
    p2 = np.ascontiguousarray(np.asarray(imagePoints, dtype=np.float32).reshape(-1, 2))
    n = P3.shape[0]
    if n != p2.shape[0]:
        raise ValueError("objectPoints / imagePoints length mismatch")
    if n < 4:
        raise ValueError("solvePnPRansac needs at least 4 points (cv2 asserts npoints >= 4)")
    cam = camera_from_K(cameraMatrix, distCoeffs)
    r = solve_pnp_ransac_batch(_to_cuda(P3, torch.float32, dev)[None], _to_cuda(p2, torch.float32, dev)[None],
                               torch.tensor([n], dtype=torch.int32, device=dev), camera=cam,
                               reproj_error=reprojectionError, confidence=confidence, max_iters=iterationsCount,
                               min_points=5)
    ok = int(r["status"][0].item()) == 0
    ni = int(r["n_inliers"][0].item())
    inl = r["inlier_idx"][0, :ni].cpu().numpy().astype(np.int32).reshape(-1, 1) if ok else None
    return ok, r["rvec"][0].cpu().numpy().reshape(3, 1), r["tvec"][0].cpu().numpy().reshape(3, 1), inl


def Rodrigues(rvec):
    """cv2.Rodrigues(rvec) (stereo_slam.py:298) → (R [3,3], None)."""
    dev = _dev()
    r = _to_cuda(np.asarray(rvec, dtype=np.float64).reshape(1, 3), torch.float64, dev)
    return rodrigues_batch(r)[0].cpu().numpy(), None


# ------------------------------------------------------------------------------------------- SURVEY §8(b) names
# The scope table names the drop-in functions in snake_case; they are the cv2-named wrappers above, argument for argument.
def bf_match_hamming(d0, d1, *, cross_check: bool = True, ratio: Optional[float] = None):
    """stereo_slam.py:85, :234 → (queryIdx [K] int32, trainIdx [K] int32, distance [K] float32), ascending queryIdx."""
    bf = BFMatcher(NORM_HAMMING, crossCheck=cross_check)
    ms = bf.match(d0, d1) if ratio is None else bf.ratioMatch(d0, d1, ratio)
    return (np.array([m.queryIdx for m in ms], np.int32), np.array([m.trainIdx for m in ms], np.int32),
            np.array([m.distance for m in ms], np.float32))


def find_essential_mat(p0, p1, *, focal: float, pp, prob: float = 0.999, threshold: float = 1.0,
                       max_iters: int = 1000):
    """mono_slam.py:111 → (E, mask)."""
    return findEssentialMat(p0, p1, focal=focal, pp=pp, method=RANSAC, prob=prob, threshold=threshold,
                            maxIters=max_iters)


def recover_pose(E, p0, p1, *, focal: float, pp, distance_thresh: float = 50.0):
    """mono_slam.py:112 → (n, R, t, mask)."""
    return recoverPose(E, p0, p1, focal=focal, pp=pp, distanceThresh=distance_thresh)


def solve_pnp_ransac(P3, p2, K, dist, *, reprojection_error: float = 1.0, confidence: float = 0.99,
                     iterations_count: int = 1000):
    """stereo_slam.py:294-295 → (ok, rvec, tvec, inliers)."""
    return solvePnPRansac(P3, p2, K, dist, reprojectionError=reprojection_error, confidence=confidence,
                          iterationsCount=iterations_count, flags=SOLVEPNP_ITERATIVE)


# ------------------------------------------------------------------------------------------- §8f: guided ORB matching
def orb_search_for_initialization(kpts, octave, desc, counts, pair_a, pair_b, prev_matched, *, angle=None,
                                  window_size: int = 100, nn_ratio: float = 0.9, check_orientation: bool = True,
                                  bounds=(0.0, 960.0, 0.0, 600.0)):
    """ORBmatcher(nn_ratio, check_orientation).SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
    (ORB_SLAM3/src/ORBmatcher.cc:648-762; Tracking.cc:2491-2492 uses 0.9, true, 100) for a batch of frame pairs.
    prev_matched [P, n_max, 2] is updated in place.  → dict(matches12 [P, n_max], n_matches [P], prev_matched)."""
    lib = _lib.load()
    kpts = _chk(kpts, torch.float32, "kpts")
    dev = kpts.device
    F, n_max = int(kpts.shape[0]), int(kpts.shape[1])
    octave = _chk(octave, torch.int32, "octave")
    desc = _chk(desc, torch.uint8, "desc")
    counts = _chk(counts.to(dev), torch.int32, "counts")
    if angle is not None:
        angle = _chk(angle, torch.float32, "angle")
    a, b = _pairs(pair_a, pair_b, dev)
    P = int(a.numel())
    prev_matched = _chk(prev_matched, torch.float32, "prev_matched")
    if tuple(prev_matched.shape) != (P, n_max, 2):
        raise ValueError("prev_matched must be [n_pairs, n_max, 2]")
    m12 = torch.empty((P, n_max), dtype=torch.int32, device=dev)
    nm = torch.empty((P,), dtype=torch.int32, device=dev)
    _lib.check(lib.fs_orb_search_for_initialization(_p(kpts), _p(angle), _p(octave), _p(desc), F, n_max, _p(counts),
                                                    _p(a), _p(b), P, _p(prev_matched), int(window_size),
                                                    float(nn_ratio), int(bool(check_orientation)), float(bounds[0]),
                                                    float(bounds[1]), float(bounds[2]), float(bounds[3]), _p(m12),
                                                    _p(nm), _stream()), "fs_orb_search_for_initialization")
    return {"matches12": m12, "n_matches": nm, "prev_matched": prev_matched}


def orb_stereo_matches(kpts_l, octave_l, desc_l, counts_l, kpts_r, octave_r, desc_r, counts_r, image_l, image_r, *,
                       camera: ForestCamera = ForestCamera(), baseline: Optional[float] = None):
    """Frame::ComputeStereoMatches (ORB_SLAM3/src/Frame.cc:811-981) for a batch of stereo frames: row-band Hamming search,
    SAD sub-pixel refinement on the image pair, median test.  → dict(u_right, depth, sad, disparity); `disparity`
    (uL - uR, 0 where unmatched) is what the A6 back-projection of the stereo path reads per keypoint."""
    lib = _lib.load()
    kpts_l = _chk(kpts_l, torch.float32, "kpts_l")
    kpts_r = _chk(kpts_r, torch.float32, "kpts_r")
    dev = kpts_l.device
    F, n_max = int(kpts_l.shape[0]), int(kpts_l.shape[1])
    if tuple(kpts_r.shape) != tuple(kpts_l.shape):
        raise ValueError("left and right keypoint banks must have the same [F, n_max, 2] shape")
    image_l = _chk(image_l, torch.uint8, "image_l")
    image_r = _chk(image_r, torch.uint8, "image_r")
    rows, cols = int(image_l.shape[1]), int(image_l.shape[2])
    B = camera.baseline if baseline is None else float(baseline)
    ur = torch.empty((F, n_max), dtype=torch.float32, device=dev)
    dp = torch.empty_like(ur)
    sad = torch.empty((F, n_max), dtype=torch.int32, device=dev)
    _lib.check(lib.fs_orb_stereo_matches(_p(kpts_l), _p(_chk(octave_l, torch.int32, "octave_l")),
                                         _p(_chk(desc_l, torch.uint8, "desc_l")),
                                         _p(_chk(counts_l.to(dev), torch.int32, "counts_l")), _p(kpts_r),
                                         _p(_chk(octave_r, torch.int32, "octave_r")),
                                         _p(_chk(desc_r, torch.uint8, "desc_r")),
                                         _p(_chk(counts_r.to(dev), torch.int32, "counts_r")), F, n_max, _p(image_l),
                                         _p(image_r), rows, cols, float(B), float(camera.fx * B), _p(ur), _p(dp),
                                         _p(sad), _stream()), "fs_orb_stereo_matches")
    disp = torch.where(ur >= 0, kpts_l[..., 0] - ur, torch.zeros_like(ur))
    return {"u_right": ur, "depth": dp, "sad": sad, "disparity": disp}


def render_forest(T_wc, trunks: torch.Tensor, *, camera: ForestCamera = ForestCamera(), baseline: Optional[float] = None,
                  ground_y: float = 1.2, width: int = 960, height: int = 600, want_disparity: bool = True):
    """CUDA ray caster (csrc/render.cu): grey image [H, W] u8 of the procedural forest (ground plane + trunk cylinders,
    solid hashed texture) from camera pose T_wc, and the true dense disparity map [H, W] f32."""
    lib = _lib.load()
    trunks = _chk(trunks, torch.float64, "trunks")
    dev = trunks.device
    T = np.ascontiguousarray(T_wc, dtype=np.float64)
    img = torch.empty((height, width), dtype=torch.uint8, device=dev)
    disp = torch.empty((height, width), dtype=torch.float32, device=dev) if want_disparity else None
    cam = camera.as_struct()
    B = camera.baseline if baseline is None else float(baseline)
    _lib.check(lib.fs_render_forest(T.ctypes.data_as(C.c_void_p), C.byref(cam), B, float(ground_y), _p(trunks),
                                    int(trunks.shape[0]), int(width), int(height), _p(img), _p(disp), _stream()),
               "fs_render_forest")
    return img, disp
