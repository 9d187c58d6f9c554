"""Batched drivers of the hot path: what the frame loops of mono_slam.py:85-182 and stereo_slam.py:177-345 do per
frame pair, run for a whole list of frame pairs in a handful of launches, and sharded by frame-pair index across the
ranks of a torch.distributed job (one process per GPU) with ONE gather of fixed-size result records at the end
(the path has no other cross-GPU dependency).

Everything here enqueues CUDA work through the C ABI (api.py); there is no CPU path.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, Optional

import torch

from . import api
from .api import ForestCamera

RECORD_F64 = 20  # n_matches, n_inliers, status, n_good, R[9], t[3], pad -> 20 doubles = 160 B per frame pair


@dataclasses.dataclass
class FrameBank:
    """Device-resident frame bank (SURVEY.md §8a 'types on the path'): everything the hot path reads."""
    kpts: torch.Tensor                      # [F, n_max, 2] f32 pixels
    counts: torch.Tensor                    # [F] i32
    desc_f32: Optional[torch.Tensor] = None  # [F, 256, n_max] f32 (SuperPoint layout)
    desc_bin: Optional[torch.Tensor] = None  # [F, n_max, 32] u8 (ORB layout)
    disparity: Optional[torch.Tensor] = None  # [F, n_max] f32 per-keypoint disparity (uL - uR), or the frames' dense
    #                                           disparity maps [F, H, W] f32 (SGBM / 16, stereo_slam.py:262-271)
    operands: Optional[torch.Tensor] = None  # [F, n_max + 1, 512] fp16 operand bank (api.float_prepare) of the tensor-core matcher

    @staticmethod
    def from_numpy(kpts, counts, desc_f32=None, desc_bin=None, disparity=None, device="cuda") -> "FrameBank":
        def dev(a, dt):
            return None if a is None else torch.as_tensor(a, dtype=dt).to(device, non_blocking=True).contiguous()
        return FrameBank(dev(kpts, torch.float32), dev(counts, torch.int32), dev(desc_f32, torch.float32),
                         dev(desc_bin, torch.uint8), dev(disparity, torch.float32))


def shard_pairs(n_pairs: int, rank: int, world: int) -> torch.Tensor:
    """Pair indices owned by `rank`: i mod world == rank (balances varying correspondence counts)."""
    return torch.arange(rank, n_pairs, world, dtype=torch.int64)


class StageTimer:
    """Optional CUDA-event bracket around every stage of a batched driver (bench.py reads it for the roofline).
    Events are recorded on torch's current stream — the stream every launch of api.py is enqueued on."""

    def __init__(self):
        self.spans = []  # (name, start_event, end_event)

    def run(self, name, fn):
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        out = fn()
        b.record()
        self.spans.append((name, a, b))
        return out

    def totals_ms(self) -> Dict[str, float]:
        tot: Dict[str, float] = {}
        for name, a, b in self.spans:
            tot[name] = tot.get(name, 0.0) + a.elapsed_time(b)
        return tot


def _run(timer, name, fn):
    return fn() if timer is None else timer.run(name, fn)


def _match(bank: FrameBank, pa, pb, descriptor: str, threshold, ratio, exact_fp32: bool, scale: float = 1.0,
           timer=None, split3: bool = False):
    if descriptor == "float":
        if bank.desc_f32 is None and bank.operands is None:
            raise ValueError("the frame bank holds no float descriptors")
        operands = bank.operands
        n_max = int(bank.desc_f32.shape[2]) if bank.desc_f32 is not None else int(bank.operands.shape[1]) - 1
        if operands is None and not exact_fp32 and api.TENSOR_CORE_MATCHER and n_max % 128 == 0:
            # fp32 channel-major descriptors -> split-fp16 operand rows, once per frame (not per pair)
            operands = _run(timer, "prepare", lambda: api.float_prepare(bank.desc_f32, bank.counts))
        r = _run(timer, "match", lambda: api.float_match_batch(bank.desc_f32, bank.counts, pa, pb, scale=scale,
                                                               threshold=threshold, ratio=ratio, operands=operands,
                                                               exact_fp32=exact_fp32, split3=split3, want_diag=True))
    elif descriptor == "binary":
        if bank.desc_bin is None:
            raise ValueError("the frame bank holds no binary descriptors")
        n_max = int(bank.desc_bin.shape[1])
        ops = None
        if api.TENSOR_CORE_MATCHER and n_max <= api.HAMMING_TC_MAX_COLS:
            # descriptor bits -> +-1 fp16 operand rows of the tensor-core Hamming matcher, once per frame
            ops = _run(timer, "prepare", lambda: api.hamming_prepare(bank.desc_bin, bank.counts))
        r = _run(timer, "match", lambda: api.hamming_match_batch(bank.desc_bin, bank.counts, pa, pb, cross_check=True,
                                                                 ratio=ratio, want_matches1=False, operands=ops))
    else:
        raise ValueError("descriptor must be 'float' or 'binary'")
    return r


def mono_pairs(bank: FrameBank, pair_a, pair_b, *, descriptor: str = "float", camera: ForestCamera = ForestCamera(),
               threshold: Optional[float] = None, ratio: Optional[float] = None, prob: float = 0.999,
               ransac_threshold: float = 1.0, max_iters: int = 1000, exact_fp32: bool = False, split3: bool = False, scale: float = 1.0,
               timer: Optional[StageTimer] = None) -> Dict[str, torch.Tensor]:
    """mono_slam.py:100-118 for every (pair_a[i], pair_b[i]): match -> gather -> findEssentialMat(RANSAC) -> recoverPose.
    The single `focal` for both axes and recoverPose without the RANSAC mask are the reference's behaviour."""
    dev = bank.kpts.device
    pa = api._to_cuda(pair_a, torch.int32, dev)
    pb = api._to_cuda(pair_b, torch.int32, dev)
    m = _match(bank, pa, pb, descriptor, threshold, ratio, exact_fp32, scale, timer, split3)
    g = _run(timer, "gather", lambda: api.gather_matches_batch(bank.kpts, bank.counts, pa, pb, m["matches0"]))
    pp = (camera.cx, camera.cy)
    e = _run(timer, "essential_ransac",
             lambda: api.find_essential_batch(g["mkpts0"], g["mkpts1"], g["n_corr"], focal=camera.fx, pp=pp, prob=prob,
                                              threshold=ransac_threshold, max_iters=max_iters))
    r = _run(timer, "recover_pose",
             lambda: api.recover_pose_batch(e["E"], g["mkpts0"], g["mkpts1"], g["n_corr"], focal=camera.fx, pp=pp,
                                            status=e["status"]))
    extra = {"match_diag": m["diag"]} if "diag" in m else {}
    return {**extra, "matches0": m["matches0"], "n_matches": g["n_corr"], "mkpts0": g["mkpts0"], "mkpts1": g["mkpts1"],
            "E": e["E"], "inlier_mask": e["mask"], "n_inliers": e["n_inliers"], "status": e["status"],
            "iters": e["iters"], "R": r["R"], "t": r["t"], "pose_mask": r["mask"], "n_good": r["n_good"]}


def stereo_pairs(bank: FrameBank, pair_a, pair_b, *, descriptor: str = "binary", camera: ForestCamera = ForestCamera(),
                 threshold: Optional[float] = None, ratio: Optional[float] = None, reproj_error: float = 1.0,
                 confidence: float = 0.99, max_iters: int = 1000, exact_fp32: bool = False, split3: bool = False, scale: float = 1.0,
                 timer: Optional[StageTimer] = None) -> Dict[str, torch.Tensor]:
    """stereo_slam.py:210-306 for every pair: temporal match (left) -> gather -> depth back-projection of the previous
    frame's keypoints + joint depth filter -> solvePnPRansac (>= 6 gate) -> Rodrigues."""
    if bank.disparity is None:
        raise ValueError("the frame bank holds no disparity")
    dev = bank.kpts.device
    pa = api._to_cuda(pair_a, torch.int32, dev)
    pb = api._to_cuda(pair_b, torch.int32, dev)
    m = _match(bank, pa, pb, descriptor, threshold, ratio, exact_fp32, scale, timer, split3)
    g = _run(timer, "gather", lambda: api.gather_matches_batch(bank.kpts, bank.counts, pa, pb, m["matches0"]))
    # disparity of the previous-left keypoint: looked up inside the kernel, per keypoint or in the frame's dense map
    # at (int(y), int(x)) as the reference does (stereo_slam.py:274-281)
    b = _run(timer, "triangulate",
             lambda: api.backproject_filter_bank(g["mkpts0"], g["mkpts1"], g["n_corr"], bank.disparity, pa,
                                                 g["src_idx"], camera=camera))
    s = _run(timer, "pnp_ransac",
             lambda: api.solve_pnp_ransac_batch(b["points3d"], b["pts2d"], b["n_out"], camera=camera,
                                                reproj_error=reproj_error, confidence=confidence, max_iters=max_iters,
                                                min_points=6))
    R = api.rodrigues_batch(s["rvec"])
    return {"matches0": m["matches0"], "n_matches": g["n_corr"], "points3d": b["points3d"], "pts2d": b["pts2d"],
            "n_points": b["n_out"], "rvec": s["rvec"], "tvec": s["tvec"], "R": R, "t": s["tvec"],
            "inlier_mask": s["mask"], "inlier_idx": s["inlier_idx"], "n_inliers": s["n_inliers"],
            "status": s["status"], "iters": s["iters"]}


def pack_records(res: Dict[str, torch.Tensor]) -> torch.Tensor:
    """Fixed-size per-pair result record [P, 20] f64: n_matches, n_inliers, status, n_good, R(9), t(3), iters, pad
    (one launch of k_pack_records)."""
    P = res["R"].shape[0]
    return api.pack_records(res["n_matches"], res["n_inliers"], res["status"], res["R"].reshape(P, 9),
                            res["t"].reshape(P, 3), n_good=res.get("n_good"), iters=res.get("iters"))


def gather_records(rec_local: torch.Tensor, n_pairs_total: int) -> torch.Tensor:
    """The one collective of the path: all ranks' records -> [n_pairs_total, 20] in pair-index order on every rank
    (NCCL all_gather over NVLink; pair i lives on rank i mod W at slot i // W)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return rec_local
    W = dist.get_world_size()
    per = (n_pairs_total + W - 1) // W
    if rec_local.shape[0] == per:
        pad = rec_local
    else:
        pad = torch.zeros((per, rec_local.shape[1]), dtype=rec_local.dtype, device=rec_local.device)
        pad[: rec_local.shape[0]] = rec_local
    out = torch.empty((W, per, rec_local.shape[1]), dtype=rec_local.dtype, device=rec_local.device)
    dist.all_gather_into_tensor(out.view(W * per, -1), pad)
    # [W, per] -> pair index i = slot * W + rank
    return out.permute(1, 0, 2).reshape(W * per, -1)[:n_pairs_total]


def trajectory(records: torch.Tensor) -> torch.Tensor:
    """T_cum = T_cum @ T_rel over the gathered records (mono_slam.py:115-118, stereo_slam.py:301-306)."""
    P = records.shape[0]
    R = records[:, 4:13].reshape(P, 3, 3).contiguous()
    t = records[:, 13:16].contiguous()
    status = records[:, 2].to(torch.int32).contiguous()
    return api.compose_trajectory(R, t, status)


class CapturedPipeline:
    """A batched driver captured once in a CUDA graph and replayed per step.

    One step of `mono_pairs` / `stereo_pairs` is ~25 launches of a few microseconds to a few milliseconds; issuing them
    from Python costs more than several of the kernels take, so the launch sequence (including the record gather, which
    NCCL supports under capture) is recorded once on static buffers and replayed.  The frame bank tensors are the
    graph's inputs: refresh them in place (`bank.kpts.copy_(...)`) and call `run()`; the returned dict / record tensor
    are static too and are overwritten by every replay.
    """

    def __init__(self, bank: FrameBank, pair_a, pair_b, kind: str = "mono", n_pairs_total: Optional[int] = None,
                 warmup: int = 3, gather: bool = True, **kw):
        """gather=False keeps the collective out of the graph (records stay local): needed when several pipelines replay
        concurrently on different streams under torch.distributed — collectives of one communicator must be issued in
        the same order on every rank, which concurrent graph replays cannot promise."""
        if kind not in ("mono", "stereo"):
            raise ValueError("kind must be 'mono' or 'stereo'")
        dev = bank.kpts.device
        self.bank = bank
        self.pa = api._to_cuda(pair_a, torch.int32, dev)
        self.pb = api._to_cuda(pair_b, torch.int32, dev)
        fn = mono_pairs if kind == "mono" else stereo_pairs
        n_total = int(n_pairs_total if n_pairs_total is not None else self.pa.numel())

        scope = api.workspace_scope(f"pipeline-{id(self)}")  # private workspaces: pipelines may replay concurrently

        def body():
            with scope:
                res = fn(self.bank, self.pa, self.pb, **kw)
                rec = pack_records(res)
                return res, (gather_records(rec, n_total) if gather else rec)

        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):  # sizes every workspace / shared-memory opt-in before capture
                body()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.results, self.records = body()

    def run(self):
        self.graph.replay()
        return self.results, self.records


# ------------------------------------------------------------------------------------------------ loop closure (config 5)
def triangular_index(n_frames: int, lin: torch.Tensor):
    """Linear index of the strict upper triangle (row-major: (0,1), (0,2), ..., (0,n-1), (1,2), ...) -> (a, b), a < b."""
    lin = lin.to(torch.int64)
    n = int(n_frames)
    # row a starts at s(a) = a * (2n - a - 1) / 2; invert with the quadratic formula in float64, then repair by +-1
    a = torch.floor(((2 * n - 1) - torch.sqrt(((2 * n - 1) ** 2 - 8.0 * lin.double()).clamp(min=0))) / 2).to(torch.int64)
    start = lambda r: r * (2 * n - r - 1) // 2
    a = torch.where(start(a) > lin, a - 1, a)
    a = torch.where(start(a + 1) <= lin, a + 1, a)
    b = lin - start(a) + a + 1
    return a, b


def shard_all_pairs(n_frames: int, rank: int, world: int, block: int = 1024) -> torch.Tensor:
    """Linear triangular indices owned by `rank`: blocks of `block` consecutive pairs dealt round-robin, so that every rank
    touches every keyframe about equally."""
    total = n_frames * (n_frames - 1) // 2
    lin = torch.arange(total, dtype=torch.int64)
    return lin[(lin // block) % world == rank]


def loop_closure_pairs(bank: FrameBank, *, descriptor: str = "binary", min_matches: int = 30, max_distance: int = 64,
                       rank: int = 0,
                       world: int = 1, chunk: int = 4096, block: int = 1024, camera: ForestCamera = ForestCamera(),
                       prob: float = 0.999, ransac_threshold: float = 1.0, max_iters: int = 1000,
                       threshold: Optional[float] = None, scale: float = 1.0) -> torch.Tensor:
    """All-pairs keyframe matching: every (a, b), a < b, of the bank is matched (mutual nearest neighbour); pairs with
    more than `min_matches` mutual matches go through essential-matrix RANSAC.  Binary descriptors: a mutual match
    counts only when its Hamming distance is <= `max_distance` (ORBmatcher.cc:704 style gate, 0 = off) — crossCheck alone
    pairs about n/2 random descriptors of two unrelated keyframes.  This rank's share of the pairs only.
    Returns records [n_local, 5] int64 on the device: linear pair index, a, b, n_matches, n_inliers (-1: not verified)."""
    dev = bank.kpts.device
    F = int(bank.kpts.shape[0])
    lin_all = shard_all_pairs(F, rank, world, block)
    out = torch.empty((lin_all.numel(), 5), dtype=torch.int64, device=dev)
    operands = bank.operands
    if descriptor == "float" and operands is None:
        operands = api.float_prepare(bank.desc_f32, bank.counts)
    ops_bin = None
    if descriptor == "binary" and api.TENSOR_CORE_MATCHER and int(bank.desc_bin.shape[1]) <= api.HAMMING_TC_MAX_COLS:
        ops_bin = api.hamming_prepare(bank.desc_bin, bank.counts)  # once per keyframe, reused by all its pairs
    pp = (camera.cx, camera.cy)
    for o in range(0, lin_all.numel(), chunk):
        lin = lin_all[o:o + chunk]
        a, b = triangular_index(F, lin)
        pa, pb = a.to(torch.int32).to(dev), b.to(torch.int32).to(dev)
        if descriptor == "binary":
            m = api.hamming_match_batch(bank.desc_bin, bank.counts, pa, pb, cross_check=True, want_matches1=False,
                                        operands=ops_bin)
            if max_distance > 0:
                m["n_matches"] = api.filter_matches_distance(m["matches0"], m["dist0"], max_distance)
        else:
            m = api.float_match_batch(bank.desc_f32, bank.counts, pa, pb, scale=scale, threshold=threshold,
                                      operands=operands)
        nm = m["n_matches"].to(torch.int64)
        rec = out[o:o + lin.numel()]
        rec[:, 0] = lin.to(dev)
        rec[:, 1] = a.to(dev)
        rec[:, 2] = b.to(dev)
        rec[:, 3] = nm
        rec[:, 4] = -1
        sel = torch.nonzero(nm > min_matches).flatten()
        if sel.numel():
            g = api.gather_matches_batch(bank.kpts, bank.counts, pa[sel], pb[sel], m["matches0"][sel].contiguous())
            e = api.find_essential_batch(g["mkpts0"], g["mkpts1"], g["n_corr"], focal=camera.fx, pp=pp, prob=prob,
                                         threshold=ransac_threshold, max_iters=max_iters)
            rec[sel, 4] = torch.where(e["status"] == 0, e["n_inliers"].to(torch.int64), torch.full_like(sel, -1))
    return out


def gather_loop_records(rec_local: torch.Tensor, n_frames: int) -> torch.Tensor:
    """One all_gather of the loop-closure records of all ranks, restored to linear pair order."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return rec_local[torch.argsort(rec_local[:, 0])]
    W = dist.get_world_size()
    total = n_frames * (n_frames - 1) // 2
    n_loc = torch.tensor([rec_local.shape[0]], dtype=torch.int64, device=rec_local.device)
    n_max = n_loc.clone()
    dist.all_reduce(n_max, op=dist.ReduceOp.MAX)
    per = int(n_max.item())
    pad = torch.full((per, rec_local.shape[1]), -1, dtype=rec_local.dtype, device=rec_local.device)
    pad[: rec_local.shape[0]] = rec_local
    out = torch.empty((W * per, rec_local.shape[1]), dtype=rec_local.dtype, device=rec_local.device)
    dist.all_gather_into_tensor(out, pad)
    out = out[out[:, 0] >= 0]
    assert out.shape[0] == total
    return out[torch.argsort(out[:, 0])]
