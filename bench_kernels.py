#!/usr/bin/env python
"""Per-kernel roofline table: every kernel of the path timed ALONE at a saturating batch size.

Not the contract bench (that is bench.py); this answers BASELINE.json's "each choice is evidenced ... achieved HBM GB/s
for the Hamming, scoring and triangulation kernels against B200 peak" with one JSON line per kernel:
  {"kernel", "config", "ms", "algorithmic": {...work figures of DESIGN.md §4...}, "bound", "achieved", "peak", "frac"}
Timing: CUDA events on the launching stream, 3 warm-ups, median of `--reps`, a 256 MB write between launches (L2 flush).
Peaks: MEASURED_PEAKS.json (HBM, bf16 tensor) when present; CUDA-core pipe peaks (fp64 without FMA, popc, MUFU ex2) are
the measured rates in profiles/pipe_peaks.json (scratch/pipe_peaks.cu run on a B200 of this pool).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from forest_slam_b200 import api as fsb  # noqa: E402
from forest_slam_b200 import synth  # noqa: E402

SMS = 148


def peaks():
    p = {"hbm_gbs": 6500.0, "tensor_tflops": 1600.0, "source": "fallback (B200_PROFILING.md)"}
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            j = json.load(open(path))
            flat = json.dumps(j)
            def find(d, keys):
                for k, v in (d.items() if isinstance(d, dict) else []):
                    if isinstance(v, (int, float)) and all(s in k.lower() for s in keys):
                        return float(v)
                    if isinstance(v, dict):
                        r = find(v, keys)
                        if r:
                            return r
                return None
            h = find(j, ("hbm", "burst")) or find(j, ("hbm",)) or find(j, ("copy",))
            t = find(j, ("bf16", "burst")) or find(j, ("bf16",)) or find(j, ("tflop",))
            if h:
                p["hbm_gbs"] = h
            if t:
                p["tensor_tflops"] = t
            p["source"] = "measured (MEASURED_PEAKS.json)"
            del flat
        except Exception:
            pass
    return p


def pipe_peaks():
    """Measured CUDA-core pipe rates (profiles/pipe_peaks.json, scratch/pipe_peaks.cu on a B200 of this pool)."""
    path = os.path.join(ROOT, "profiles", "pipe_peaks.json")
    if os.path.exists(path):
        j = json.load(open(path))
        j["source"] = "measured (profiles/pipe_peaks.json)"
        return j
    return {"fp64_tflops_no_fma": 18.6, "fp64_tflops_fma": 37.2, "popc_gops": 4653.0, "ex2_gops": 4653.0,
            "source": "nominal (148 SMs x lanes x 1.965 GHz)"}


def sm_clock_ghz() -> float:
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(0)
        return pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM) / 1000.0
    except Exception:
        return 1.965


class Timer:
    def __init__(self, reps: int):
        self.reps = reps
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def __call__(self, fn) -> float:
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(self.reps):
            self.flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return statistics.median(ts)


def line(kernel, config, ms, alg, bound, achieved, peak, unit, note=""):
    out = {"kernel": kernel, "config": config, "ms": round(ms, 5), "algorithmic": alg, "bound": bound,
           "achieved": round(achieved, 3), "peak": round(peak, 1), "unit": unit, "frac": round(achieved / peak, 4)}
    if note:
        out["note"] = note
    print(json.dumps(out), flush=True)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=15)
    ap.add_argument("--pairs", type=int, default=256)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("bench_kernels.py needs a GPU (there is no CPU path)")
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    pk = peaks()
    pipes = pipe_peaks()
    ghz = sm_clock_ghz()
    T = Timer(args.reps)
    P = args.pairs
    want = set(args.only.split(",")) if args.only else None
    g = torch.Generator(device="cpu").manual_seed(7)

    def on(name):
        return want is None or name in want

    print(json.dumps({"peaks": pk, "pipe_peaks": pipes, "sm_clock_ghz": ghz, "sms": SMS,
                      "device": torch.cuda.get_device_name(0)}))

    # ------------------------------------------------------------------ A2 Hamming (POPC pipe; HBM figure asked too)
    if on("hamming"):
        for n in (2000, 2048):
            F = P + 1
            desc = torch.randint(0, 256, (F, n, 32), dtype=torch.uint8, generator=g).to(dev)
            counts = torch.full((F,), n, dtype=torch.int32, device=dev)
            a = torch.arange(P, dtype=torch.int32, device=dev)
            b = a + 1
            ms = T(lambda: fsb.hamming_match_batch(desc, counts, a, b, popc=True))
            popc = P * n * n * 8.0
            byts = P * ((n + n) * 32 + n * 12.0)
            peak_popc = pipes["popc_gops"]  # measured G popc32/s (16 lanes/clk/SM)
            line("k_hamming_top2 (+finalize)", {"pairs": P, "n0": n, "n1": n, "crossCheck": True}, ms,
                 {"popc32": popc, "bytes": byts, "hbm_gbs": byts / ms / 1e6}, "popc pipe",
                 popc / ms / 1e6, peak_popc, "Gpopc32/s",
                 "both directions from ONE distance matrix; HBM fraction = %.4f of %.0f GB/s (operands live on chip)"
                 % (byts / ms / 1e6 / pk["hbm_gbs"], pk["hbm_gbs"]))

            # tensor-core variant: the same N0 x N1 x 256 bit comparisons as a +-1 fp16 contraction (hamming_tc.cu)
            ops = fsb.hamming_prepare(desc, counts)
            ms_p = T(lambda: fsb.hamming_prepare(desc, counts))
            ms_t = T(lambda: fsb.hamming_match_batch(None, counts, a, b, operands=ops))
            flop = P * 2.0 * n * n * 256
            line("fs_hamming_match_tc (k_float_top3_tc on +-1 operands + finalize)",
                 {"pairs": P, "n0": n, "n1": n, "crossCheck": True}, ms_t,
                 {"flop": flop, "popc32_equivalent": popc, "equiv_gpopc32_per_s": popc / ms_t / 1e6,
                  "speedup_vs_popc_kernel": ms / ms_t, "prepare_ms_per_%d_frames" % F: ms_p},
                 "tensor", flop / ms_t / 1e9, pk["tensor_tflops"], "TFLOP/s",
                 "single fp16 tcgen05 pass (crossCheck from the slab maxima of the column side); exact integers")

    # ------------------------------------------------------------------ config 5: all-pairs loop-closure matching
    if on("loopclosure"):
        from forest_slam_b200 import sequence as seq
        F, n = 320, 2000  # 51 040 keyframe pairs (the 2000-keyframe job is 1 999 000 pairs of the same shape)
        desc = torch.randint(0, 256, (F, n, 32), dtype=torch.uint8, generator=g).to(dev)
        counts = torch.full((F,), n, dtype=torch.int32, device=dev)
        kp = (torch.rand((F, n, 2), generator=g) * torch.tensor([950.0, 590.0])).to(dev)
        bank = seq.FrameBank(kpts=kp, counts=counts, desc_bin=desc)
        npairs = F * (F - 1) // 2
        for tc in (True, False):
            fsb.TENSOR_CORE_MATCHER = tc
            # gate above n: matching + counting only (random descriptors give ~n/2 chance mutual matches per pair, so a
            # 30-match gate would send every pair through a 1000-iteration E-RANSAC on garbage and time only that)
            ms = Timer(3)(lambda: seq.loop_closure_pairs(bank, descriptor="binary", min_matches=n + 1))
            flop = npairs * 2.0 * n * n * 256
            line("loop_closure_pairs (all-pairs Hamming crossCheck, %s)" % ("tensor cores" if tc else "POPC pipe"),
                 {"keyframes": F, "kpts": n, "pairs": npairs, "descriptors": "random; verification gate closed (matching only)"},
                 ms, {"flop": flop, "pairs_per_s": npairs / ms * 1e3,
                      "est_s_for_2000_keyframes_1gpu": 1999000 / (npairs / ms * 1e3)},
                 "tensor", flop / ms / 1e9, pk["tensor_tflops"], "TFLOP/s",
                 "whole driver (pair index generation, chunked launches, record assembly), one GPU")
        fsb.TENSOR_CORE_MATCHER = True

    # ------------------------------------------------------------------ A6 back-projection (HBM, 25 B/point)
    if on("backproject"):
        for npts in (1 << 20, 1 << 24):
            pts = (torch.rand((npts, 2), generator=g) * torch.tensor([950.0, 590.0])).to(dev)
            disp = (torch.rand((npts,), generator=g) * 60.0 + 0.5).to(dev)
            ms = T(lambda: fsb.backproject_stereo(pts, disp))
            byts = npts * 25.0
            line("k_backproject", {"points": npts}, ms, {"bytes": byts}, "hbm", byts / ms / 1e6, pk["hbm_gbs"],
                 "GB/s", "12 B in (x, y, d) + 13 B out (XYZ f32 + valid u8) per point; includes two torch.empty")

    # ------------------------------------------------------------------ A1 operand preparation (HBM)
    if on("prepare"):
        n = 2048
        F = P + 1
        d = torch.randn((F, 256, n), generator=g)
        d = (d / d.norm(dim=1, keepdim=True)).to(dev)
        counts = torch.full((F,), n, dtype=torch.int32, device=dev)
        ms = T(lambda: fsb.float_prepare(d, counts))
        byts = F * n * 256 * 8.0
        line("k_float_prepare", {"frames": F, "n": n}, ms, {"bytes": byts}, "hbm", byts / ms / 1e6, pk["hbm_gbs"],
             "GB/s", "4 B in + 2 x 2 B out (hi | lo fp16) per descriptor component")
        # ------------------------------------------------------------------ A1 matcher at config 2 and config 3
        for (n, pp) in ((2048, P), (4096, 64)):
            F = pp + 1
            d = torch.randn((F, 256, n), generator=g)
            d = (d / d.norm(dim=1, keepdim=True)).to(dev)
            counts = torch.full((F,), n, dtype=torch.int32, device=dev)
            ops = fsb.float_prepare(d, counts)
            a = torch.arange(pp, dtype=torch.int32, device=dev)
            b = a + 1
            ms = T(lambda: fsb.float_match_batch(None, counts, a, b, operands=ops))
            fl = pp * 2.0 * n * n * 256
            line("fs_float_match (single-pass top3_tc + colmerge + resolve + slabscan + colcheck + finalize)", {"pairs": pp, "n": n,
                 "descriptors": "random unit vectors (worst case for certification: no inlier margin)"}, ms,
                 {"flop": fl}, "tensor", fl / ms / 1e9, pk["tensor_tflops"], "TFLOP/s")
            ms3 = T(lambda: fsb.float_match_batch(None, counts, a, b, operands=ops, split3=True))
            line("fs_float_match_split3 (k_float_match_tc)", {"pairs": pp, "n": n}, ms3, {"flop": fl}, "tensor",
                 fl / ms3 / 1e9, pk["tensor_tflops"], "TFLOP/s", "3 MMAs per k-step (fp32-grade scores)")
            del d, ops

    # ------------------------------------------------------------------ A4 / A7 scoring sweep (config 4)
    if on("score"):
        K = 1200
        cam = fsb.ForestCamera()
        rng = np.random.default_rng(3)
        Pn = 8
        x0 = rng.uniform([0, 0], [960, 600], (Pn, K, 2)).astype(np.float32)
        x1 = (x0 + rng.normal(0, 3.0, x0.shape)).astype(np.float32)
        P3 = rng.uniform([-10, -3, 2], [10, 3, 40], (Pn, K, 3)).astype(np.float32)
        nc = torch.full((Pn,), K, dtype=torch.int32, device=dev)
        x0t, x1t, P3t = (torch.from_numpy(v).to(dev) for v in (x0, x1, P3))
        fp64_peak = pipes["fp64_tflops_no_fma"]  # measured DMUL + DADD rate (the scorers run without FMA contraction)
        for H in (1024, 2048, 4096, 8192, 16384, 32768, 65536):
            E = torch.from_numpy(rng.normal(0, 1, (Pn, H, 3, 3))).to(dev)
            ms = T(lambda: fsb.score_essential_batch(x0t, x1t, nc, E, focal=cam.fx, pp=(cam.cx, cam.cy)))
            ev = Pn * H * K * 1.0
            byts = Pn * (K * 16 + H * 72 + H * 4.0)
            line("k_score_essential", {"pairs": Pn, "H": H, "K": K}, ms,
                 {"evals": ev, "flop": ev * 44, "bytes": byts, "hbm_gbs": byts / ms / 1e6,
                  "gevals_per_s": ev / ms / 1e6}, "fp64 pipe", ev * 44 / ms / 1e9, fp64_peak, "TFLOP/s",
                 "Sampson error in OpenCV's operation order, fp64, no FMA contraction (44 flop/eval)")
            rt = torch.from_numpy(np.concatenate([rng.normal(0, 0.05, (Pn, H, 3)),
                                                  rng.normal(0, 0.3, (Pn, H, 3))], axis=2)).to(dev)
            ms = T(lambda: fsb.score_pnp_batch(P3t, x0t, nc, rt))
            byts = Pn * (K * 20 + H * 48 + H * 4.0)
            line("k_score_pnp", {"pairs": Pn, "H": H, "K": K}, ms,
                 {"evals": ev, "flop": ev * 40, "bytes": byts, "hbm_gbs": byts / ms / 1e6,
                  "gevals_per_s": ev / ms / 1e6}, "fp64 pipe", ev * 40 / ms / 1e9, fp64_peak, "TFLOP/s",
                 "projectPoints with k1,k2,p1,p2,k3 in fp64 (40 flop/eval), float32 compare")
            del E, rt

    # ------------------------------------------------------------------ A3 gather (HBM)
    if on("gather"):
        n = 2048
        F = P + 1
        kp = (torch.rand((F, n, 2), generator=g) * 900).to(dev)
        counts = torch.full((F,), n, dtype=torch.int32, device=dev)
        a = torch.arange(P, dtype=torch.int32, device=dev)
        b = a + 1
        m0 = torch.randint(-1, n, (P, n), dtype=torch.int32, generator=g).to(dev)
        ms = T(lambda: fsb.gather_matches_batch(kp, counts, a, b, m0))
        kk = float((m0 >= 0).sum())
        byts = P * n * 4.0 + kk * 36.0
        line("k_gather_matches", {"pairs": P, "n": n}, ms, {"bytes": byts}, "hbm", byts / ms / 1e6, pk["hbm_gbs"],
             "GB/s", "latency-bound at this size (a few MB per launch)")

    # ------------------------------------------------------------------ guided ORB matching + ray caster (§8f ranks 1, 4)
    if on("orb"):
        from tests import util as tu
        cam = fsb.ForestCamera()
        world = synth.make_world(synth.SEED0)
        # SearchForInitialization: 256 temporal pairs of 2000 keypoints, window 100
        Pn, n = min(P, 256), 2000
        frames = synth.make_sequence(Pn + 1, n, world=world)
        kp = torch.from_numpy(np.stack([f.kpts for f in frames])).to(dev)
        desc = torch.from_numpy(np.stack([f.desc_bin for f in frames])).to(dev)
        ang = torch.from_numpy(np.stack([synth.keypoint_angles(world.points.shape[0], f.landmark, i)
                                         for i, f in enumerate(frames)])).to(dev)
        octv = torch.zeros((Pn + 1, n), dtype=torch.int32, device=dev)
        counts = torch.full((Pn + 1,), n, dtype=torch.int32, device=dev)
        a = torch.arange(Pn, dtype=torch.int32, device=dev)
        b = a + 1
        prev0 = kp[:Pn].clone()
        r = fsb.orb_search_for_initialization(kp, octv, desc, counts, a, b, prev0.clone(), angle=ang)
        ms = T(lambda: fsb.orb_search_for_initialization(kp, octv, desc, counts, a, b, prev0.clone(), angle=ang))
        cand = Pn * n * (200.0 * 200.0) / (960.0 * 600.0) * n
        line("k_orb_search_init (ORBmatcher::SearchForInitialization, window 100)", {"pairs": Pn, "kpts": n}, ms,
             {"candidate_distances": cand, "popc32": cand * 8, "pairs_per_s": Pn / ms * 1e3,
              "mean_matches": float(r["n_matches"].float().mean())}, "latency (sequential greedy loop per pair)",
             cand * 8 / ms / 1e6, pipes["popc_gops"], "Gpopc32/s",
             "one warp per frame pair walks the queries IN ORDER (later queries depend on earlier assignments); "
             "the batch supplies the parallelism")
        # ray caster: stereo pair at 960 x 600, 400 trunks
        tr = torch.from_numpy(world.trunks).to(dev)
        T_l = frames[3].T_wc
        ms = T(lambda: fsb.render_forest(T_l, tr, ground_y=world.ground_y))
        px = 960 * 600.0
        line("k_render_forest (ray caster, 400 trunk cylinders + ground plane)", {"image": "960x600"}, ms,
             {"ray_cylinder_tests": px * 400, "flop": px * 400 * 14, "bytes_out": px * 5}, "fp64 pipe",
             px * 400 * 14 / ms / 1e9, pipes["fp64_tflops_no_fma"], "TFLOP/s",
             "one thread per pixel, every trunk tested (fp64, no FMA: bit-identical to the NumPy restatement)")
        # ComputeStereoMatches on 16 rendered stereo frames
        Fn = 16
        imL = torch.empty((Fn, 600, 960), dtype=torch.uint8, device=dev)
        imR = torch.empty_like(imL)
        kL, kR, dL, dR = [], [], [], []
        for f in range(Fn):
            T_l = frames[2 + f].T_wc
            T_r = synth.right_pose(T_l)
            imL[f] = fsb.render_forest(T_l, tr, ground_y=world.ground_y, want_disparity=False)[0]
            imR[f] = fsb.render_forest(T_r, tr, ground_y=world.ground_y, want_disparity=False)[0]
            fl = synth.observe(world, T_l, n, 500 + f, border=16)
            fr = synth.observe(world, T_r, n, 900 + f, border=16)
            kL.append(fl.kpts); kR.append(fr.kpts); dL.append(fl.desc_bin); dR.append(fr.desc_bin)
        tt = lambda x, dt: torch.from_numpy(np.stack(x)).to(dev).to(dt)
        kL, kR, dL, dR = tt(kL, torch.float32), tt(kR, torch.float32), tt(dL, torch.uint8), tt(dR, torch.uint8)
        oz = torch.zeros((Fn, n), dtype=torch.int32, device=dev)
        cn = torch.full((Fn,), n, dtype=torch.int32, device=dev)
        r = fsb.orb_stereo_matches(kL, oz, dL, cn, kR, oz, dR, cn, imL, imR)
        ms = T(lambda: fsb.orb_stereo_matches(kL, oz, dL, cn, kR, oz, dR, cn, imL, imR))
        nm = float((r["u_right"] >= 0).float().sum(1).mean())
        byts = Fn * (2 * n * (8 + 32 + 4) + n * 12.0)
        line("k_orb_stereo + k_orb_stereo_median (Frame::ComputeStereoMatches)", {"frames": Fn, "kpts": n}, ms,
             {"row_band_tests": Fn * n * n * 1.0, "bytes": byts, "frames_per_s": Fn / ms * 1e3, "mean_matched": nm},
             "hbm", byts / ms / 1e6, pk["hbm_gbs"], "GB/s",
             "one warp per left keypoint: row-band scan of the right keypoints, 11 x 11 SAD x 11 shifts, parabola fit; "
             "latency-bound at this size")

    # ------------------------------------------------------------------ SuperGlue score head (Sinkhorn, §8f rank 2)
    if on("sinkhorn"):
        for (m, n, pp) in ((1024, 1024, 64), (2048, 2048, 16)):
            sc = (torch.randn((pp, m, n), generator=g) * 2.0).to(dev)
            cm = torch.full((pp,), m, dtype=torch.int32, device=dev)
            cn = torch.full((pp,), n, dtype=torch.int32, device=dev)
            iters = 20
            ms = T(lambda: fsb.sinkhorn_match_batch(sc, cm, cn, iters=iters))
            ex2 = pp * m * n * 2.0 * iters
            byts = pp * m * n * 4.0
            peak_mufu = pipes["ex2_gops"]  # measured G ex2/s (16 MUFU lanes/clk/SM)
            line("k_sk_resident (cooperative, 1 launch per group)", {"pairs": pp, "m": m, "n": n, "iters": iters}, ms,
                 {"ex2": ex2, "bytes_hbm_once": byts, "bytes_streamed": byts * (iters + 1),
                  "stream_gbs": byts * (iters + 1) / ms / 1e6, "pairs_per_s": pp / ms * 1e3}, "mufu pipe",
                 ex2 / ms / 1e6, peak_mufu, "Gex2/s",
                 "2 exp per matrix element per iteration (row and column log-sum-exp); "
                 "resident kernel: the matrix is read from HBM once and stays in shared memory for all "
                 "iterations; two pair-wide barriers per iteration")
            ms_s = T(lambda: fsb.sinkhorn_match_batch(sc, cm, cn, iters=iters, pairs_per_group=-1))
            line("k_sk_band x (iters + 1) [streaming variant]", {"pairs": pp, "m": m, "n": n, "iters": iters}, ms_s,
                 {"ex2": ex2, "bytes_streamed": byts * (iters + 1), "stream_gbs": byts * (iters + 1) / ms_s / 1e6,
                  "pairs_per_s": pp / ms_s * 1e3}, "mufu pipe", ex2 / ms_s / 1e6, peak_mufu, "Gex2/s",
                 "one launch per iteration, matrices streamed from L2 (groups of pairs sized to 96 MB)")


if __name__ == "__main__":
    main()
