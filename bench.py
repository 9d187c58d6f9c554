#!/usr/bin/env python
"""bench.py — frame-pairs/s of the Forest-SLAM matching + pose hot path on N B200s, next to the host-CPU reference.

Contract (driver): `python bench.py --gpus N --steps K --warmup W [--impl reference]` prints ONE JSON line on rank 0.
A step = one pass of the hot path over one batch of synthetic frame pairs.  The headline workload is BASELINE.json
configs[1] (mono_slam.py temporal matching + essential-matrix RANSAC + recoverPose at 2048 keypoints/frame); the same
line carries, under `workloads`, the other BASELINE configs measured in the same run:
  stereo      config 1: stereo_slam.py path (Hamming match + depth back-projection + PnP RANSAC, 2000 kpts/frame)
  float4096   config 3: float 256-d mutual-NN + ratio test at 4096 kpts/frame, 64 pairs per launch
  pnp_sweep   config 4: back-projection + PnP RANSAC with H = 1k .. 64k hypotheses per frame pair (adaptive stop off)
  loopclosure config 5: all pairs of 2000 keyframes (1 999 000 pairs) sharded over the ranks, STRONG scaling
(`--workload X` runs one of them alone; `--headline-only` skips the extras).
value   = pairs/s with the frame bank resident in HBM, CUDA events, max over ranks;
e2e     = the same through the public batched API with HOST (pinned) buffers: H2D of the frame bank and D2H of the
          per-pair result records inside the timed region;
roofline= the dominant kernel's algorithmic work / its CUDA-event time vs MEASURED_PEAKS.json (tensor, HBM) or
          profiles/pipe_peaks.json (measured fp64 / popc / MUFU pipe rates, scratch/pipe_peaks.cu);
cpu_baseline = the reference's own cv2/torch-CPU calls (oracle/cv_literal.py) timed on a bounded sample on rank 0;
parity  = every pair of the timed sequence (rank 0's) compared with oracle/cv_literal.py outside the timed region.
"""
from __future__ import annotations

import argparse
import copy
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "frame-pairs/sec (match+RANSAC+triangulate)"
UNIT = "frame-pairs/s"
LN_THRESH = math.log(0.2)  # superglue.py:275 exp(score) > 0.2 on the scaled score
SEQ_PAIRS = 256            # frame pairs per synthetic sequence (a batch of more pairs is made of several sequences)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="mono", choices=["mono", "stereo", "float4096", "pnp_sweep", "loopclosure"],
                    help="mono = the headline (BASELINE configs[1]); a plain `bench.py` run also carries the other "
                         "workloads under `workloads` unless --headline-only")
    ap.add_argument("--headline-only", action="store_true", help="skip the `workloads` object (tuning runs)")
    ap.add_argument("--no-verify", action="store_true", help="skip the parity leg (all pairs of the timed sequence "
                                                            "against oracle/cv_literal.py, outside the timed region)")
    ap.add_argument("--keyframes", type=int, default=2000, help="loopclosure: keyframes (config 5: 2000)")
    ap.add_argument("--pairs", type=int, default=512, help="frame pairs per step per GPU (sequences of 256 pairs)")
    ap.add_argument("--kpts", type=int, default=0, help="keypoints per frame (default 2048 mono / 2000 stereo)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="pairs timed for cpu_baseline (0 = auto)")
    ap.add_argument("--exact-fp32", action="store_true", help="CUDA-core fp32 matcher instead of tcgen05")
    ap.add_argument("--split3", action="store_true", help="split-fp16 three-MMA matcher instead of the certified pass")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--lanes", type=int, default=5,
                    help="captured pipelines (own buffers, own stream) the device loop rotates through when --overlap-steps")
    ap.add_argument("--overlap-steps", type=int, default=1,
                    help="1: consecutive steps alternate between streams (software pipelining across steps)")
    ap.add_argument("--pos-outlier", type=float, default=-1.0,
                    help="probability that an observation is a gross positional outlier (default 0.1 mono, 0 stereo)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------ workload
def make_workload(args, rank: int):
    """Seeded synthetic sequence (procedural forest): `pairs + 1` frames of `kpts` keypoints for this rank."""
    from forest_slam_b200 import synth
    n_kpts = args.kpts or (2048 if args.workload == "mono" else 2000)
    world = synth.make_world(synth.SEED0)
    # every observation is a gross positional outlier with probability 0.1 => ~19 % corrupted matches per pair
    # (mono = BASELINE config 2, a RANSAC benchmark; the stereo sequence of config 1 is left as generated)
    # a step's batch = consecutive frame pairs of one or more independent sequences (SEQ_PAIRS pairs each: the forest is
    # 80 m deep and the rig moves 0.13 m per frame, so one sequence stays inside it for ~450 frames)
    frames, pair_a, pair_b = [], [], []
    left, k = args.pairs, 0
    while left > 0:
        n = min(left, SEQ_PAIRS)
        base = len(frames)
        frames += synth.make_sequence(n + 1, n_kpts, seed=synth.SEED0 + 1000 * rank + 37 * k, world=world,
                                      pos_outlier=pos_outlier(args))
        pair_a += list(range(base, base + n))
        pair_b += list(range(base + 1, base + n + 1))
        left -= n
        k += 1
    kpts, dbin, df32, disp, counts = synth.stack_frames(frames)
    return {"frames": frames, "kpts": kpts, "desc_bin": dbin, "desc_f32": df32, "disparity": disp, "counts": counts,
            "n_kpts": n_kpts, "pair_a": np.array(pair_a, np.int32), "pair_b": np.array(pair_b, np.int32),
            "n_sequences": k}


def pos_outlier(args) -> float:
    if args.pos_outlier >= 0:
        return args.pos_outlier
    return 0.1 if args.workload == "mono" else 0.0


def config_dict(args, wl, world_size):
    name = ("mono_slam.py temporal matching (float 256-d mutual-NN) + essential-matrix RANSAC + recoverPose"
            if args.workload == "mono" else
            "stereo_slam.py temporal matching (256-bit Hamming crossCheck) + depth back-projection + PnP RANSAC")
    return {"workload": name, "kpts_per_frame": int(wl["n_kpts"]), "pairs_per_step_per_gpu": int(args.pairs),
            "sequences_per_step": int(wl["n_sequences"]),
            "image": "960x600", "gross_outlier_prob_per_observation": pos_outlier(args),
            "sharding": f"frame-pair index over {world_size} rank(s), one all_gather of records",
            "l2": "L2 flushed (256 MB write) between timed iterations; frame bank per step %.0f MB" % (
                bank_bytes(args, wl) / 1e6),
            "steps_in_flight": (max(2, args.lanes) if args.overlap_steps else 1)}


def bank_bytes(args, wl):
    F = len(wl["frames"])
    n = wl["n_kpts"]
    per = n * 8 + (n * 256 * 4 if args.workload == "mono" else n * 32 + n * 4)
    return F * per


def load_peaks():
    """Roofline denominators: MEASURED_PEAKS.json (driver-written: HBM, bf16 tensor) and profiles/pipe_peaks.json
    (measured on a B200 of this pool with scratch/pipe_peaks.cu: fp64 / popc / MUFU pipe rates)."""
    pk = {"hbm_gbs": 6650.0, "bf16_tflops": 1600.0, "bf16_tflops_sustained": 1400.0,
          "source": "fallback (B200_PROFILING.md)"}
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        pk.update({k: v for k, v in json.load(open(p)).items() if isinstance(v, (int, float))})
        pk["source"] = "measured (MEASURED_PEAKS.json)"
    pp = os.path.join(ROOT, "profiles", "pipe_peaks.json")
    pk["pipes"] = {"fp64_tflops_no_fma": 18.6, "fp64_tflops_fma": 37.2, "popc_gops": 4653.0, "ex2_gops": 4653.0,
                   "source": "nominal (148 SMs x lanes x 1.965 GHz)"}
    if os.path.exists(pp):
        pk["pipes"] = json.load(open(pp))
        pk["pipes"]["source"] = "measured (profiles/pipe_peaks.json, scratch/pipe_peaks.cu)"
    return pk


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU reference
def _cpu_worker(job):
    import cv2
    cv2.setNumThreads(1)
    import torch
    torch.set_num_threads(1)
    from oracle import cv_literal as lit
    kind, a = job
    t0 = time.perf_counter()
    if kind == "mono":
        lit.mono_frame_pair(*a)
    else:
        lit.stereo_frame_pair(*a)
    return time.perf_counter() - t0


def _pool_map(fn, jobs, procs, warm=True):
    if procs <= 1:
        t0 = time.perf_counter()
        out = [fn(j) for j in jobs]
        return out, time.perf_counter() - t0
    import multiprocessing as mp
    ctx = mp.get_context("spawn")  # the parent holds a CUDA context and OpenMP pools: never fork it
    with ctx.Pool(procs) as pool:
        if warm:
            pool.map(fn, jobs[:procs])  # warm the workers (imports, page-in)
        t0 = time.perf_counter()
        out = pool.map(fn, jobs, chunksize=1)
        return out, time.perf_counter() - t0


def cpu_reference(args, wl, n_sample: int, procs: int):
    """The reference's own CPU implementation of the path (oracle/cv_literal.py = the literal cv2 / torch calls of
    mono_slam.py:100-118 / stereo_slam.py:232-306) over `n_sample` frame pairs, data-parallel over `procs` processes
    with one library thread each (findEssentialMat / recoverPose / solvePnPRansac are single-threaded in OpenCV)."""
    fr = wl["frames"]
    n_sample = max(1, min(n_sample, len(wl["pair_a"])))
    jobs = []
    for i in range(n_sample):
        a, b = fr[wl["pair_a"][i]], fr[wl["pair_b"][i]]
        if args.workload == "mono":
            jobs.append(("mono", (a.kpts, a.desc_f32, b.kpts, b.desc_f32)))
        else:
            jobs.append(("stereo", (a.kpts, a.desc_bin, a.disparity, b.kpts, b.desc_bin)))
    _, dt = _pool_map(_cpu_worker, jobs, procs)
    return n_sample / dt, dt, n_sample


# ------------------------------------------------------------------------------------------------ parity leg
def _verify_worker(job):
    """CHECKER (outside every timed region): one frame pair of the timed sequence, the device results against the
    reference's literal calls.  Stages are compared one by one on identical inputs (the device's own gathered
    correspondences go into cv2), and `chain` says whether the whole pair agrees end to end."""
    import cv2
    cv2.setNumThreads(1)
    import torch
    torch.set_num_threads(1)
    from oracle import cv_literal as lit
    kind, fa, fb, g = job
    out = {"match_rows": 0, "ambiguous_rows": 0, "gather_bad": 0, "mask_bytes": 0, "model_exact": 0, "R": 0.0, "t": 0.0,
           "pose_mask_bytes": 0, "points_bad": 0, "chain": 1, "skipped": 0}
    n0 = fa["kpts"].shape[0]
    gm = g["matches0"][:n0].astype(np.int64)
    if kind == "mono":
        lm = lit.superglue_match_tail(fa["desc"], fb["desc"])
        mis = np.nonzero(lm != gm)[0]
        out["match_rows"] = int(mis.size)
        if mis.size:
            # rows whose decision is within 1e-5 of flipping in float64: the reference's own float32
            # contraction is not decisive there either
            o = lit.float_mutual_nn(fa["desc"], fb["desc"], scale=0.0625, threshold=LN_THRESH)
            out["ambiguous_rows"] = int(o["ambiguous0"][mis].sum())
        mk0, mk1 = lit.gather_matches(fa["kpts"], fb["kpts"], gm)
        K = mk0.shape[0]
        out["gather_bad"] = int(K != g["n_matches"] or not np.array_equal(mk0, g["mkpts0"][:K])
                                or not np.array_equal(mk1, g["mkpts1"][:K]))
        E, mask = lit.find_essential(mk0, mk1)
        if E is None or E.shape != (3, 3):
            out["skipped"] = 1
            return out
        out["mask_bytes"] = int((mask.ravel() != g["inlier_mask"][:K]).sum())
        out["model_exact"] = int(np.array_equal(E, g["E"].reshape(3, 3)))
        ng, R, t, pm = lit.recover_pose(E, mk0, mk1)
        out["R"] = lit.angle_error_mat(R, g["R"].reshape(3, 3))
        out["t"] = lit.angle_error_vec(t.ravel(), g["t"].ravel())
        out["pose_mask_bytes"] = int((pm.ravel() != g["pose_mask"][:K]).sum()) + int(ng != g["n_good"])
    else:
        q, tt, _ = lit.bf_hamming_match(fa["desc"], fb["desc"])
        lm = np.full(n0, -1, np.int64)
        lm[q] = tt
        out["match_rows"] = int((lm != gm).sum())
        mk0, mk1 = lit.gather_matches(fa["kpts"], fb["kpts"], gm)
        disp = fa["disparity"][gm > -1]
        P3, valid = lit.backproject(mk0, disp)
        P3, m1 = P3[valid], mk1[valid]
        K = P3.shape[0]
        out["gather_bad"] = int(K != g["n_points"])
        if K == g["n_points"]:
            out["points_bad"] = int(not np.array_equal(P3, g["points3d"][:K]) or not np.array_equal(m1, g["pts2d"][:K]))
        if K < 6:
            out["skipped"] = 1
            return out
        ok, rvec, tvec, inl = lit.solve_pnp_ransac(P3, m1)
        gi = g["inlier_idx"][:g["n_inliers"]]
        li = inl.ravel() if inl is not None else np.zeros(0, np.int32)
        out["mask_bytes"] = int(np.setxor1d(li, gi).size)
        R, _ = cv2.Rodrigues(rvec)
        Rg, _ = cv2.Rodrigues(np.ascontiguousarray(g["rvec"], dtype=np.float64).reshape(3, 1))
        out["R"] = lit.angle_error_mat(R, Rg)
        out["t"] = float(np.linalg.norm(tvec.ravel() - g["tvec"].ravel()) / max(np.linalg.norm(tvec), 1e-12))
        out["model_exact"] = int(out["R"] < 1e-4 and out["t"] < 1e-4)
    if out["match_rows"] or out["gather_bad"] or out["mask_bytes"] or out["pose_mask_bytes"] or out["points_bad"]:
        out["chain"] = 0
    return out


def verify_sequence(args, wl, res, procs: int):
    """All pairs of the timed sequence (this rank's) against oracle/cv_literal.py.  `res` = the device results of the
    LAST timed step (static buffers of the captured pipeline)."""
    import cv2
    fr = wl["frames"]
    P = args.pairs
    host = {k: v.detach().cpu().numpy() for k, v in res.items() if hasattr(v, "detach")}
    jobs = []
    for i in range(P):
        a, b = fr[wl["pair_a"][i]], fr[wl["pair_b"][i]]
        if args.workload == "mono":
            fa = {"kpts": a.kpts, "desc": a.desc_f32}
            fb = {"kpts": b.kpts, "desc": b.desc_f32}
            keys = ("matches0", "n_matches", "mkpts0", "mkpts1", "E", "inlier_mask", "R", "t", "pose_mask", "n_good")
        else:
            fa = {"kpts": a.kpts, "desc": a.desc_bin, "disparity": a.disparity}
            fb = {"kpts": b.kpts, "desc": b.desc_bin}
            keys = ("matches0", "n_matches", "points3d", "pts2d", "n_points", "rvec", "tvec", "inlier_idx", "n_inliers")
        g = {k: (host[k][i] if host[k][i].ndim else host[k][i].item()) for k in keys}
        jobs.append((args.workload, fa, fb, g))
    t0 = time.perf_counter()
    outs, _ = _pool_map(_verify_worker, jobs, procs, warm=False)
    sk = sum(o["skipped"] for o in outs)
    par = {"pairs": P, "against": f"oracle/cv_literal.py (cv2 {cv2.__version__} + torch CPU literal calls)",
           "match_mismatch_rows": sum(o["match_rows"] for o in outs),
           "ambiguous_rows": sum(o["ambiguous_rows"] for o in outs),
           "gather_mismatch_pairs": sum(o["gather_bad"] for o in outs),
           "mask_mismatch_pairs": sum(1 for o in outs if o["mask_bytes"]),
           "mask_mismatch_bytes": sum(o["mask_bytes"] for o in outs),
           "pairs_equal_end_to_end": sum(o["chain"] for o in outs), "pairs_without_model_in_cv2": sk,
           "worst_R_rad": max(o["R"] for o in outs), "worst_t": max(o["t"] for o in outs),
           "seconds": round(time.perf_counter() - t0, 1)}
    if args.workload == "mono":
        par["E_bit_identical_pairs"] = sum(o["model_exact"] for o in outs)
        par["pose_mask_mismatch_pairs"] = sum(1 for o in outs if o["pose_mask_bytes"])
    else:
        par["points3d_mismatch_pairs"] = sum(o["points_bad"] for o in outs)
        par["pose_within_1e-4_pairs"] = sum(o["model_exact"] for o in outs)
    return par


def run_reference(args):
    """`--impl reference`: the reference's CPU path on the box's host cores, same config / metric / unit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.workload not in ("mono", "stereo"):
        args.workload = "mono"
    wl = make_workload(args, 0)
    procs = os.cpu_count() or 1
    for _ in range(max(args.warmup, 0)):
        cpu_reference(args, wl, min(args.pairs, 2 * procs), procs)
    tot_pairs, tot_t = 0, 0.0
    for _ in range(args.steps):
        _, dt, n = cpu_reference(args, wl, args.pairs, procs)
        tot_pairs += n
        tot_t += dt
    value = tot_pairs / tot_t
    import cv2
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64 solvers / f32 residuals", "data": "synthetic (procedural forest, seeded)",
            "config": config_dict(args, wl, 1),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": "reference",
                             "sample": f"{args.pairs} frame pairs per step, {procs} processes x 1 library thread, "
                                       f"cv2 {cv2.__version__} + torch CPU (oracle/cv_literal.py literal calls)"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm
# stage of the step -> the kernel that is (all but a few microseconds of) it; ncu launch lists under profiles/ use
# these names
STAGE_KERNEL = {"match": {"mono": "k_float_top3_tc", "stereo": "k_float_top3_tc (+-1 Hamming operands)"},
                "prepare": {"mono": "k_float_prepare", "stereo": "k_hamming_expand"},
                "gather": "k_gather_matches", "essential_ransac": "k_rg_select/solve/score<EPolicy> rounds",
                "recover_pose": "k_rp_cheirality", "triangulate": "k_backproject_filter",
                "pnp_ransac": "k_rg_select/solve/score<PPolicy> rounds + k_pnp_refine"}


def kernel_of(stage: str, workload: str) -> str:
    k = STAGE_KERNEL.get(stage, stage)
    return k[workload] if isinstance(k, dict) else k


def measured_traffic(kernel: str, args, wl):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of the same
    command (profiles/traffic.json: kernel -> {"pairs": P, "kpts": n, "bytes": B}); None when no capture matches."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(path):
        return None
    try:
        t = json.load(open(path)).get(kernel)
    except Exception:
        return None
    if not t or t.get("pairs") != args.pairs or t.get("kpts") != wl["n_kpts"]:
        return None
    return t.get("bytes")


# fp64 operations (every one its own instruction: the solvers run without FMA contraction) of one minimal solve, counted
# from the code: five-point = 9x9 Jacobi SVD + getCoeffMat (5748 terms) + 10x20 LU + det B(z) + 300 Durand-Kerner sweeps
# of a degree-10 polynomial (10 Horner chains + 90 denominator factors + 10 complex divisions a sweep) + <= 10 models;
# EPnP-on-5 = the 12x12 Jacobi SVD (up to 30 sweeps x 66 rotations x ~150) + betas + Gauss-Newton + Horn
FLOP_PER_E_SOLVE = 20e3 + 17.2e3 + 3e3 + 1.4e3 + 300 * 1.62e3 + 3e3
FLOP_PER_PNP_SOLVE = 160e3
FLOP_PER_SAMPSON = 44.0
FLOP_PER_REPROJ = 40.0


def stage_work(args, wl, res, stage: str):
    """Algorithmic work of one step of `stage` (SURVEY §8d figures): (bound, amount, unit)."""
    P, n = args.pairs, wl["n_kpts"]
    k = float(res["n_matches"].float().mean().item()) if "n_matches" in res else 0.0
    it = float(res["iters"].float().mean().item()) if "iters" in res else 0.0
    if stage == "match" and args.workload == "mono":
        return "tensor", P * 2.0 * n * n * 256, "TFLOP/s"
    if stage == "prepare" and args.workload != "mono":
        return "hbm", (P + 1) * n * (32.0 + 1024.0), "GB/s"  # 32 B of bits in, one 1 KB fp16 operand row out, per descriptor
    if stage == "prepare":
        return "hbm", (P + 1) * n * 256 * (4.0 + 4.0), "GB/s"  # fp32 in, fp16 hi + lo out, once per frame
    if stage == "match":
        return "tensor", P * 2.0 * n * n * 256, "TFLOP/s"  # Hamming as a +-1 fp16 contraction
    if stage == "essential_ransac":
        # the iterations OpenCV's serial loop consumes (speculative solves past the adaptive cap are not counted):
        # one five-point solve and ~6 candidates x K Sampson residuals each
        return "fp64", P * it * (FLOP_PER_E_SOLVE + 6.0 * k * FLOP_PER_SAMPSON), "TFLOP/s"
    if stage == "pnp_ransac":
        return "fp64", P * it * (FLOP_PER_PNP_SOLVE + k * FLOP_PER_REPROJ), "TFLOP/s"
    if stage == "recover_pose":
        return "fp64", P * 4.0 * k * 1.1e3, "TFLOP/s"  # 4 K triangulations = 4x4 Jacobi SVDs (~1.1 kflop each)
    if stage == "triangulate":
        return "hbm", P * k * (12.0 + 8.0 + 20.0), "GB/s"
    return "hbm", P * k * (4.0 + 32.0), "GB/s"  # gather: matches in, two point pairs out


def roofline_of(args, wl, res, stages, dom, dom_s, peaks, timed, kernel_ms=None):
    bound, work, unit = stage_work(args, wl, res, dom)
    if bound == "tensor":
        # a kernel timed alone is held against the burst figure, a whole stage against the sustained one
        key = "bf16_tflops" if kernel_ms is not None else "bf16_tflops_sustained"
        achieved, peak, src = work / dom_s / 1e12, float(peaks.get(key, 1400.0)), peaks["source"]
    elif bound == "fp64":
        achieved, peak = work / dom_s / 1e12, float(peaks["pipes"]["fp64_tflops_no_fma"])
        src = peaks["pipes"]["source"] + "; fp64 add/mul issue rate (the solvers restate OpenCV without FMA contraction)"
    else:
        achieved, peak, src = work / dom_s / 1e9, float(peaks.get("hbm_gbs", 6650.0)), peaks["source"]
    return {"kernel": kernel_of(dom, args.workload), "stage": dom,
            "bound": {"tensor": "tensor", "hbm": "hbm", "fp64": "fp64 pipe"}[bound], "achieved": achieved, "peak": peak,
            "unit": unit, "frac": achieved / peak, "traffic": measured_traffic(kernel_of(dom, args.workload), args, wl),
            "peak_source": src, "timed": timed, "kernel_ms": kernel_ms,
            "stage_ms_per_step": {k: v / args.steps for k, v in stages.items()}}


class Ctx:
    """Process-wide state shared by the workloads of one bench run."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        from forest_slam_b200 import _lib
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise RuntimeError("bench.py needs a CUDA device: forest_slam_b200 has no CPU path")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            # NCCL_DEBUG is honoured as the caller set it.  NCCL writes its log to stdout; the ONE JSON line of this
            # bench is printed last, after a flush, on a line of its own.
            dist.init_process_group("nccl", device_id=self.dev)
        self.lib = _lib.load()
        self.peaks = load_peaks()
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2

    def barrier(self):
        import torch
        import torch.distributed as dist
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, vals):
        import torch
        import torch.distributed as dist
        tt = torch.tensor(vals, dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return [float(x) for x in tt]


def run_sequence(args, ctx: Ctx):
    """mono / stereo: the batched frame-loop drivers (sequence.mono_pairs / stereo_pairs), weak scaling."""
    import torch
    from forest_slam_b200 import _lib, sequence as seq
    lib, dev, rank, world = ctx.lib, ctx.dev, ctx.rank, ctx.world
    wl = make_workload(args, rank)
    P = args.pairs
    pa = torch.from_numpy(wl["pair_a"]).to(dev)
    pb = torch.from_numpy(wl["pair_b"]).to(dev)
    host = {"kpts": torch.from_numpy(wl["kpts"]).pin_memory(), "counts": torch.from_numpy(wl["counts"]).pin_memory()}
    if args.workload == "mono":
        host["desc_f32"] = torch.from_numpy(wl["desc_f32"]).pin_memory()
    else:
        host["desc_bin"] = torch.from_numpy(wl["desc_bin"]).pin_memory()
        host["disparity"] = torch.from_numpy(wl["disparity"]).pin_memory()
    devbuf = {k: torch.empty_like(v, device=dev) for k, v in host.items()}
    h2d_bytes = sum(v.numel() * v.element_size() for v in host.values())

    def upload():
        for k, v in host.items():
            devbuf[k].copy_(v, non_blocking=True)
        return seq.FrameBank(kpts=devbuf["kpts"], counts=devbuf["counts"], desc_f32=devbuf.get("desc_f32"),
                             desc_bin=devbuf.get("desc_bin"), disparity=devbuf.get("disparity"))

    def step(bank, timer=None):
        if args.workload == "mono":
            res = seq.mono_pairs(bank, pa, pb, descriptor="float", scale=0.0625, threshold=LN_THRESH,
                                 exact_fp32=args.exact_fp32, split3=args.split3, timer=timer)
        else:
            res = seq.stereo_pairs(bank, pa, pb, descriptor="binary", timer=timer)
        rec = seq.pack_records(res)
        allrec = seq.gather_records(rec, P * world)
        return res, allrec

    rec_host = torch.empty((P * world, seq.RECORD_F64), dtype=torch.float64).pin_memory()
    flush = ctx.flush
    barrier = ctx.barrier

    bank = upload()
    torch.cuda.synchronize()
    for _ in range(max(args.warmup, 3)):
        step(bank)
    # ---- launches of OUR kernels in one step (library-side counter) and per-stage device times.  The stage pass runs
    # eagerly behind a long spin kernel so that every launch is already queued when the GPU gets to it: the event
    # brackets then measure kernel time, not Python launch latency.
    l0 = lib.fs_launch_count()
    step(bank)
    launches_per_step = lib.fs_launch_count() - l0
    timer = seq.StageTimer()
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda._sleep(int(4e7))
        res, allrec = step(bank, timer)
    torch.cuda.synchronize()
    # ---- the dominant kernel alone (mono, certified tensor-core matcher): the same launch fs_float_match makes
    # (same grid, same operands) through the measurement hook, CUDA events on the launching stream, L2 flushed before
    # every launch
    kernel_ms = None
    if args.workload == "mono" and not args.exact_fp32 and not args.split3:
        from forest_slam_b200 import api
        ops = api.float_prepare(bank.desc_f32, bank.counts)
        n_max = int(bank.desc_f32.shape[2])
        ws = torch.empty(int(lib.fs_float_match_tc_workspace_bytes(P, n_max)), dtype=torch.uint8, device=dev)
        cs = torch.cuda.current_stream(dev)
        spans = []
        for i in range(max(args.warmup, 3) + args.steps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(cs)
            _lib.check(lib.fs_float_top3_keys(api._p(ops), int(ops.shape[0]), n_max, api._p(bank.counts), api._p(pa),
                                              api._p(pb), P, int(os.environ.get("FS_TOP3_PROBE", "0")), None, api._p(ws), ws.numel(),
                                              api._stream()), "fs_float_top3_keys")
            b.record(cs)
            spans.append((a, b))
        torch.cuda.synchronize()
        kernel_ms = sum(a.elapsed_time(b) for a, b in spans[-args.steps:]) / args.steps
        del ops, ws
    # ---- the timed path: the whole step (all launches + the record gather) captured once in a CUDA graph
    kw = (dict(descriptor="float", scale=0.0625, threshold=LN_THRESH, exact_fp32=args.exact_fp32, split3=args.split3)
          if args.workload == "mono" else dict(descriptor="binary"))
    # multi-rank: the record gather stays OUT of the graphs and is issued on one stream in step order (same order on
    # every rank); single rank: gather_records is the identity and stays inside
    kw["gather"] = world == 1
    pipe = seq.CapturedPipeline(bank, pa, pb, kind=args.workload, n_pairs_total=P * world, **kw)
    # further pipelines on their own input buffers: consecutive steps rotate over them, each on its own stream, so that
    # the latency-bound RANSAC rounds of step i run beside the tensor-core matcher of step i+1 (this is how a streaming
    # caller drives the path; the e2e loop below uses the same pipelines with uploads in between)
    n_lanes = max(2, args.lanes)
    pipes, bufs = [pipe], [devbuf]
    for _ in range(n_lanes - 1):
        dbuf = {k: v.clone() for k, v in devbuf.items()}
        bk = seq.FrameBank(kpts=dbuf["kpts"], counts=dbuf["counts"], desc_f32=dbuf.get("desc_f32"),
                           desc_bin=dbuf.get("desc_bin"), disparity=dbuf.get("disparity"))
        pipes.append(seq.CapturedPipeline(bk, pa, pb, kind=args.workload, n_pairs_total=P * world, **kw))
        bufs.append(dbuf)
    main = torch.cuda.current_stream(dev)
    if args.overlap_steps:
        lanes = [torch.cuda.Stream(device=dev) for _ in range(n_lanes)]
        flushes = [flush] + [torch.empty_like(flush) for _ in range(n_lanes - 1)]
    else:
        n_lanes, lanes, flushes = 1, [main], [flush]

    comm = torch.cuda.Stream(device=dev)
    gathered = [torch.cuda.Event() for _ in range(len(pipes))]

    def device_loop(n_steps):
        """n_steps steps with the frame bank resident in HBM; returns (results, records) of the last step."""
        out = None
        for st_ in lanes + [comm]:
            st_.wait_stream(main)
        for i in range(n_steps):
            k = i % n_lanes
            with torch.cuda.stream(lanes[k]):
                lanes[k].wait_event(gathered[k])  # the previous gather of this pipeline's record buffer is done
                flushes[k].zero_()  # L2 flush before every step (a 256 MB write, > 126 MB L2)
                res_, rec_ = pipes[k].run()
            if world > 1:
                comm.wait_stream(lanes[k])
                with torch.cuda.stream(comm):
                    rec_ = seq.gather_records(rec_, P * world)  # the one collective, in step order on every rank
                    gathered[k].record(comm)
            out = (res_, rec_)
        for st_ in lanes + [comm]:
            main.wait_stream(st_)
        return out

    device_loop(max(args.warmup, 3))
    clocks = ClockSampler(ctx.local)
    barrier()
    clocks.start()
    wall0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res, allrec = device_loop(args.steps)
    e1.record()
    barrier()
    wall = time.perf_counter() - wall0
    clk = clocks.stop()
    launches = launches_per_step * args.steps
    t_dev = e0.elapsed_time(e1) / 1e3
    # ---- end-to-end through the public API with host buffers.  Every step copies ITS frame bank from pinned host
    # memory into the pipeline's input tensors, replays the step and reads the gathered records back to the host —
    # all inside the timed region.  Pipelines (own input buffers) rotate so that the upload of step i+1 runs on a copy
    # stream while step i computes; the per-step L2 flush stays on the compute stream.
    n_slots = min(len(pipes), 3)
    rec_hosts = [rec_host] + [torch.empty_like(rec_host).pin_memory() for _ in range(n_slots - 1)]
    copy_stream = torch.cuda.Stream(device=dev)

    def e2e_loop(n_steps):
        copied = [torch.cuda.Event() for _ in range(n_slots)]
        computed = [torch.cuda.Event() for _ in range(n_slots)]

        def upload_to(slot):
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(computed[slot])  # the previous step that read this slot has finished
                for k, v in host.items():
                    bufs[slot][k].copy_(v, non_blocking=True)
                copied[slot].record(copy_stream)

        for s in range(n_slots):
            computed[s].record(main)
        for s in range(min(n_slots - 1, n_steps)):
            upload_to(s)
        for i in range(n_steps):
            cur = i % n_slots
            nxt = i + n_slots - 1
            if nxt < n_steps:
                upload_to(nxt % n_slots)
            main.wait_event(copied[cur])
            flush.zero_()
            _, ar = pipes[cur].run()
            if world > 1:
                ar = seq.gather_records(ar, P * world)
            rec_hosts[cur].copy_(ar, non_blocking=True)
            computed[cur].record(main)

    e2e_loop(2)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_loop(args.steps)
    e1.record()
    barrier()
    t_e2e = e0.elapsed_time(e1) / 1e3
    t_dev, t_e2e = ctx.max_over_ranks([t_dev, t_e2e])
    if rank != 0:
        return None

    total_pairs = P * world * args.steps
    value = total_pairs / t_dev
    stages = timer.totals_ms()
    peaks = ctx.peaks
    dom = max(stages, key=stages.get)
    roof = roofline_of(args, wl, res, stages, dom, stages[dom] / 1e3 / args.steps, peaks,
                       "stage bracket (all launches of the stage)")
    if kernel_ms is not None:
        # the tensor-core matcher timed alone (the north star's named contraction kernel); the stage with the largest
        # bracket is reported next to it
        roof_k = roofline_of(args, wl, res, stages, "match", kernel_ms / 1e3, peaks,
                             "kernel alone (fs_float_top3_keys), events on the launching stream", kernel_ms)
        roof_k["largest_stage"] = {k: roof[k] for k in ("kernel", "stage", "bound", "achieved", "peak", "unit", "frac",
                                                         "peak_source", "timed")}
        roof = roof_k
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "f16 tensor-core scores (f32 accumulate, f32 exact rescoring) + f64 solvers / f32 residuals" if (
                args.workload == "mono" and not args.exact_fp32) else
            "f16 tensor-core Hamming (exact integers) / f32 back-projection + f64 solvers / f32 residuals",
            "data": "synthetic (procedural forest, seeded)", "config": config_dict(args, wl, world),
            "clocks": clk,
            "e2e": {"value": total_pairs / t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d_bytes),
                    "d2h_bytes_per_step": int(rec_host.numel() * 8), "upload_slots": n_slots},
            "gpu_launches": int(launches),
            "roofline": roof,
            "wall_s": wall,
            "stats": {"mean_matches": float(res["n_matches"].float().mean()),
                      "mean_inliers": float(res["n_inliers"].float().mean()),
                      "mean_iters": float(res["iters"].float().mean()),
                      "ok_pairs": int((res["status"] == 0).sum()),
                      **({"matcher_rows_rescanned_exactly": int(res["match_diag"][0]),
                          "matcher_rows_with_rescored_candidates": int(res["match_diag"][1]),
                          "matcher_columns_with_rescored_slabs": int(res["match_diag"][2]),
                          "matcher_columns_rescanned_exactly": int(res["match_diag"][3])}
                         if "match_diag" in res else {})}}
    procs = os.cpu_count() or 1
    if not args.no_cpu_baseline:
        n_s = args.cpu_sample or min(args.pairs, 8 * procs)
        import cv2
        v, dt, n = cpu_reference(args, wl, n_s, procs)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": procs, "kind": "reference",
                                "sample": f"{n} frame pairs of this workload in {dt:.1f} s, {procs} processes x 1 "
                                          f"library thread, cv2 {cv2.__version__} + torch CPU literal calls"}
    if not args.no_verify:
        line["parity"] = verify_sequence(args, wl, res, procs)
    return line


# ------------------------------------------------------------------------------------------------ config 3
def run_float4096(args, ctx: Ctx):
    """BASELINE config 3: float 256-d mutual-NN + ratio test at 4096 kpts/frame, 64 frame pairs per launch, on forest
    descriptors (and, as the worst case for the certification stage, on random unit vectors)."""
    import torch
    from forest_slam_b200 import api, synth
    dev = ctx.dev
    if ctx.rank != 0:
        return None
    P, n = 64, 4096
    world = synth.make_world(synth.SEED0)
    frames = synth.make_sequence(P + 1, n, seed=synth.SEED0 + 77, world=world)
    d = torch.from_numpy(np.stack([f.desc_f32 for f in frames])).to(dev)
    counts = torch.full((P + 1,), n, dtype=torch.int32, device=dev)
    a = torch.arange(P, dtype=torch.int32, device=dev)
    b = a + 1
    out = {}
    flop = P * 2.0 * n * n * 256
    for name, desc in (("forest", d), ("random_unit_vectors", None)):
        if desc is None:
            g = torch.Generator(device="cpu").manual_seed(7)
            desc = torch.randn((P + 1, 256, n), generator=g)
            desc = (desc / desc.norm(dim=1, keepdim=True)).to(dev)
        ops = api.float_prepare(desc, counts)
        fn = lambda: api.float_match_batch(None, counts, a, b, operands=ops, ratio=0.8, want_diag=True)
        for _ in range(3):
            r = fn()
        ts = []
        for _ in range(max(args.steps, 5)):
            ctx.flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = float(np.median(ts))
        out[name] = {"ms_per_launch": ms, "pairs_per_s": P / ms * 1e3, "tflops": flop / ms / 1e9,
                     "frac_of_bf16_burst": flop / ms / 1e9 / ctx.peaks["bf16_tflops"],
                     "mean_matches": float(r["n_matches"].float().mean()),
                     "rows_rescanned_exactly": int(r["diag"][0]) if "diag" in r else None,
                     "columns_with_rescored_slabs": int(r["diag"][2]) if "diag" in r else None,
                     "columns_rescanned_exactly": int(r["diag"][3]) if "diag" in r else None}
        if name == "forest" and not args.no_verify:
            from oracle import cv_literal as lit
            mis = amb = 0
            for i in range(2):   # two pairs against the float64 oracle (18 GFLOP of float64 each on the host)
                o = lit.float_mutual_nn(frames[i].desc_f32, frames[i + 1].desc_f32, ratio=0.8)
                gm = r["matches0"][i].cpu().numpy().astype(np.int64)
                bad = np.nonzero(o["matches0"] != gm)[0]
                mis += int(bad.size)
                amb += int(o["ambiguous0"][bad].sum())
            out[name]["parity"] = {"pairs": 2, "against": "oracle/cv_literal.py float_mutual_nn (float64)",
                                   "match_mismatch_rows": mis, "ambiguous_rows": amb}
        del ops
    f = out["forest"]
    return {"metric": METRIC, "value": f["pairs_per_s"], "unit": UNIT, "n_gpus": 1, "ms_per_step": f["ms_per_launch"],
            "scaling": "weak", "config": {"workload": "float 256-d descriptor mutual-NN + ratio test (0.8), whole "
                                          "fs_float_match call (top-3 pass + certification + finalize)",
                                          "kpts_per_frame": n, "pairs_per_launch": P},
            "roofline": {"kernel": "fs_float_match (single-pass k_float_top3_tc + colmerge + resolve + slabscan + colcheck + finalize)", "bound": "tensor",
                         "achieved": f["tflops"], "peak": ctx.peaks["bf16_tflops"], "unit": "TFLOP/s",
                         "frac": f["frac_of_bf16_burst"], "peak_source": ctx.peaks["source"]},
            "cases": out}


# ------------------------------------------------------------------------------------------------ config 4
def run_pnp_sweep(args, ctx: Ctx):
    """BASELINE config 4: stereo back-projection + PnP RANSAC with H = 1k .. 64k hypotheses per frame pair (adaptive stop
    disabled through the C ABI's `confidence > 1` switch: every one of the H hypotheses is solved and scored)."""
    import torch
    from forest_slam_b200 import api, synth, sequence as seq
    if ctx.rank != 0:
        return None
    dev = ctx.dev
    P, n = 16, 2000
    world = synth.make_world(synth.SEED0)
    frames = synth.make_sequence(P + 1, n, seed=synth.SEED0 + 33, world=world)
    kp, dbin, _, disp, counts = synth.stack_frames(frames)
    bank = seq.FrameBank.from_numpy(kp, counts, desc_bin=dbin, disparity=disp, device=dev)
    a = torch.arange(P, dtype=torch.int32, device=dev)
    b = a + 1
    m = api.hamming_match_batch(bank.desc_bin, bank.counts, a, b, cross_check=True, want_matches1=False)
    g = api.gather_matches_batch(bank.kpts, bank.counts, a, b, m["matches0"])
    cam = api.ForestCamera()
    K = float(g["n_corr"].float().mean())
    sweep = []
    fp64_peak = float(ctx.peaks["pipes"]["fp64_tflops_no_fma"])
    for H in (1000, 2000, 4000, 8000, 16000, 32000, 64000):
        def fn():
            bp = api.backproject_filter_bank(g["mkpts0"], g["mkpts1"], g["n_corr"], bank.disparity, a, g["src_idx"],
                                             camera=cam)
            return api.solve_pnp_ransac_batch(bp["points3d"], bp["pts2d"], bp["n_out"], camera=cam, reproj_error=1.0,
                                              confidence=2.0, max_iters=H, min_points=6), bp
        s, bp = fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(3):
            ctx.flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            s, bp = fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = float(np.median(ts))
        iters = float(s["iters"].float().mean())
        Kp = float(bp["n_out"].float().mean())
        flop = P * iters * (FLOP_PER_PNP_SOLVE + Kp * FLOP_PER_REPROJ)
        # the scoring kernel alone on the same correspondences (one warp per hypothesis)
        rt = torch.cat([s["rvec"], s["tvec"]], dim=1)[:, None, :].expand(P, H, 6).contiguous()
        for _ in range(2):
            api.score_pnp_batch(bp["points3d"], bp["pts2d"], bp["n_out"], rt, camera=cam)
        ctx.flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        api.score_pnp_batch(bp["points3d"], bp["pts2d"], bp["n_out"], rt, camera=cam)
        e1.record()
        torch.cuda.synchronize()
        ms_sc = e0.elapsed_time(e1)
        sweep.append({"H": H, "iterations_run": iters, "ms": ms, "pairs_per_s": P / ms * 1e3,
                      "hypotheses_per_s": P * iters / ms * 1e3, "fp64_tflops": flop / ms / 1e9,
                      "frac_fp64_pipe": flop / ms / 1e9 / fp64_peak,
                      "score_kernel_ms": ms_sc, "score_kernel_tflops": P * H * Kp * FLOP_PER_REPROJ / ms_sc / 1e9,
                      "score_kernel_frac_fp64_pipe": P * H * Kp * FLOP_PER_REPROJ / ms_sc / 1e9 / fp64_peak,
                      "mean_inliers": float(s["n_inliers"].float().mean())})
        del rt
    par = None
    if not args.no_verify:
        # parity at H = 1000 with the reference's own arguments (confidence 0.99: adaptive semantics)
        from oracle import cv_literal as lit
        bp = api.backproject_filter_bank(g["mkpts0"], g["mkpts1"], g["n_corr"], bank.disparity, a, g["src_idx"],
                                         camera=cam)
        s = api.solve_pnp_ransac_batch(bp["points3d"], bp["pts2d"], bp["n_out"], camera=cam, min_points=6)
        bad = 0
        for i in range(P):
            k = int(bp["n_out"][i])
            ok, rv, tv, inl = lit.solve_pnp_ransac(bp["points3d"][i, :k].cpu().numpy(), bp["pts2d"][i, :k].cpu().numpy())
            gi = s["inlier_idx"][i, :int(s["n_inliers"][i])].cpu().numpy()
            bad += int(not np.array_equal(gi, inl.ravel()))
        par = {"pairs": P, "against": "cv2.solvePnPRansac (reference arguments, H = 1000 adaptive)",
               "inlier_set_mismatch_pairs": bad}
    s0 = sweep[0]
    return {"metric": METRIC, "value": s0["pairs_per_s"], "unit": UNIT, "n_gpus": 1, "ms_per_step": s0["ms"],
            "scaling": "weak",
            "config": {"workload": "stereo depth back-projection + PnP RANSAC, H hypotheses per frame pair all solved "
                                   "and scored (value = H 1000)", "pairs_per_launch": P, "kpts_per_frame": n,
                       "mean_correspondences": K},
            "roofline": {"kernel": "k_rg_solve/score<PPolicy> rounds", "bound": "fp64 pipe",
                         "achieved": sweep[-1]["fp64_tflops"], "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": sweep[-1]["frac_fp64_pipe"], "at_H": sweep[-1]["H"],
                         "peak_source": ctx.peaks["pipes"]["source"]},
            "sweep": sweep, "parity": par}


# ------------------------------------------------------------------------------------------------ config 5
def run_loopclosure(args, ctx: Ctx):
    """BASELINE config 5: all C(F, 2) pairs of F = 2000 synthetic keyframes (2000 binary descriptors each) matched
    (Hamming crossCheck on tensor cores), pairs with more than 30 mutual matches verified by essential-matrix RANSAC;
    the pairs are sharded block-cyclically over the ranks (STRONG scaling: the job is fixed), the keyframe bank is
    replicated (one NCCL broadcast), one all_gather of the per-pair records at the end."""
    import torch
    import torch.distributed as dist
    from forest_slam_b200 import synth, sequence as seq
    dev, rank, world = ctx.dev, ctx.rank, ctx.world
    F, n = args.keyframes, 2000
    kp = torch.empty((F, n, 2), dtype=torch.float32, device=dev)
    db = torch.empty((F, n, 32), dtype=torch.uint8, device=dev)
    counts = torch.full((F,), n, dtype=torch.int32, device=dev)
    frames = None
    t0 = time.perf_counter()
    if rank == 0:
        wld = synth.make_world(synth.SEED0)
        frames = synth.make_loop_sequence(F, n, world=wld)
        kp.copy_(torch.from_numpy(np.stack([f.kpts for f in frames])))
        db.copy_(torch.from_numpy(np.stack([f.desc_bin for f in frames])))
    t_gen = time.perf_counter() - t0
    if world > 1:
        dist.broadcast(kp, 0)   # the replicated keyframe bank (160 MB): once per job, outside the timed region
        dist.broadcast(db, 0)
    bank = seq.FrameBank(kpts=kp, counts=counts, desc_bin=db)
    total = F * (F - 1) // 2
    # warm-up on a small bank of the same shape (sizes the workspaces)
    small = seq.FrameBank(kpts=kp[:96], counts=counts[:96], desc_bin=db[:96])
    seq.gather_loop_records(seq.loop_closure_pairs(small, rank=rank, world=world), 96)
    ctx.barrier()
    clocks = ClockSampler(ctx.local)
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    rec_local = seq.loop_closure_pairs(bank, descriptor="binary", min_matches=30, rank=rank, world=world)
    rec = seq.gather_loop_records(rec_local, F)
    e1.record()
    ctx.barrier()
    clk = clocks.stop()
    t, = ctx.max_over_ranks([e0.elapsed_time(e1) / 1e3])
    if rank != 0:
        return None
    r = rec.cpu().numpy()
    ver = r[:, 4] >= 0
    loops = ver & (np.abs(r[:, 2] - r[:, 1]) > 200) & (r[:, 4] > 50)
    flop = total * 2.0 * n * n * 256
    out = {"metric": METRIC, "value": total / t, "unit": UNIT, "n_gpus": world, "ms_per_step": 1e3 * t, "steps": 1,
           "scaling": "strong", "clocks": clk,
           "config": {"workload": "all-pairs loop-closure matching (Hamming crossCheck) + essential-matrix RANSAC "
                                  "verification of pairs with > 30 matches (a mutual match counts when its Hamming distance is <= 64)", "keyframes": F, "kpts_per_frame": n,
                      "pairs": total, "trajectory": "two laps of a 25 m circle in the 80 m forest (synth.make_loop_sequence)",
                      "sharding": f"block-cyclic (1024 pairs) over {world} rank(s); bank replicated by one broadcast; "
                                  "one all_gather of the records"},
           "stats": {"pairs_verified": int(ver.sum()), "loop_candidates_gt50_inliers_gt200_apart": int(loops.sum()),
                     "keyframe_generation_s_host": round(t_gen, 1)},
           "roofline": {"kernel": "k_float_top3_tc (+-1 Hamming operands)", "bound": "tensor",
                        "achieved": flop / t / 1e12 / world, "peak": ctx.peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
                        "frac": flop / t / 1e12 / world / ctx.peaks["bf16_tflops_sustained"],
                        "peak_source": ctx.peaks["source"], "timed": "whole job (matching + verification + gather) per GPU"}}
    if not args.no_verify:
        import cv2
        from oracle import cv_literal as lit
        rng = np.random.default_rng(1)
        vi = np.nonzero(ver)[0]
        idx = np.concatenate([rng.choice(vi, min(24, vi.size), replace=False), rng.choice(total, 8, replace=False)])
        bad_m = bad_i = 0
        for j in idx:
            _, a, b, nm, ni = (int(x) for x in r[j])
            q, tt, d = lit.bf_hamming_match(frames[a].desc_bin, frames[b].desc_bin)
            q, tt = q[d <= 64], tt[d <= 64]      # the job's distance gate (sequence.loop_closure_pairs max_distance)
            bad_m += int(q.size != nm)
            if nm > 30:
                E, mask = cv2.findEssentialMat(frames[a].kpts[q], frames[b].kpts[tt], focal=lit.K0[0, 0],
                                               pp=(lit.K0[0, 2], lit.K0[1, 2]), method=cv2.RANSAC, prob=0.999, threshold=1.0)
                bad_i += int((int(mask.sum()) if mask is not None else -1) != ni)
            else:
                bad_i += int(ni != -1)
        out["parity"] = {"pairs": int(idx.size), "against": f"cv2 {cv2.__version__} BFMatcher + findEssentialMat",
                         "match_count_mismatch_pairs": bad_m, "inlier_count_mismatch_pairs": bad_i}
    return out


# ------------------------------------------------------------------------------------------------ main
def run_b200(args):
    ctx = Ctx()
    t_start = time.perf_counter()
    if args.workload in ("mono", "stereo"):
        line = run_sequence(args, ctx)
        if args.workload == "mono" and not args.headline_only:
            extras = {}
            sa = copy.copy(args)
            sa.workload, sa.kpts = "stereo", 0
            r = run_sequence(sa, ctx)
            if r is not None:
                extras["stereo"] = {k: r[k] for k in ("value", "unit", "ms_per_step", "scaling", "config", "e2e",
                                                      "gpu_launches", "roofline", "stats", "cpu_baseline", "parity")
                                    if k in r}
            if ctx.world == 1:
                extras["float4096"] = run_float4096(args, ctx)
                extras["pnp_sweep"] = run_pnp_sweep(args, ctx)
            lc = run_loopclosure(args, ctx)
            if lc is not None:
                extras["loopclosure"] = lc
            if line is not None:
                line["workloads"] = extras
    elif args.workload == "float4096":
        line = run_float4096(args, ctx)
    elif args.workload == "pnp_sweep":
        line = run_pnp_sweep(args, ctx)
    else:
        line = run_loopclosure(args, ctx)
    if ctx.rank == 0 and line is not None:
        line.setdefault("n_gpus", ctx.world)
        line.setdefault("steps", args.steps)
        line.setdefault("warmup", max(args.warmup, 3))
        line.setdefault("higher_is_better", True)
        line.setdefault("vs_baseline", None)
        line.setdefault("data", "synthetic (procedural forest, seeded)")
        line["bench_wall_s"] = round(time.perf_counter() - t_start, 1)
        sys.stdout.flush()
        print("\n" + json.dumps(line), flush=True)
    _finish(ctx.world)


def _finish(world: int):
    """Leave without tearing the NCCL communicator down: destroy_process_group() can block forever when collectives
    of that communicator live inside captured CUDA graphs (a bench-only shortcut).  All ranks meet at a
    barrier first, so nobody is killed mid-collective; stdout is already flushed."""
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
