#!/usr/bin/env python
"""bench.py — frame-pairs/s of the Forest-SLAM matching + pose hot path on N B200s, next to the host-CPU reference.

Contract (driver): `python bench.py --gpus N --steps K --warmup W [--impl reference]` prints ONE JSON line on rank 0.
A step = one pass of the hot path over one batch of synthetic frame pairs (BASELINE.json configs[1]: mono_slam.py
temporal matching + essential-matrix RANSAC + recoverPose at 2048 keypoints/frame; `--workload stereo` runs the
stereo_slam.py path: Hamming match + depth back-projection + PnP RANSAC at 2000 keypoints/frame).
value   = pairs/s with the frame bank resident in HBM, CUDA events, max over ranks;
e2e     = the same through the public batched API with HOST (pinned) buffers: H2D of the frame bank and D2H of the
          per-pair result records inside the timed region;
roofline= the dominant kernel's algorithmic work / its CUDA-event time inside the timed steps vs MEASURED_PEAKS.json;
cpu_baseline = the reference's own cv2/torch-CPU calls (oracle/cv_literal.py) timed on a bounded sample on rank 0.
"""
from __future__ import annotations

import argparse
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="mono", choices=["mono", "stereo"])
    ap.add_argument("--pairs", type=int, default=256, help="frame pairs per step per GPU")
    ap.add_argument("--kpts", type=int, default=0, help="keypoints per frame (default 2048 mono / 2000 stereo)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="pairs timed for cpu_baseline (0 = auto)")
    ap.add_argument("--exact-fp32", action="store_true", help="CUDA-core fp32 matcher instead of tcgen05")
    ap.add_argument("--split3", action="store_true", help="split-fp16 three-MMA matcher instead of the certified pass")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--lanes", type=int, default=5,
                    help="captured pipelines (own buffers, own stream) the device loop rotates through when --overlap-steps")
    ap.add_argument("--overlap-steps", type=int, default=1,
                    help="1: consecutive steps alternate between two streams (software pipelining across steps)")
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
    frames = synth.make_sequence(args.pairs + 1, n_kpts, seed=synth.SEED0 + 1000 * rank, world=world,
                                 pos_outlier=pos_outlier(args))
    kpts, dbin, df32, disp, counts = synth.stack_frames(frames)
    return {"frames": frames, "kpts": kpts, "desc_bin": dbin, "desc_f32": df32, "disparity": disp, "counts": counts,
            "n_kpts": n_kpts}


def pos_outlier(args) -> float:
    if args.pos_outlier >= 0:
        return args.pos_outlier
    return 0.1 if args.workload == "mono" else 0.0


def config_dict(args, wl, world_size):
    name = ("mono_slam.py temporal matching (float 256-d mutual-NN) + essential-matrix RANSAC + recoverPose"
            if args.workload == "mono" else
            "stereo_slam.py temporal matching (256-bit Hamming crossCheck) + depth back-projection + PnP RANSAC")
    return {"workload": name, "kpts_per_frame": int(wl["n_kpts"]), "pairs_per_step_per_gpu": int(args.pairs),
            "image": "960x600", "gross_outlier_prob_per_observation": pos_outlier(args), "sharding": f"frame-pair index over {world_size} rank(s), one all_gather of records",
            "l2": "L2 flushed (256 MB write) between timed iterations; frame bank per step %.0f MB" % (
                bank_bytes(args, wl) / 1e6),
            "steps_in_flight": (max(2, args.lanes) if args.overlap_steps else 1)}


def bank_bytes(args, wl):
    F = args.pairs + 1
    n = wl["n_kpts"]
    per = n * 8 + (n * 256 * 4 if args.workload == "mono" else n * 32 + n * 4)
    return F * per


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


def cpu_reference(args, wl, n_sample: int, procs: int):
    """The reference's own CPU implementation of the path (oracle/cv_literal.py = the literal cv2 / torch calls of
    mono_slam.py:100-118 / stereo_slam.py:232-306) over `n_sample` frame pairs, data-parallel over `procs` processes
    with one library thread each (findEssentialMat / recoverPose / solvePnPRansac are single-threaded in OpenCV)."""
    fr = wl["frames"]
    n_sample = max(1, min(n_sample, len(fr) - 1))
    jobs = []
    for i in range(n_sample):
        a, b = fr[i], fr[i + 1]
        if args.workload == "mono":
            jobs.append(("mono", (a.kpts, a.desc_f32, b.kpts, b.desc_f32)))
        else:
            jobs.append(("stereo", (a.kpts, a.desc_bin, a.disparity, b.kpts, b.desc_bin)))
    if procs <= 1:
        t0 = time.perf_counter()
        for j in jobs:
            _cpu_worker(j)
        dt = time.perf_counter() - t0
    else:
        import multiprocessing as mp
        ctx = mp.get_context("spawn")  # the parent holds a CUDA context and OpenMP pools: never fork it
        with ctx.Pool(procs) as pool:
            pool.map(_cpu_worker, jobs[:procs])  # warm the workers (imports, page-in)
            t0 = time.perf_counter()
            pool.map(_cpu_worker, jobs, chunksize=1)
            dt = time.perf_counter() - t0
    return n_sample / dt, dt, n_sample


def run_reference(args):
    """`--impl reference`: the reference's CPU path on the box's host cores, same config / metric / unit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
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
                "gather": "k_gather_matches", "essential_ransac": "k_rg_select/solve/tail/score<EPolicy> rounds",
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
        return "hbm", P * ((n + n) * 32.0 + n * 12.0), "GB/s"
    if stage == "essential_ransac":
        return "hbm", P * (k * 16.0 + it * (20.0 + 72.0 + 4.0) + k + 72.0), "GB/s"
    if stage == "pnp_ransac":
        return "hbm", P * (k * 20.0 + it * (20.0 + 48.0 + 4.0) + k * 5.0 + 48.0), "GB/s"
    if stage == "recover_pose":
        return "hbm", P * (k * 16.0 + k + 72.0 + 96.0), "GB/s"
    if stage == "triangulate":
        return "hbm", P * k * (12.0 + 8.0 + 20.0), "GB/s"
    return "hbm", P * k * (4.0 + 32.0), "GB/s"  # gather: matches in, two point pairs out


def run_b200(args):
    import torch
    import torch.distributed as dist
    from forest_slam_b200 import _lib, sequence as seq

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: forest_slam_b200 has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL prints its version banner on STDOUT when NCCL_DEBUG=VERSION/INFO: the bench's stdout is ONE JSON line
        os.environ["NCCL_DEBUG"] = os.environ.get("FS_NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()

    wl = make_workload(args, rank)
    P = args.pairs
    pa = torch.arange(0, P, dtype=torch.int32, device=dev)
    pb = pa + 1
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
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
    # second pipeline on its own input buffers: consecutive steps alternate between the two, each on its own stream,
    # so that the latency-bound RANSAC rounds of step i run beside the tensor-core matcher of step i+1 (this is how a
    # streaming caller drives the path; the e2e loop below uses the same pair of pipelines with uploads in between)
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
    clocks = ClockSampler(local)
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
    # all inside the timed region.  Two pipelines (A/B input buffers) alternate so that the upload of step i+1 runs on
    # a copy stream while step i computes; the per-step L2 flush stays on the compute stream.
    rec_hosts = [rec_host, torch.empty_like(rec_host).pin_memory()]
    copy_stream = torch.cuda.Stream(device=dev)

    def e2e_loop(n_steps):
        copied = [torch.cuda.Event() for _ in range(2)]
        computed = [torch.cuda.Event() for _ in range(2)]

        def upload_to(slot):
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(computed[slot])  # the previous step that read this slot has finished
                for k, v in host.items():
                    bufs[slot][k].copy_(v, non_blocking=True)
                copied[slot].record(copy_stream)

        computed[0].record(main)
        computed[1].record(main)
        upload_to(0)
        for i in range(n_steps):
            cur = i & 1
            if i + 1 < n_steps:
                upload_to(cur ^ 1)
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
    tt = torch.tensor([t_dev, t_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_dev, t_e2e = float(tt[0]), float(tt[1])
    if rank != 0:
        _finish(world)
        return

    total_pairs = P * world * args.steps
    value = total_pairs / t_dev
    stages = timer.totals_ms()
    dom = max(stages, key=stages.get)
    bound, work, unit = stage_work(args, wl, res, dom)
    peaks = {}
    pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak_src = "measured (MEASURED_PEAKS.json)"
    if os.path.exists(pk_path):
        peaks = json.load(open(pk_path))
    else:
        peaks = {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0}
        peak_src = "fallback (B200_PROFILING.md)"
    dom_s = stages[dom] / 1e3 / args.steps
    timed = "stage bracket (all launches of the stage)"
    if kernel_ms is not None:
        # the single kernel with the largest share of the step (profiles/: ncu launch list); multi-launch stages
        # (RANSAC rounds) are dozens of short latency-bound fp64 launches with no HBM / tensor roofline of their own
        dom, dom_s, timed = "match", kernel_ms / 1e3, "kernel alone (fs_float_top3_keys), events on the launching stream"
        bound, work, unit = stage_work(args, wl, res, dom)
    if bound == "tensor":
        # a kernel timed alone is held against the burst figure, a whole stage against the sustained one
        key = "bf16_tflops" if kernel_ms is not None else "bf16_tflops_sustained"
        achieved, peak = work / dom_s / 1e12, float(peaks.get(key, 1400.0))
    else:
        achieved, peak = work / dom_s / 1e9, float(peaks.get("hbm_gbs", 6650.0))
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "f16 tensor-core scores (f32 accumulate, f32 exact rescoring) + f64 solvers / f32 residuals" if (
                args.workload == "mono" and not args.exact_fp32) else "u8/f32 + f64 solvers / f32 residuals",
            "data": "synthetic (procedural forest, seeded)", "config": config_dict(args, wl, world),
            "clocks": clk,
            "e2e": {"value": total_pairs / t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d_bytes),
                    "d2h_bytes_per_step": int(rec_host.numel() * 8)},
            "gpu_launches": int(launches),
            "roofline": {"kernel": kernel_of(dom, args.workload), "stage": dom, "bound": bound, "achieved": achieved,
                         "peak": peak, "unit": unit, "frac": achieved / peak,
                         "traffic": measured_traffic(kernel_of(dom, args.workload), args, wl),
                         "peak_source": peak_src, "timed": timed, "kernel_ms": kernel_ms,
                         "stage_ms_per_step": {k: v / args.steps for k, v in stages.items()}},
            "wall_s": wall,
            "stats": {"mean_matches": float(res["n_matches"].float().mean()),
                      "mean_inliers": float(res["n_inliers"].float().mean()),
                      "mean_iters": float(res["iters"].float().mean()),
                      "ok_pairs": int((res["status"] == 0).sum()),
                      **({"matcher_rows_rescanned_exactly": int(res["match_diag"][0]),
                          "matcher_rows_with_rescored_candidates": int(res["match_diag"][1])}
                         if "match_diag" in res else {})}}
    if world == 1 and not args.no_cpu_baseline:
        procs = os.cpu_count() or 1
        n_s = args.cpu_sample or min(args.pairs, 8 * procs)
        import cv2
        v, dt, n = cpu_reference(args, wl, n_s, procs)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": procs, "kind": "reference",
                                "sample": f"{n} frame pairs of this workload in {dt:.1f} s, {procs} processes x 1 "
                                          f"library thread, cv2 {cv2.__version__} + torch CPU literal calls"}
    print(json.dumps(line), flush=True)
    _finish(world)


def _finish(world: int):
    """Leave without tearing the NCCL communicator down: destroy_process_group() can block forever when collectives
    of that communicator live inside captured CUDA graphs.  All ranks meet at a barrier first, so nobody is killed
    mid-collective; stdout is already flushed."""
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
