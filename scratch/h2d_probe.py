#!/usr/bin/env python
"""Copy-only probe: what the host side of the box delivers to N GPUs at once.
Every rank copies one pinned 543 MB buffer (the mono step's frame bank) to its GPU `reps` times, all ranks together
(barrier first); prints per-rank GB/s, the aggregate and — with --chunks — the same bytes as C chunked cudaMemcpyAsync.
  python -m torch.distributed.run --nproc-per-node N scratch/h2d_probe.py      (or plain python for N = 1)"""
import argparse, json, os, time
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--mb", type=int, default=543)
ap.add_argument("--reps", type=int, default=12)
ap.add_argument("--chunks", type=int, default=1)
args = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = args.mb << 20
host = torch.empty(n, dtype=torch.uint8).pin_memory()
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
res = {}
for chunks in sorted({1, args.chunks}):
    parts = [(i * n // chunks, (i + 1) * n // chunks) for i in range(chunks)]
    for _ in range(2):
        dev.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.reps):
        for a, b in parts:
            dev[a:b].copy_(host[a:b], non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    gbs = args.reps * n / (e0.elapsed_time(e1) / 1e3) / 1e9
    t = torch.tensor([gbs], dtype=torch.float64, device="cuda")
    if world > 1:
        allv = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allv, t)
        vals = [float(v) for v in allv]
    else:
        vals = [gbs]
    res[f"chunks_{chunks}"] = {"per_rank_gbs": [round(v, 2) for v in vals], "aggregate_gbs": round(sum(vals), 2)}
if rank == 0:
    print(json.dumps({"n_gpus": world, "mb_per_copy": args.mb, "reps": args.reps, **res}), flush=True)
if world > 1:
    dist.barrier()
    os._exit(0)
