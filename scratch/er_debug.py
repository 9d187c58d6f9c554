"""scratch: run the mono step once on bench data and dump the RANSAC engine's control block + per-pair iterations"""
import sys, os, argparse
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from forest_slam_b200 import api, sequence as seq, _lib
args = argparse.Namespace(pairs=int(sys.argv[1]) if len(sys.argv) > 1 else 128, kpts=0, workload="mono", pos_outlier=-1.0)
wl = bench.make_workload(args, 0)
dev = torch.device("cuda")
bank = seq.FrameBank.from_numpy(wl["kpts"], wl["counts"], desc_f32=wl["desc_f32"])
P = args.pairs
pa = torch.arange(0, P, dtype=torch.int32, device=dev); pb = pa + 1
res = seq.mono_pairs(bank, pa, pb, descriptor="float", scale=0.0625, threshold=bench.LN_THRESH)
torch.cuda.synchronize()
ws = api._workspaces[(dev.index if dev.index is not None else torch.cuda.current_device(), "ransac")]
n_max = res["mkpts0"].shape[1]
a256 = lambda x: (x + 255) // 256 * 256
off_st = a256(P * 4 * n_max * 8)
off_ctl = a256(off_st + P * 120)
ctl = ws[off_ctl:off_ctl + 41 * 16].cpu().numpy().view(np.int32).reshape(41, 4)
print("round: n_used n_active n_tail")
for r in range(8): print(r, ctl[r][:3])
it = res["iters"].cpu().numpy(); print("iters: mean %.1f max %d; hist >32: %d, >64: %d, >128: %d, >288: %d" % (it.mean(), it.max(), (it > 32).sum(), (it > 64).sum(), (it > 128).sum(), (it > 288).sum()))
print("n_corr mean", res["n_matches"].float().mean().item(), "inliers mean", res["n_inliers"].float().mean().item())
