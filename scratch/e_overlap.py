"""Does the E-RANSAC stage overlap with itself?  512 two-view pairs (K ~ 1600, 27 % outliers) through
fs_find_essential_ransac on 1 stream back to back vs the same calls spread over 2 / 3 / 5 streams (own workspaces)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from forest_slam_b200 import api, _lib
sys.path.insert(0, "tests")
import util
lib = _lib.load()
P, n_max = 512, 2048
rng = np.random.default_rng(0)
mk0 = np.zeros((P, n_max, 2), np.float32); mk1 = np.zeros_like(mk0); nc = np.zeros(P, np.int32)
for p in range(P):
    k = int(rng.integers(1500, 1750))
    a, b, _, _, _ = util.two_view(k, 1000 + p, outliers=0.27)
    mk0[p, :k], mk1[p, :k], nc[p] = a, b, k
dev = torch.device("cuda")
t0, t1, tn = torch.from_numpy(mk0).to(dev), torch.from_numpy(mk1).to(dev), torch.from_numpy(nc).to(dev)
need = int(lib.fs_find_essential_workspace_bytes(P, n_max, 1000))
def make():
    return dict(E=torch.empty((P, 9), dtype=torch.float64, device=dev), mask=torch.empty((P, n_max), dtype=torch.uint8, device=dev),
                ni=torch.empty(P, dtype=torch.int32, device=dev), it=torch.empty(P, dtype=torch.int32, device=dev),
                st=torch.empty(P, dtype=torch.int32, device=dev), ws=torch.empty(need, dtype=torch.uint8, device=dev))
def call(b, stream):
    _lib.check(lib.fs_find_essential_ransac(t0.data_ptr(), t1.data_ptr(), tn.data_ptr(), P, n_max, util.FOCAL, util.PP[0], util.PP[1],
                                            0.999, 1.0, 1000, b["E"].data_ptr(), b["mask"].data_ptr(), b["ni"].data_ptr(),
                                            b["it"].data_ptr(), b["st"].data_ptr(), b["ws"].data_ptr(), need, stream.cuda_stream), "E")
for ns in (1, 2, 3, 5):
    streams = [torch.cuda.Stream() for _ in range(ns)]
    bufs = [make() for _ in range(ns)]
    for rep in range(2):
        torch.cuda.synchronize()
        t = time.perf_counter()
        N = 20
        for i in range(N):
            call(bufs[i % ns], streams[i % ns])
        torch.cuda.synchronize()
        dt = time.perf_counter() - t
    print(f"{ns} stream(s): {dt / N * 1e3:.3f} ms per 512-pair call, mean iters {bufs[0]['it'].float().mean().item():.1f}", flush=True)
