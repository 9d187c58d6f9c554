"""Smallest matcher cases for compute-sanitizer --tool memcheck: ragged float and Hamming banks through the single-pass
kernels (out-of-bounds slab / key / staging writes would show here)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from forest_slam_b200 import api as fsb
rng = np.random.default_rng(3)
n = 384
counts = [n, n - 37, 1, 0, 129]
F = len(counts)
desc = rng.standard_normal((F, 256, n)).astype(np.float32)
desc /= np.linalg.norm(desc, axis=1, keepdims=True)
desc[1, :, :100] = desc[0, :, 50:150]
for f, c in enumerate(counts):
    desc[f, :, c:] = 0
pa, pb = [0, 1, 1, 4, 2, 3, 0], [1, 0, 4, 1, 0, 1, 3]
cnt = torch.tensor(counts, dtype=torch.int32).cuda()
r = fsb.float_match_batch(torch.from_numpy(desc).cuda(), cnt, pa, pb, ratio=0.9, want_diag=True)
print("float", r["n_matches"].tolist(), r["diag"].tolist())
n2 = 200  # not a multiple of the 128-column tile
bank = rng.integers(0, 256, (F, n2, 32), dtype=np.uint8)
cnt2 = torch.tensor([n2, n2 - 7, 1, 0, 129], dtype=torch.int32).cuda()
dev = torch.from_numpy(bank).cuda()
ops = fsb.hamming_prepare(dev, cnt2)
h = fsb.hamming_match_batch(None, cnt2, pa, pb, cross_check=True, operands=ops)
ref = fsb.hamming_match_batch(dev, cnt2, pa, pb, cross_check=True, popc=True)
print("hamming", h["n_matches"].tolist(), all(torch.equal(h[k], ref[k]) for k in ("matches0", "matches1", "dist0")))
torch.cuda.synchronize()
