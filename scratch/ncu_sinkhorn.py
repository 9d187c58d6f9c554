"""ncu target: one cooperative launch of the resident Sinkhorn kernel (6 pairs of 1024 x 1024, 20 iterations)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from forest_slam_b200 import api as fsb
P, m, n = 6, 1024, 1024
g = torch.Generator().manual_seed(1)
sc = (torch.randn((P, m, n), generator=g) * 2.0).cuda()
cm = torch.full((P,), m, dtype=torch.int32)
cn = torch.full((P,), n, dtype=torch.int32)
for _ in range(2):
    r = fsb.sinkhorn_match_batch(sc, cm, cn)
torch.cuda.synchronize()
print("matched", int((r["matches0"] >= 0).sum()))
