"""Pipe-overlap experiment for k_float_top3_tc: time the real kernel, MMAs only, epilogue only (fs_float_top3_keys probe)."""
import sys, torch
sys.path.insert(0, ".")
from forest_slam_b200 import api, _lib
lib = _lib.load()
P, n, F = 256, 2048, 257
dev = torch.device("cuda")
desc = torch.randn(F, 256, n, device=dev)
desc /= desc.norm(dim=1, keepdim=True)
counts = torch.full((F,), n, dtype=torch.int32, device=dev)
ops = api.float_prepare(desc, counts)
pa = torch.arange(P, dtype=torch.int32, device=dev); pb = pa + 1
ws = torch.empty(int(lib.fs_float_match_tc_workspace_bytes(P, n)), dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
trace = torch.zeros(160, dtype=torch.int64, device=dev)
for probe in (0, 3):
    for pair in (0, 200):
        for rep in range(3):
            flush.zero_()
            trace.zero_()
            _lib.check(lib.fs_float_top3_keys(api._p(ops), F, n, api._p(counts), api._p(pa), api._p(pb), P, probe | (pair << 8),
                                              api._p(trace), api._p(ws), ws.numel(), api._stream()), "x")
            torch.cuda.synchronize()
        t = trace.cpu().numpy()
        t0 = t[0]
        rel = lambda i: int(t[i] - t0) if t[i] else -1
        print(f"probe {probe} pair {pair}: sync {rel(1)} a_full {rel(3)} end {rel(4)}")
        print("  tma  issued :", [rel(10 + i) for i in range(16)])
        print("  mma  start  :", [rel(30 + i) for i in range(16)])
        print("  mma  issued :", [rel(50 + i) for i in range(16)])
        print("  epi  full   :", [rel(70 + i) for i in range(16)])
        print("  epi  loaded :", [rel(110 + i) for i in range(16)])
        print("  epi  done   :", [rel(90 + i) for i in range(16)], flush=True)
