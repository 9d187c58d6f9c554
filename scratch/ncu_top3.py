"""k_float_top3_tc alone (the bench's roofline workload: 256 pairs x 2048^2, forest-like descriptors): timings with the
probe bits, then left to ncu (`-k regex:k_float_top3`)."""
import sys, torch
sys.path.insert(0, ".")
from forest_slam_b200 import api, _lib
lib = _lib.load()
P, n, F = 256, 2048, 257
dev = torch.device("cuda")
g = torch.Generator(device="cpu").manual_seed(5)
desc = torch.randn(F, 256, n, generator=g).to(dev)
desc /= desc.norm(dim=1, keepdim=True)
counts = torch.full((F,), n, dtype=torch.int32, device=dev)
ops = api.float_prepare(desc, counts)
pa = torch.arange(P, dtype=torch.int32, device=dev); pb = pa + 1
ws = torch.empty(int(lib.fs_float_match_tc_workspace_bytes(P, n)), dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
probes = [int(x) for x in sys.argv[1:]] or [0]
for probe in probes:
    ts = []
    for rep in range(4):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.fs_float_top3_keys(api._p(ops), F, n, api._p(counts), api._p(pa), api._p(pb), P, probe, None,
                                          api._p(ws), ws.numel(), api._stream()), "x")
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    print(f"probe {probe}: {min(ts[1:]):.4f} ms", flush=True)
