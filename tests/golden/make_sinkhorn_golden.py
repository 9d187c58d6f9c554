"""Generates tests/golden/sinkhorn_golden.npz by running the REFERENCE's own code: log_optimal_transport and the
match tail of SuperGlue.forward (SuperGluePretrainedNetwork/models/superglue.py:143-172, :268-285), imported from
/root/reference (pure torch; no weights are needed for these functions) on seeded score matrices, with the
hyper-parameters of mono_slam.py:59-63 (20 iterations, threshold 0.2).  The reference tree does not travel to the GPU
box, so the outputs are committed.  Run in the build container: python tests/golden/make_sinkhorn_golden.py"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402

REF = "/root/reference/ros_ws/src/SuperGluePretrainedNetwork/models/superglue.py"

CASES = [  # (m, n, seed, gain, bin_score, iters, threshold)
    (37, 52, 500, 20.0, 1.0, 20, 0.2),
    (128, 96, 501, 20.0, 1.0, 20, 0.2),
    (1, 5, 502, 20.0, 1.0, 20, 0.2),
    (64, 64, 503, 8.0, 2.37, 20, 0.2),
    (200, 333, 504, 30.0, 1.0, 100, 0.2),
    (90, 77, 505, 20.0, -0.5, 3, 0.0),
]


def reference_head(sg, scores, alpha, iters, thr):
    """superglue.py:263-285 with the reference's own functions (forward() itself needs the GNN weights)."""
    s = torch.from_numpy(scores)[None]
    Z = sg.log_optimal_transport(s, torch.tensor(float(alpha)), iters=iters)
    max0, max1 = Z[:, :-1, :-1].max(2), Z[:, :-1, :-1].max(1)
    indices0, indices1 = max0.indices, max1.indices
    mutual0 = sg.arange_like(indices0, 1)[None] == indices1.gather(1, indices0)
    mutual1 = sg.arange_like(indices1, 1)[None] == indices0.gather(1, indices1)
    zero = Z.new_tensor(0)
    mscores0 = torch.where(mutual0, max0.values.exp(), zero)
    mscores1 = torch.where(mutual1, mscores0.gather(1, indices1), zero)
    valid0 = mutual0 & (mscores0 > thr)
    valid1 = mutual1 & valid0.gather(1, indices1)
    indices0 = torch.where(valid0, indices0, indices0.new_tensor(-1))
    indices1 = torch.where(valid1, indices1, indices1.new_tensor(-1))
    return Z[0].numpy(), indices0[0].numpy(), indices1[0].numpy(), mscores0[0].numpy(), mscores1[0].numpy()


def main():
    spec = importlib.util.spec_from_file_location("ref_superglue", REF)
    sg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sg)
    torch.set_num_threads(1)
    out = {"n_cases": np.int32(len(CASES)), "torch_version": np.array(torch.__version__)}
    for i, (m, n, seed, gain, alpha, iters, thr) in enumerate(CASES):
        sc = util.sinkhorn_scores(m, n, seed, gain)
        if i == 3:  # exact duplicates: ties must fall on the lowest index
            sc[5] = sc[4]
            sc[:, 9] = sc[:, 8]
        Z, i0, i1, s0, s1 = reference_head(sg, sc, alpha, iters, thr)
        out[f"c{i}_params"] = np.array([m, n, seed, gain, alpha, iters, thr], np.float64)
        out[f"c{i}_scores"] = sc
        if m * n <= 16384:  # the full transported matrix only for the small cases (fixture size)
            out[f"c{i}_Z"] = Z
        out[f"c{i}_matches0"], out[f"c{i}_matches1"] = i0, i1
        out[f"c{i}_mscores0"], out[f"c{i}_mscores1"] = s0, s1
        print(i, (m, n), "matched", int((i0 >= 0).sum()))
    np.savez_compressed(os.path.join(HERE, "sinkhorn_golden.npz"), **out)
    print("wrote sinkhorn_golden.npz")


if __name__ == "__main__":
    main()
