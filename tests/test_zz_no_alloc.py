"""fs_recover_pose_static (no counterpart in the reference): the
allocation-free variant of fs_recover_pose whose per-pair scratch is a file-scope __device__ array.  Called through the
raw C ABI, twice in a row on different batches, and compared bit for bit with fs_recover_pose (same kernels, so the
tolerance of the recoverPose tests — mask / count exact — tightens to equality for R and t as well).

Collected last on purpose."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


def _skew(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]], np.float64)


def _batch(sizes, n_max, seed):
    """Ragged correspondence batch + one essential matrix per pair (true pose of the synthetic two-view scene)."""
    P = len(sizes)
    mk0 = np.zeros((P, n_max, 2), np.float32)
    mk1 = np.zeros((P, n_max, 2), np.float32)
    E = np.zeros((P, 9), np.float64)
    for p, n in enumerate(sizes):
        rvec = (0.01 + 0.002 * p, 0.05 - 0.003 * p, -0.02)
        t = np.array((0.1, -0.02 + 0.01 * p, 0.5))
        a, b, _, _, _ = util.two_view(max(n, 1), seed + p, outliers=0.2, rvec=rvec, t=tuple(t))
        mk0[p, :n], mk1[p, :n] = a[:n], b[:n]
        E[p] = (_skew(t / np.linalg.norm(t)) @ util.rodrigues(np.array(rvec))).ravel()
    return mk0, mk1, E


def _run(lib, entry, E, mk0, mk1, sizes, status, n_max):
    import torch
    from forest_slam_b200 import _lib
    P = len(sizes)
    dev = torch.device("cuda:0")
    d = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)
    tE, t0, t1 = d(E, torch.float64), d(mk0, torch.float32), d(mk1, torch.float32)
    nc, st = d(np.array(sizes), torch.int32), d(np.array(status), torch.int32)
    R = torch.full((P, 9), 7.0, dtype=torch.float64, device=dev)
    t = torch.full((P, 3), 7.0, dtype=torch.float64, device=dev)
    mask = torch.full((P, n_max), 3, dtype=torch.uint8, device=dev)
    ng = torch.full((P,), -5, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    args = [tE.data_ptr(), t0.data_ptr(), t1.data_ptr(), nc.data_ptr(), st.data_ptr(), P, n_max, util.FOCAL,
            util.PP[0], util.PP[1], 50.0, R.data_ptr(), t.data_ptr(), mask.data_ptr(), ng.data_ptr()]
    if entry == "static":
        before = torch.cuda.memory_allocated()
        rc = lib.fs_recover_pose_static(*args, stream)
        assert torch.cuda.memory_allocated() == before
    else:
        nbytes = lib.fs_recover_pose_workspace_bytes(P)
        ws = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
        rc = lib.fs_recover_pose(*args, ws.data_ptr(), nbytes, stream)
    _lib.check(rc, f"fs_recover_pose[{entry}]")
    torch.cuda.synchronize()
    return R.cpu().numpy(), t.cpu().numpy(), mask.cpu().numpy(), ng.cpu().numpy()


def test_recover_pose_static_equals_workspace_variant_twice():
    import torch
    assert torch.cuda.is_available()
    from forest_slam_b200 import _lib
    lib = _lib.load()
    assert lib.fs_recover_pose_static_max_pairs() == 4096
    # two different batches back to back: the second call finds the first call's contents in the shared scratch
    cases = [([700, 40, 0, 333, 5, 700], [0, 0, 0, 0, 0, 2], 704, 4100),
             ([64, 512, 300, 9], [0, 1, 0, 0], 512, 4200)]
    for sizes, status, n_max, seed in cases:
        mk0, mk1, E = _batch(sizes, n_max, seed)
        ref = _run(lib, "workspace", E, mk0, mk1, sizes, status, n_max)
        got = _run(lib, "static", E, mk0, mk1, sizes, status, n_max)
        for name, a, b in zip(("R", "t", "mask", "n_good"), ref, got):
            assert np.array_equal(a, b), (name, sizes)
        # the result is a real pose, not two equal failures: pairs that ran have inliers and a unit translation
        R, t, mask, ng = got
        for p, (n, s) in enumerate(zip(sizes, status)):
            if s != 0:
                assert ng[p] == 0 and not mask[p].any() and np.array_equal(R[p], np.eye(3).ravel())
            elif n >= 40:
                assert ng[p] > n // 4 and abs(np.linalg.norm(t[p]) - 1) < 1e-12
                assert int((mask[p] == 255).sum()) == ng[p] and not mask[p, n:].any()


def test_recover_pose_static_rejects_oversize_batches():
    from forest_slam_b200 import _lib
    lib = _lib.load()
    too_many = lib.fs_recover_pose_static_max_pairs() + 1
    rc = lib.fs_recover_pose_static(None, None, None, None, None, too_many, 16, util.FOCAL, util.PP[0], util.PP[1],
                                    50.0, None, None, None, None, None)
    assert rc == -1 and lib.fs_status_string(rc).startswith(b"FS_ERR_BAD_ARG")
    msg = lib.fs_last_error_string().decode()
    assert "exceeds the static scratch" in msg and "4096" in msg
