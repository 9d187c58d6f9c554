"""N > 1 host logic on CPU (gloo, world_size 2 and 3): pair-index sharding and the single gather of result records
(SURVEY §8e).  The records are synthetic — no compute kernels run here."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_pairs, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from forest_slam_b200 import sequence as seq
    mine = seq.shard_pairs(n_pairs, rank, world)
    rec = torch.zeros((mine.numel(), seq.RECORD_F64), dtype=torch.float64)
    rec[:, 0] = mine.double()            # n_matches slot carries the global pair index
    rec[:, 4] = mine.double() * 10 + rank
    allrec = seq.gather_records(rec, n_pairs)
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), allrec.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_pairs", [(2, 9), (2, 8), (3, 10)])
def test_shard_and_gather_order(tmp_path, world, n_pairs):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_pairs, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        got = np.load(tmp_path / f"rank{r}.npy")
        assert got.shape == (n_pairs, 20)
        assert np.array_equal(got[:, 0], np.arange(n_pairs))                       # pair-index order restored
        assert np.array_equal(got[:, 4], np.arange(n_pairs) * 10 + np.arange(n_pairs) % world)  # owner = i mod W


def test_shard_pairs_partition():
    from forest_slam_b200 import sequence as seq
    for world in (1, 2, 4, 8):
        parts = [seq.shard_pairs(37, r, world) for r in range(world)]
        allp = torch.cat(parts).sort().values
        assert torch.equal(allp, torch.arange(37))
        assert max(p.numel() for p in parts) - min(p.numel() for p in parts) <= 1


def test_single_rank_gather_is_identity():
    from forest_slam_b200 import sequence as seq
    rec = torch.arange(40, dtype=torch.float64).reshape(2, 20)
    assert seq.gather_records(rec, 2) is rec


def _loop_worker(rank, world, port, n_frames, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from forest_slam_b200 import sequence as seq
    lin = seq.shard_all_pairs(n_frames, rank, world, block=4)
    a, b = seq.triangular_index(n_frames, lin)
    rec = torch.stack([lin, a, b, a * 100 + b, torch.full_like(lin, rank)], dim=1)
    allrec = seq.gather_loop_records(rec, n_frames)
    np.save(os.path.join(out_dir, f"loop{rank}.npy"), allrec.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_frames", [(2, 9), (3, 11)])
def test_loop_closure_gather(tmp_path, world, n_frames):
    """Config 5 host logic: block-cyclic triangular sharding (uneven shares) + the one gather, back in pair order."""
    import itertools
    port = _free_port()
    mp.spawn(_loop_worker, args=(world, port, n_frames, str(tmp_path)), nprocs=world, join=True)
    ref = np.array(list(itertools.combinations(range(n_frames), 2)))
    for r in range(world):
        got = np.load(tmp_path / f"loop{r}.npy")
        assert got.shape == (len(ref), 5)
        assert np.array_equal(got[:, 0], np.arange(len(ref)))
        assert np.array_equal(got[:, 1:3], ref) and np.array_equal(got[:, 3], ref[:, 0] * 100 + ref[:, 1])
        assert np.array_equal(got[:, 4], (np.arange(len(ref)) // 4) % world)
