"""Multi-GPU host logic on CPU: scene sharding + the single histogram all-reduce, world_size 2 over gloo.
Reference: run/validation.py:269-286 (manual scene split), :441-450 (dead all_reduce the reduction replaces)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from geopurify_b200 import validation as V

C, SCENES = 20, 11


def scene_hist(i):
    """Deterministic fake per-scene [3,C] histogram block (intersection <= target <= union shape not required)."""
    rng = np.random.default_rng(5557 + i)
    inter = rng.integers(0, 4000, C)
    tgt = inter + rng.integers(0, 3000, C)
    out = inter + rng.integers(0, 3000, C)
    return np.stack([inter, out + tgt - inter, tgt]).astype(np.int64)


def test_shard_scenes_partitions_every_scene_once():
    for world in (1, 2, 3, 8):
        got = sorted(i for r in range(world) for i in V.shard_scenes(SCENES, r, world))
        assert got == list(range(SCENES))
    sizes = [100, 5, 70, 30, 30, 90]
    a, b = V.shard_scenes(6, 0, 2, sizes), V.shard_scenes(6, 1, 2, sizes)
    assert a == [0, 2, 4] and b == [5, 3, 1]          # longest first, ties by index
    with pytest.raises(ValueError):
        V.shard_scenes(4, 2, 2)


def test_reduce_is_noop_without_process_group():
    h = torch.ones((3, C), dtype=torch.int64)
    assert V.reduce_histograms(h).sum() == 3 * C
    with pytest.raises(TypeError):
        V.reduce_histograms(torch.ones((3, C)))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    total = torch.zeros((3, C), dtype=torch.int64)
    for i in V.shard_scenes(SCENES, rank, world):
        total += torch.as_tensor(scene_hist(i))
    V.reduce_histograms(total)
    q.put((rank, total.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_histogram_allreduce_equals_single_process_sum():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = sum(scene_hist(i) for i in range(SCENES))
    assert (got[0] == want).all() and (got[1] == want).all()
    assert V.mean_iou(*got[0]) == V.mean_iou(*want)
