"""N>1 bookkeeping of the sharded path on CPU: world_size-2 gloo processes.
Each rank processes its contiguous slice (here with the oracle standing in for the GPU
kernels, as the checker of the slicing logic only); rank 0 reassembles per-image counts."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from shared.processing.sharding import gather_counts, max_over_ranks, my_shard, shard_bounds


def test_shard_bounds_cover_batch_once():
    for total in (0, 1, 7, 8, 255, 256, 1000):
        for world in (1, 2, 3, 4, 8):
            b = shard_bounds(total, world)
            assert b[0][0] == 0 and b[-1][1] == total
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank: int, world: int, port: int, total: int, out_path: str) -> None:
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import pipeline as O
    from workloads import synth

    lo, hi = my_shard(total)
    heads = synth.head_v5(total, seed=5, anchors=600, dup=8)  # same seed everywhere
    local = torch.tensor([len(O.detect_v5(heads[i], 0.25, 0.45)[1]) for i in range(lo, hi)], dtype=torch.int32)
    counts = gather_counts(local, total)
    slowest = max_over_ranks(float(rank + 1))
    if rank == 0:
        np.save(out_path, np.concatenate([counts.numpy(), [int(slowest)]]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [5, 8])
def test_two_rank_gloo_shards_match_single_process(tmp_path, total):
    from oracle import pipeline as O
    from workloads import synth

    out = str(tmp_path / "counts.npy")
    mp.spawn(_worker, args=(2, _free_port(), total, out), nprocs=2, join=True)
    got = np.load(out)
    heads = synth.head_v5(total, seed=5, anchors=600, dup=8)
    want = [len(O.detect_v5(heads[i], 0.25, 0.45)[1]) for i in range(total)]
    assert got[:-1].tolist() == want and got[-1] == 2
