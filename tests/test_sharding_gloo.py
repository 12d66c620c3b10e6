"""N > 1 host logic on CPU: world-size-2 gloo run of the sharding + throughput reduction bench.py uses."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from oak_b200.sharding import broadcast_bytes, gather_streams, reduce_throughput, shard_range, split_batch


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    first, count = shard_range(rank, world, 1000)
    # each rank "measures" its own time and work; the job value is sum(work) / max(time)
    t, w = reduce_throughput([10.0 + 5.0 * rank, 2.0], [count * 100.0, float(first)])
    out[rank] = (first, count, t, w)
    dist.destroy_process_group()


def test_two_rank_sharding_and_reduction():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert out[0][0:2] == (0, 1000) and out[1][0:2] == (1000, 1000)
    for r in range(world):
        t, w = out[r][2], out[r][3]
        assert t == [15.0, 2.0]          # max over ranks
        assert w == [200000.0, 1000.0]   # sum over ranks
    # the logical batch is covered exactly once
    owned = np.concatenate([np.arange(out[r][0], out[r][0] + out[r][1]) for r in range(world)])
    assert np.array_equal(owned, np.arange(2000))


def test_split_batch_covers_exactly_once():
    for n in (0, 1, 7, 4096, 4099):
        for world in (1, 2, 3, 8):
            parts = [split_batch(n, r, world) for r in range(world)]
            assert sum(c for _, c in parts) == n
            pos = 0
            for f, c in parts:
                assert f == pos
                pos += c
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1


def _io_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    blob = broadcast_bytes(bytes(range(200)) * 3 if rank == 0 else b"", src=0)
    mine = np.full(1000 * rank + 17, rank + 1, np.uint8)          # ragged per-rank frame streams
    parts = gather_streams(mine, dst=0)
    empty = gather_streams(np.empty(0, np.uint8), dst=0)
    out[rank] = (blob.tobytes(), None if parts is None else [p.tobytes() for p in parts], empty if empty is None else len(empty))
    dist.destroy_process_group()


def test_weight_broadcast_and_frame_gather():
    """The two collectives of the multi-GPU path: parameter image from rank 0, frame streams to rank 0."""
    world = 2
    out = mp.Manager().dict()
    mp.spawn(_io_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    for r in range(world):
        assert out[r][0] == bytes(range(200)) * 3
    assert out[1][1] is None and out[1][2] is None
    assert out[0][1] == [bytes([1]) * 17, bytes([2]) * 1017] and out[0][2] == 2
