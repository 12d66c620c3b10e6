"""N > 1 host logic on CPU: world-size-2 gloo run of the sharding + throughput reduction bench.py uses."""
import os
import socket

import numpy as np
import torch
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


def _ring_worker(rank, world, port, out, tag):
    """Every rank pushes ragged, sometimes empty drains into its own shared-memory ring; rank 0 also runs the collector.
    No collective on the data path: the barriers only order start-up and shutdown."""
    from oak_b200.sharding import FrameCollector, FrameSender
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sender = FrameSender(rank, capacity=4096, tag=tag)  # tiny: the 60 drains below wrap it many times and fill it
    dist.barrier()
    got = {r: [] for r in range(world)}
    collector = FrameCollector(world, lambda r, view: got[r].append(view.tobytes()), tag=tag) if rank == 0 else None
    rng = np.random.default_rng(100 + rank)
    sent = []
    for i in range(60):
        n = 0 if i % 7 == 3 else int(rng.integers(1, 1500 if rank else 700))  # ragged across ranks, empty drains skipped
        a = rng.integers(0, 256, n).astype(np.uint8)
        if i % 2:
            a = torch.from_numpy(a)  # host tensors take the same path
        sender.submit(a)
        if n:
            sent.append(bytes(a.numpy().tobytes() if torch.is_tensor(a) else a.tobytes()))
    sender.flush()
    dist.barrier()
    if collector is not None:
        collector.stop()
        out["got"] = {r: got[r] for r in range(world)}
    out[rank] = sent
    dist.barrier()
    sender.close()
    dist.destroy_process_group()


def test_frame_rings_carry_ragged_and_empty_drains_without_a_collective():
    import torch  # noqa: F401
    world = 2
    out = mp.Manager().dict()
    tag = f"test{os.getpid()}"
    mp.spawn(_ring_worker, args=(world, _free_port(), out, tag), nprocs=world, join=True)
    for r in range(world):
        assert out["got"][r] == out[r] and len(out[r]) > 40  # every message, whole, in order, per rank
    assert not [f for f in os.listdir("/dev/shm") if tag in f]  # the rings are unlinked


def test_frame_ring_rejects_oversized_and_times_out_when_nobody_reads():
    import pytest
    from oak_b200.sharding import FrameRing
    ring = FrameRing(f"oakb200_test_{os.getpid()}", 1024, create=True)
    try:
        with pytest.raises(ValueError):
            ring.push(np.zeros(2000, np.uint8))
        ring.push(np.zeros(600, np.uint8))
        with pytest.raises(RuntimeError):
            ring.push(np.zeros(600, np.uint8), timeout_s=0.05)
        seen = []
        assert ring.poll(lambda v: seen.append(len(v))) == 1 and seen == [600] and ring.drained()
        ring.push(np.arange(600, dtype=np.uint8))  # wraps: 608 + 608 > 1024
        assert ring.poll(lambda v: seen.append(int(v[599]))) == 1 and seen[-1] == 599 % 256
    finally:
        ring.close()
