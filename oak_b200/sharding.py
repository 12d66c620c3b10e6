"""Battle sharding across ranks (one process per GPU). Battles are independent, so the data path has
no collective: rank r owns battles [r * per_rank, (r + 1) * per_rank) of the logical batch, and the only
communication is the barrier + (max time, sum of work) reduction used for whole-job throughput
(reference analogue: one battle stream per std::thread, cpp/src/generate.cc:527-530)."""
import torch
import torch.distributed as dist


def shard_range(rank, world, per_rank):
    """-> (first, count) of the logical battle indices owned by `rank` (weak scaling: fixed per-rank work)."""
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    return rank * per_rank, per_rank


def split_batch(n_total, rank, world):
    """Strong-scaling split of a fixed logical batch: contiguous, sizes differ by at most one."""
    base, rem = divmod(n_total, world)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def reduce_throughput(times_ms, work, device="cpu"):
    """times_ms / work: per-rank sequences. Returns (max over ranks of each time, sum over ranks of each work)."""
    t = torch.tensor(list(times_ms), dtype=torch.float64, device=device)
    w = torch.tensor(list(work), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
    return t.tolist(), w.tolist()
