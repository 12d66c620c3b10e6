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


def _group_device(device):
    return torch.device(device) if device is not None else torch.device("cpu")


def broadcast_bytes(data, src=0, device=None):
    """Broadcast a byte string (e.g. a network parameter file image) from rank `src` to every rank; returns a uint8
    numpy array on all ranks. NCCL groups need `device` = this rank's CUDA device; gloo groups use the CPU.
    (Reference analogue: every generate worker re-reads the learner's newest parameter file, cpp/src/generate.cc.)"""
    import numpy as np
    dev = _group_device(device)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return np.frombuffer(bytes(data), np.uint8).copy()
    rank = dist.get_rank()
    size = torch.tensor([len(data) if rank == src else 0], dtype=torch.int64, device=dev)
    dist.broadcast(size, src)
    if rank == src:
        buf = torch.from_numpy(np.frombuffer(bytes(data), np.uint8).copy()).to(dev)
    else:
        buf = torch.empty(int(size.item()), dtype=torch.uint8, device=dev)
    dist.broadcast(buf, src)
    return buf.cpu().numpy()


_pinned = {}
_STAGE_BYTES = 32 << 20


def _device_to_numpy(torch, src, out):
    """Device uint8 tensor -> the numpy array `out`, through one cached page-locked staging buffer of fixed size (a
    pageable .cpu() of a few hundred MB costs more than the self-play pass that produced it)."""
    stage = _pinned.get("stage")
    if stage is None:
        stage = _pinned["stage"] = torch.empty(_STAGE_BYTES, dtype=torch.uint8).pin_memory()
    flat, host = src.reshape(-1), stage.numpy()
    stream = torch.cuda.current_stream(src.device)
    for o in range(0, flat.numel(), _STAGE_BYTES):
        k = min(_STAGE_BYTES, flat.numel() - o)
        stage[:k].copy_(flat[o:o + k], non_blocking=True)
        stream.synchronize()
        out[o:o + k] = host[:k]


def gather_streams(stream, dst=0, device=None):
    """Gather every rank's frame stream (uint8 array, possibly empty, ragged across ranks) to rank `dst`. Returns the
    list of per-rank arrays on `dst` (rank order) and None elsewhere. Sizes travel first, payloads are padded to the
    largest. This is the only collective of the self-play path: the game records themselves are
    independent, so the writer just concatenates them. Only `dst` receives (dist.gather); with NCCL its copy comes back
    through page-locked memory (see below: the arrays may be views of a cached buffer, valid until the next call)."""
    import numpy as np
    dev = _group_device(device)
    stream = np.ascontiguousarray(stream, np.uint8)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [stream]
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = torch.zeros(world, dtype=torch.int64, device=dev)
    sizes[rank] = stream.size
    dist.all_reduce(sizes, op=dist.ReduceOp.SUM)
    sizes = sizes.cpu()
    cap = int(sizes.max().item())
    if cap == 0:
        return [np.empty(0, np.uint8) for _ in range(world)] if rank == dst else None
    mine = torch.empty(cap, dtype=torch.uint8, device=dev)
    mine[:stream.size] = torch.from_numpy(stream).to(dev)
    if rank != dst:
        dist.gather(mine, None, dst=dst)
        return None
    rows = torch.empty((world, cap), dtype=torch.uint8, device=dev)
    dist.gather(mine, list(rows.unbind(0)), dst=dst)
    if rows.is_cuda:
        # the writer's copy lands in one cached page-locked buffer (allocated with 4x head-room the first time: streams
        # grow until games end at a steady rate) and the returned arrays are views of it, valid until the next call; a
        # gather that has outgrown it goes through the fixed staging buffer into fresh memory instead of re-pinning
        big = _pinned.get("rows")
        if big is None:
            big = _pinned["rows"] = torch.empty(4 * world * cap, dtype=torch.uint8).pin_memory()
        if big.numel() >= world * cap:
            host = big[:world * cap].view(world, cap)
            host.copy_(rows, non_blocking=True)
            torch.cuda.current_stream(rows.device).synchronize()
            out = host.numpy()
        else:
            out = np.empty((world, cap), np.uint8)
            _device_to_numpy(torch, rows, out.reshape(-1))
    else:
        out = rows.numpy()
    return [out[r, :int(sizes[r].item())] for r in range(world)]
