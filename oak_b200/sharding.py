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

def _to_pinned(rows):
    """Device uint8 tensor [world, cap] -> numpy view of one cached page-locked buffer (grown geometrically, so re-pinning
    happens a few times early on and never in steady state); valid until the next call."""
    big = _pinned.get("rows")
    if big is None or big.numel() < rows.numel():
        big = _pinned["rows"] = torch.empty(4 * rows.numel(), dtype=torch.uint8).pin_memory()
    host = big[:rows.numel()].view(rows.shape)
    host.copy_(rows, non_blocking=True)
    torch.cuda.current_stream(rows.device).synchronize()
    return host.numpy()


def gather_streams(stream, dst=0, device=None):
    """Gather every rank's frame stream (ragged across ranks, possibly empty) to rank `dst`. `stream` is a uint8 numpy array
    or a torch uint8 CUDA tensor (SelfPlay.collect(device=True): the drained stream never visits this rank's host).
    Returns the list of per-rank numpy arrays on `dst` (rank order) and None elsewhere. Sizes travel first, payloads are
    padded to the largest. This is the only collective of the self-play path: the game records themselves
    are independent, so the writer just concatenates them. Only `dst` receives (dist.gather); with NCCL its copy comes
    back through a cached page-locked buffer and the returned arrays are views of it, valid until the next call."""
    import numpy as np
    dev = _group_device(device)
    on_device = torch.is_tensor(stream)
    if not on_device:
        stream = np.ascontiguousarray(stream, np.uint8)
    size = int(stream.numel()) if on_device else int(stream.size)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        if not on_device:
            return [stream]
        return [_to_pinned(stream.reshape(1, -1))[0]] if size else [np.empty(0, np.uint8)]
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = torch.zeros(world, dtype=torch.int64, device=dev)
    sizes[rank] = size
    dist.all_reduce(sizes, op=dist.ReduceOp.SUM)
    sizes = sizes.cpu()
    cap = int(sizes.max().item())
    if cap == 0:
        return [np.empty(0, np.uint8) for _ in range(world)] if rank == dst else None
    mine = torch.empty(cap, dtype=torch.uint8, device=dev)
    mine[:size] = stream.to(dev) if on_device else torch.from_numpy(stream).to(dev)
    if rank != dst:
        dist.gather(mine, None, dst=dst)
        return None
    rows = torch.empty((world, cap), dtype=torch.uint8, device=dev)
    dist.gather(mine, list(rows.unbind(0)), dst=dst)
    out = _to_pinned(rows) if rows.is_cuda else rows.numpy()
    return [out[r, :int(sizes[r].item())] for r in range(world)]


class StreamGatherer:
    """gather_streams one drain behind the generator (CUDA/NCCL groups): submit() exchanges the sizes (a few bytes, the only
    point where the host waits for the other ranks), copies the drained stream aside and queues the payload gather and the
    writer's device -> page-locked-host copy on a side stream; the generator's next steps run meanwhile. The result of a
    submit is returned by the NEXT submit() or by flush(): on the writer rank a list of per-rank numpy arrays (views of one
    of two page-locked buffers, valid until the submit after next), elsewhere None."""

    def __init__(self, dst=0, device=None):
        self.dst, self.dev = dst, _group_device(device)
        self.multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        self.world = dist.get_world_size() if self.multi else 1
        self.rank = dist.get_rank() if self.multi else 0
        self.side = torch.cuda.Stream(self.dev)
        self.pinned = [None, None]
        self.turn = 0
        self.pending = None

    def _pinned(self, nbytes):
        buf = self.pinned[self.turn]
        if buf is None or buf.numel() < nbytes:
            buf = self.pinned[self.turn] = torch.empty(4 * nbytes, dtype=torch.uint8).pin_memory()
        return buf[:nbytes]

    def _finish(self):
        p, self.pending = self.pending, None
        if p is None:
            return None
        done, host, sizes, keep = p
        done.synchronize()
        del keep
        if self.multi and self.rank != self.dst:
            return None
        import numpy as np
        if host is None:  # every rank's drain was empty
            return [np.empty(0, np.uint8) for _ in range(self.world)]
        out = host.numpy()
        return [out[r, :sizes[r]] if sizes[r] else np.empty(0, np.uint8) for r in range(self.world)]

    def submit(self, stream):
        """stream: torch uint8 CUDA tensor (SelfPlay.collect(device=True)). -> the previous submit's result."""
        prev = self._finish()
        main = torch.cuda.current_stream(self.dev)
        size = int(stream.numel())
        if self.multi:
            sizes_t = torch.zeros(self.world, dtype=torch.int64, device=self.dev)
            sizes_t[self.rank] = size
            dist.all_reduce(sizes_t, op=dist.ReduceOp.SUM)
            sizes = [int(x) for x in sizes_t.cpu().tolist()]
        else:
            sizes = [size]
        cap = max(sizes)
        done = torch.cuda.Event()
        if cap == 0:
            done.record(main)
            self.pending = (done, None, sizes, None)
            return prev
        mine = torch.empty(cap, dtype=torch.uint8, device=self.dev)
        mine[:size] = stream  # on the generator's stream: the drain buffer is free again when submit returns
        ready = torch.cuda.Event()
        ready.record(main)
        host = None
        with torch.cuda.stream(self.side):
            self.side.wait_event(ready)
            mine.record_stream(self.side)
            rows = None
            if not self.multi:
                rows = mine.view(1, cap)
            elif self.rank == self.dst:
                rows = torch.empty((self.world, cap), dtype=torch.uint8, device=self.dev)
                dist.gather(mine, list(rows.unbind(0)), dst=self.dst)
            else:
                dist.gather(mine, None, dst=self.dst)
            if rows is not None:
                host = self._pinned(rows.numel()).view(rows.shape)
                host.copy_(rows, non_blocking=True)
            done.record(self.side)
        self.turn ^= 1
        self.pending = (done, host, sizes, (mine, rows))
        return prev

    def flush(self):
        return self._finish()
