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


# ---------------------------------------------------------------------------------------------------------------
# Frames to the writer without a collective (the per-GPU form): every rank copies its drained streams device ->
# host over ITS OWN PCIe link into a page-locked ring that lives in shared memory; the writer process reads all the rings.
# No size exchange, no padding, no rank-0 link funnel, and no rank ever waits for another one.
class FrameRing:
    """Single-producer single-consumer byte ring in a /dev/shm file, messages = [u64 length][payload padded to 8 bytes].

    Layout: 128-byte header (u64 head @0 = bytes published by the producer, u64 tail @64 = bytes released by the consumer,
    both monotonic) + `capacity` ring bytes. A message never wraps: when the space to the end is too small the producer
    writes the wrap marker (length 2^64-1) and continues at offset 0. The producer publishes `head` only once the payload
    is in memory (for device sources: after the copy's CUDA event), so the consumer never sees partial messages; the
    consumer advances `tail` only after it is done with the bytes. x86-64 total store order keeps the counter and the
    payload in order without fences; the counters are aligned 8-byte words.
    """
    HEADER = 128
    WRAP = (1 << 64) - 1

    def __init__(self, name, capacity=None, create=False, pin=False):
        import mmap
        import os
        import numpy as np
        self.path = os.path.join("/dev/shm", name)
        self.owner = create
        if create:
            capacity = (int(capacity) + 7) & ~7
            fd = os.open(self.path, os.O_CREAT | os.O_TRUNC | os.O_RDWR, 0o600)
            os.ftruncate(fd, self.HEADER + capacity)
        else:
            fd = os.open(self.path, os.O_RDWR)
            capacity = os.fstat(fd).st_size - self.HEADER
        self.capacity = capacity
        self.mm = mmap.mmap(fd, self.HEADER + capacity)
        os.close(fd)
        self.words = np.frombuffer(self.mm, np.uint64, count=self.HEADER // 8)
        self.bytes = np.frombuffer(self.mm, np.uint8, offset=self.HEADER)
        self.lengths = np.frombuffer(self.mm, np.uint64, offset=self.HEADER)
        self.reserved = int(self.words[0])   # producer side: bytes handed out (>= head)
        self.pending = []                    # producer side: (event, head after the message, keep-alive) in order
        self.pinned = None
        if pin:  # page-lock the mapping so that device -> ring copies are asynchronous DMA
            self.pinned = torch.frombuffer(self.mm, dtype=torch.uint8, offset=self.HEADER)
            rc = torch.cuda.cudart().cudaHostRegister(self.pinned.data_ptr(), self.capacity, 0)
            if int(rc) != 0:
                raise RuntimeError(f"cudaHostRegister of the frame ring failed: {rc}")

    # ---- producer
    def _publish(self, block=False):
        while self.pending:
            event, head, _ = self.pending[0]
            if event is not None:
                if block:
                    event.synchronize()
                elif not event.query():
                    return
            self.words[0] = head
            self.pending.pop(0)

    def push(self, stream, copy_stream=None, timeout_s=60.0):
        """Append one message. `stream`: uint8 numpy array (copied now) or torch CUDA tensor (copied by DMA on `copy_stream`,
        which must already be ordered after the tensor's producer; the tensor is kept alive until the copy has finished).
        Blocks only while the ring is full (the writer is behind), raising after `timeout_s`."""
        import time
        import numpy as np
        on_device = torch.is_tensor(stream)
        n = int(stream.numel()) if on_device else int(stream.size)
        need = 8 + ((n + 7) & ~7)
        if need + 8 > self.capacity:
            raise ValueError(f"a {n}-byte message does not fit a {self.capacity}-byte frame ring")
        self._publish()
        pos = self.reserved % self.capacity
        skip = self.capacity - pos if pos + need > self.capacity else 0
        deadline = None
        while self.reserved + skip + need - int(self.words[8]) > self.capacity:  # full: wait for the writer
            self._publish()
            if deadline is None:
                deadline = time.monotonic() + timeout_s
            elif time.monotonic() > deadline:
                raise RuntimeError("frame ring full: the writer has not released space for %.0f s" % timeout_s)
            time.sleep(0.0005)
        if skip:
            self.lengths[pos // 8] = self.WRAP
            self.reserved += skip
            pos = 0
        self.lengths[pos // 8] = n
        self.reserved += need
        if n == 0 or not on_device:
            if n:
                self.bytes[pos + 8:pos + 8 + n] = np.ascontiguousarray(stream, np.uint8).reshape(-1)
            self.pending.append((None, self.reserved, None))
        else:
            dst = self.pinned[pos + 8:pos + 8 + n]
            with torch.cuda.stream(copy_stream):
                dst.copy_(stream.reshape(-1), non_blocking=True)
                event = torch.cuda.Event()
                event.record(copy_stream)
            self.pending.append((event, self.reserved, stream))
        self._publish()

    def flush(self):
        """Publish everything pushed so far (waits for outstanding device copies)."""
        self._publish(block=True)

    # ---- consumer
    def poll(self, sink):
        """Hand every published message to sink(memoryview-like uint8 array), then release it. -> messages consumed."""
        count = 0
        tail = int(self.words[8])
        while tail < int(self.words[0]):
            pos = tail % self.capacity
            n = int(self.lengths[pos // 8])
            if n == self.WRAP:
                tail += self.capacity - pos
            else:
                sink(self.bytes[pos + 8:pos + 8 + n])
                tail += 8 + ((n + 7) & ~7)
                count += 1
            self.words[8] = tail
        return count

    def drained(self):
        return int(self.words[8]) == int(self.words[0])

    def close(self):
        import os
        if self.pinned is not None:
            torch.cuda.cudart().cudaHostUnregister(self.pinned.data_ptr())
            self.pinned = None
        self.words = self.bytes = self.lengths = None
        try:
            self.mm.close()
        except BufferError:  # a caller still holds a view: the mapping goes with the process
            pass
        if self.owner:
            try:
                os.unlink(self.path)
            except FileNotFoundError:
                pass


def _ring_name(rank, tag=None):
    import os
    tag = tag or os.environ.get("OAKB200_RING_TAG") or os.environ.get("MASTER_PORT") or f"pid{os.getpid()}"
    return f"oakb200_frames_{os.getuid()}_{tag}_{rank}"


class FrameSender:
    """This rank's end of the frame path: submit() takes a drained stream (torch uint8 CUDA tensor from
    SelfPlay.collect(device=True), or a numpy array) and returns at once; the bytes reach the rank's FrameRing by DMA on a
    side stream while the generator's next steps run. Nothing here talks to another rank."""

    def __init__(self, rank=None, device=None, capacity=256 << 20, tag=None):
        if rank is None:
            rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
        self.dev = _group_device(device)
        self.cuda = self.dev.type == "cuda"
        self.ring = FrameRing(_ring_name(rank, tag), capacity, create=True, pin=self.cuda)
        self.side = torch.cuda.Stream(self.dev) if self.cuda else None
        self.bytes = 0

    def submit(self, stream):
        if torch.is_tensor(stream) and stream.is_cuda:
            if stream.numel() == 0:
                return
            main = torch.cuda.current_stream(self.dev)
            mine = stream.clone()  # on the generator's stream: the drain buffer is free again when submit returns
            ready = torch.cuda.Event()
            ready.record(main)
            self.side.wait_event(ready)
            mine.record_stream(self.side)
            self.ring.push(mine, self.side)
            self.bytes += int(mine.numel())
        else:
            a = stream.numpy() if torch.is_tensor(stream) else stream
            if a.size == 0:
                return
            self.ring.push(a)
            self.bytes += int(a.size)

    def flush(self):
        self.ring.flush()

    def close(self):
        self.ring.flush()
        self.ring.close()


class FrameCollector:
    """The writer's end: one thread that polls every rank's FrameRing and hands each message to sink(rank, bytes) — a
    uint8 array VIEW of the ring, valid only during the call (write it to a file or copy it). Create it after every rank
    has created its FrameSender (a barrier at start-up; not on the data path). stop() returns once every ring is empty:
    call it after the senders have flushed."""

    def __init__(self, world, sink, tag=None, poll_s=0.0005):
        import threading
        self.rings = [FrameRing(_ring_name(r, tag)) for r in range(world)]
        self.sink, self.poll_s = sink, poll_s
        self.messages = self.bytes = 0
        self.error = None
        self._stop = False
        self.thread = threading.Thread(target=self._run, name="oakb200-frame-writer", daemon=True)
        self.thread.start()

    def _sweep(self):
        got = 0
        for r, ring in enumerate(self.rings):
            got += ring.poll(lambda view, r=r: self._take(r, view))
        self.messages += got
        return got

    def _take(self, r, view):
        self.bytes += len(view)
        self.sink(r, view)

    def wait_idle(self, timeout_s=120.0):
        """Return once every ring is empty (meaningful after the senders have flushed and a barrier)."""
        import time
        deadline = time.monotonic() + timeout_s
        while not all(r.drained() for r in self.rings):
            if self.error is not None:
                raise self.error
            if time.monotonic() > deadline:
                raise RuntimeError("frame collector: rings not drained after %.0f s" % timeout_s)
            time.sleep(0.0002)

    def _run(self):
        import time
        try:
            while True:
                if self._sweep() == 0:
                    if self._stop:
                        return
                    time.sleep(self.poll_s)
        except BaseException as e:  # surfaced by stop()
            self.error = e

    def stop(self):
        self._stop = True
        self.thread.join()
        if self.error is None and not all(r.drained() for r in self.rings):
            self._sweep()
        for r in self.rings:
            r.close()
        if self.error is not None:
            raise self.error
