#!/usr/bin/env python3
"""Timing of the training-row builder on one GPU: replay_kernel and rows_encode_kernel separately (CUDA events on the
context's stream), the whole build with the stream resident, and end to end from pinned host memory."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oak_ref import Oracle, oracle_training_rows, play_frame_games  # noqa: E402
from oak_b200.engine import BattleEngine, rows_index_host  # noqa: E402
from oak_b200.training_rows import EncodedBattleFrames  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--games", type=int, default=2048)
ap.add_argument("--tile", type=int, default=4)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()

orc = Oracle()
t0 = time.time()
base = play_frame_games(orc, first=0, games=a.games, seed=1)
stream = np.tile(base, a.tile)
off, row_base, rows = rows_index_host(stream)
print(f"stream {stream.size / 1e6:.1f} MB, {len(off)} games, {rows} rows ({time.time() - t0:.1f}s to make)", file=sys.stderr)

f = EncodedBattleFrames(rows)
eng = f.engine
dev = f.device
d_stream = torch.from_numpy(stream).to(dev)
d_off = torch.from_numpy(off.view(np.int64)).to(dev)
d_base = torch.from_numpy(row_base.view(np.int32)).to(dev)
out = f._out()


def timed(fn, reps):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    fn()
    torch.cuda.synchronize()
    ev[0].record()
    for i in range(reps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return [ev[i].elapsed_time(ev[i + 1]) for i in range(reps)]


res = {}
t = timed(lambda: eng.rows_build_device(d_stream, stream.size, d_off, d_base, len(off), rows, out), a.reps)
res["resident_ms"] = float(np.median(t))
t = timed(lambda: eng.rows_build_device(d_stream, stream.size, d_off, d_base, len(off), rows, {}), a.reps)
res["replay_only_ms"] = float(np.median(t))  # all outputs NULL: the encode kernel still reads states but writes nothing
pinned = torch.from_numpy(stream).pin_memory()
t = timed(lambda: eng.rows_build_host(pinned, out, rows), a.reps)
res["e2e_ms"] = float(np.median(t))
row_bytes = 4 * (2376 + 458 + 12 + 18 + 18 + 3) + 8 * 18 + 4 + 4
res.update(rows=int(rows), games=int(len(off)), row_bytes=row_bytes,
           resident_rows_per_s=rows / res["resident_ms"] * 1e3, e2e_rows_per_s=rows / res["e2e_ms"] * 1e3,
           write_GBps=rows * row_bytes / res["resident_ms"] / 1e6)
t0 = time.time()
oracle_training_rows(orc, base)
res["cpu_rows_per_s"] = (rows / a.tile) / (time.time() - t0)
res["cpu_threads"] = os.cpu_count()
print(json.dumps(res))
