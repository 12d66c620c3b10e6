"""Diagnostic: resident playout throughput of rollout_kernel for several batch sizes (CUDA events on the engine's stream).
The difference between 2^18 and 2^22 battles is the tail: lanes that found no ticket wait for the longest battles."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oak_b200  # noqa: E402
from oak_ref import sample_seeds, sample_teams  # noqa: E402

dev = torch.device("cuda:0")
for lg in (17, 18, 20, 22):
    n = 1 << lg
    eng = oak_b200.BattleEngine(n, device=0)
    stream = torch.cuda.current_stream(dev)
    eng.set_stream(stream.cuda_stream)
    seeds_np = sample_seeds(0, n)
    eng.build_battles(sample_teams(0, n), seeds_np)
    hb, hd, hr = eng.download()
    b, d, r = (torch.from_numpy(x).to(dev) for x in (hb, hd, hr))
    seeds = torch.from_numpy(seeds_np.view(np.int64)).to(dev)
    score = torch.empty(n, dtype=torch.uint8, device=dev)
    turns = torch.empty(n, dtype=torch.int16, device=dev)
    steps = torch.empty(n, dtype=torch.int32, device=dev)
    for it in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        eng.rollout_device(b, d, r, seeds, score2_out=score, turns_out=turns, steps_out=steps)
        e1.record(stream)
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1)
    total = int(steps.sum().item())
    print(f"2^{lg} playouts: {ms:.3f} ms, {total / ms / 1e6:.3f} G battle-steps/s, longest {int(steps.max().item())} steps, mean {total / n:.1f}", flush=True)
    eng.close()
