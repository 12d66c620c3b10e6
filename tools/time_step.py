#!/usr/bin/env python3
"""Lockstep step_kernel timing (one update of 2^20 battles per launch): G battle-steps/s. A/B helper:
OAKB200_LIB=<variant .so> / OAKB200_STEP_SPLIT=1 select what is measured."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oak_b200.engine import BattleEngine  # noqa: E402

n = 1 << 20
dev = torch.device("cuda:0")
eng = BattleEngine(n)
eng.set_stream(torch.cuda.current_stream(dev).cuda_stream)
pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
rng = np.random.default_rng(3)
teams = torch.from_numpy(pool[rng.integers(0, pool.shape[0], (n, 2))]).to(dev)
seeds = torch.from_numpy(rng.integers(0, 1 << 62, n).astype(np.int64)).to(dev)
b = torch.zeros((n, 384), dtype=torch.uint8, device=dev)
d = torch.zeros((n, 8), dtype=torch.uint8, device=dev)
r = torch.full((n,), 1, dtype=torch.uint8, device=dev)
eng.respawn_device(b, d, r, teams, seeds)
for s in range(8):
    eng.step_random_device(b, d, r, seeds, s)
torch.cuda.synchronize()
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    live = int(((r & 15) == 0).sum().item())
    e0.record()
    for s in range(8 + 16 * rep, 8 + 16 * rep + 16):
        eng.step_random_device(b, d, r, seeds, s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 16
    print(f"{os.environ.get('OAKB200_LIB', 'default')} split={os.environ.get('OAKB200_STEP_SPLIT', '0')}: {ms:.3f} ms/step, "
          f"{n / ms / 1e6:.3f} G battle-steps/s ({live} live at start)")
