"""Throughput of the batched-leaf tree search (config C4): roots x particles sweep; oracle (CPU, all cores) beside it."""
import ctypes
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oak_b200  # noqa: E402
from oak_b200.search import TreeSearch  # noqa: E402
from oak_ref import sample_seeds, sample_teams  # noqa: E402

roots = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
eng = oak_b200.BattleEngine(roots)
seeds = sample_seeds(0, roots)
eng.build_battles(sample_teams(0, roots), seeds)
eng.rollout(seeds, max_steps=12, write_state=True, track_chance=True)  # mid-game roots
b, d, r = eng.download()
live = (r & 15) == 0
b, d, r, seeds = b[live], d[live], r[live], seeds[live]
n = b.shape[0]
print("roots", n)
for P in (8, 32, 64, 128):
    ts = TreeSearch(eng, n, particles=P)
    ts.set_roots(b, d, r, seeds)
    ts.run(P)  # warm-up wave
    iters = 4 * P
    t0 = time.time()
    ts.run(iters)
    dt = time.time() - t0
    w, lv = ts.stats()
    out = ts.output()
    print(f"P={P:4d}: {n * iters / dt / 1e6:.3f} M iterations/s  ({dt / (iters // P) * 1e3:.1f} ms/wave, {lv / w:.1f} levels/wave, "
          f"nodes/root {out['nodes'].mean():.0f}, mean value {out['empirical_value'].mean():.3f})", flush=True)
    ts.close()
eng.close()
