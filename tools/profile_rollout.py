"""Short rollout run for ncu: 2^18 sample-team playouts, two launches."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oak_b200  # noqa: E402
from oak_ref import sample_seeds, sample_teams  # noqa: E402

n = 1 << 18
eng = oak_b200.BattleEngine(n)
seeds = sample_seeds(0, n)
eng.build_battles(sample_teams(0, n), seeds)
for _ in range(2):
    score, turns, steps = eng.rollout(seeds)
print("steps", int(steps.sum()), "turns", int(turns.sum()))
eng.close()
