"""A/B timing of the two rollout kernels on config C2 (run on a GPU box): OAKB200_ROLLOUT_GROUPED=0/1."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oak_b200  # noqa: E402
from oak_ref import sample_seeds, sample_teams  # noqa: E402

n = 1 << 20
eng = oak_b200.BattleEngine(n)
seeds = sample_seeds(0, n)
eng.build_battles(sample_teams(0, n), seeds)
for it in range(4):
    t0 = time.time()
    score, turns, steps = eng.rollout(seeds)
    dt = time.time() - t0
    print(f"grouped={os.environ.get('OAKB200_ROLLOUT_GROUPED', '1')} iter {it}: {steps.sum() / dt / 1e9:.3f} G battle-steps/s (host API, incl. copies) "
          f"sum_turns={int(turns.sum())} score_sum={int(score.sum())}", flush=True)
eng.close()
