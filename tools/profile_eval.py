"""Short leaf-evaluation run for ncu: 64 Ki mid-game states made on the GPU (sample-team battles advanced
by 20 random-playout updates), default network, a few launches of each evaluation kernel."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oak_b200  # noqa: E402
from oak_ref import make_net_blob, sample_seeds, sample_teams  # noqa: E402

n = 1 << 16
eng = oak_b200.BattleEngine(n)
eng.build_battles(sample_teams(0, n), sample_seeds(0, n))
eng.rollout(sample_seeds(0, n), max_steps=20, write_state=True, track_chance=True)
eng.load_network(make_net_blob(np.random.default_rng(1234)))
for mode in (2, 1):
    eng.set_eval_mode(mode)
    for _ in range(3):
        v, lg, cnt = eng.evaluate()
    print("mode", mode, "mean value", float(v.mean()), "legal choices", int(cnt.sum()))
eng.close()
