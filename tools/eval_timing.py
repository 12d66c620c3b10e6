"""Diagnostic: one eval_tc launch over 64 Ki start positions. With a library built by
OAKB200_NVCC_EXTRA=-DOAKB200_TC_TIMING python -m oak_b200.build --force
CTA 0's MMA issuer prints where its cycles went (wait workers / wait weights / issue / drain, by N of the stage)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oak_b200  # noqa: E402
from oak_ref import make_net_blob, sample_seeds, sample_teams  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
eng = oak_b200.BattleEngine(n, device=0)
eng.build_battles(sample_teams(0, n), sample_seeds(0, n))
eng.load_network(make_net_blob(np.random.default_rng(1)))
for i in range(3):
    t = time.perf_counter()
    eng.evaluate()
    print(f"evaluate() {i}: {1e3 * (time.perf_counter() - t):.3f} ms wall (with downloads)", flush=True)
eng.close()
