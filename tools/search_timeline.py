import os, sys, time
import numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import oak_b200
from oak_b200.search import TreeSearch
from oak_ref import sample_seeds, sample_teams
roots = int(os.environ.get("ROOTS", "4096"))
eng = oak_b200.BattleEngine(roots)
seeds = sample_seeds(0, roots)
eng.build_battles(sample_teams(0, roots), seeds)
eng.rollout(seeds, max_steps=12, write_state=True, track_chance=True)
b, d, r = eng.download()
live = (r & 15) == 0
b, d, r, seeds = b[live], d[live], r[live], seeds[live]
n = b.shape[0]
P = 64
ts = TreeSearch(eng, n, particles=P)
ts.set_roots(b, d, r, seeds)
W = int(sys.argv[2]) if len(sys.argv) > 2 else 4
ts.run(W * P)  # warm-up: the same search once (a search object is reused turn after turn; its arenas are then mapped)
ts.set_roots(b, d, r, seeds)
os.environ["OAKB200_SEARCH_DEBUG"] = sys.argv[1] if len(sys.argv) > 1 else "1"
t0 = time.time(); ts.run(W * P); dt = time.time() - t0
print(f"{n*W*P/dt/1e6:.3f} M it/s, {dt/W*1e3:.1f} ms/wave, {W} waves")
