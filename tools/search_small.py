"""Small tree search (a few hundred roots) through the playout server: quick hang / result check."""
import sys, time
import numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import oak_b200
from oak_b200.search import TreeSearch
from oak_ref import sample_seeds, sample_teams
roots = int(sys.argv[1]) if len(sys.argv) > 1 else 256
P = int(sys.argv[2]) if len(sys.argv) > 2 else 16
eng = oak_b200.BattleEngine(roots)
seeds = sample_seeds(0, roots)
eng.build_battles(sample_teams(0, roots), seeds)
eng.rollout(seeds, max_steps=12, write_state=True, track_chance=True)
b, d, r = eng.download()
ts = TreeSearch(eng, roots, particles=P)
ts.set_roots(b, d, r, seeds)
t0 = time.time(); ts.run(4 * P); dt = time.time() - t0
o = ts.output()
print(f"roots {roots} P {P}: {dt*1e3:.1f} ms, iterations {o['iterations'][:4]}, visits {o['visit_matrix'].sum(axis=(1,2))[:4]}, value {o['empirical_value'][:3]}")
