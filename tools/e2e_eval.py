import sys, time; sys.path.insert(0,"tests"); sys.path.insert(0,".")
import numpy as np, torch, oak_b200
from oak_ref import make_net_blob, sample_seeds, sample_teams
n = 1 << 16
eng = oak_b200.BattleEngine(n)
eng.build_battles(sample_teams(0, n), sample_seeds(0, n)); eng.rollout(sample_seeds(0, n), max_steps=20, write_state=True, track_chance=True)
b, d, r = eng.download()
eng.load_network(make_net_blob(np.random.default_rng(1)))
pb, pd, pr = (torch.from_numpy(x).pin_memory() for x in (b, d, r))
pv = torch.empty(n).pin_memory(); pl = torch.empty((n, 2, 9)).pin_memory(); pc = torch.empty((n, 2), dtype=torch.uint8).pin_memory()
import os
for chunk in (0, 4736, 9472, 18944, 37888):
    if chunk: os.environ["OAKB200_EVAL_CHUNK"] = str(chunk)
    for _ in range(3): eng.evaluate_host(pb, pd, pr, pv, pl, pc)
    t = time.perf_counter()
    for _ in range(10): eng.evaluate_host(pb, pd, pr, pv, pl, pc)
    dt = (time.perf_counter() - t) / 10
    print("chunk", chunk or "default", "e2e %.1f M evals/s (%.3f ms)" % (n / dt / 1e6, dt * 1e3))
t = time.perf_counter()
for _ in range(10):
    eng.upload(pb, pd, pr); eng.lib.oakb200_eval(eng._h, pv.data_ptr(), pl.data_ptr(), pc.data_ptr())
print("upload+eval %.1f M evals/s" % (n * 10 / (time.perf_counter() - t) / 1e6))
