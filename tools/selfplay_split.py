"""Diagnostic: steady-state self-play pass (64 Ki games), wall time spent in step() and in collect(device=True)."""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
from oak_ref import make_net_blob  # noqa: E402
from oak_b200.selfplay import SelfPlay  # noqa: E402
pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
G = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
sp = SelfPlay(G, make_net_blob(np.random.default_rng(1)), pool, seed=1, max_updates=512)
for p in range(3):
    ts = tc = 0.0
    nbytes = 0
    for s in range(96):
        t0 = time.perf_counter()
        sp.step()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        ts += t1 - t0
        if (s + 1) % 8 == 0:
            st, g = sp.collect(device=True)
            torch.cuda.synchronize()
            tc += time.perf_counter() - t1
            nbytes += st.numel()
    print(f"pass {p}: step {ts / 96 * 1e3:.3f} ms each, collect {tc / 12 * 1e3:.3f} ms each ({nbytes / 12 / 1e6:.1f} MB), "
          f"{G * 96 / (ts + tc) / 1e6:.1f} M game-updates/s", flush=True)
sp.close()
