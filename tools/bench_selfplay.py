"""Self-play throughput (network policy, no search): game-updates/s for a few game counts."""
import json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oak_ref import make_net_blob  # noqa: E402
from oak_b200.selfplay import SelfPlay  # noqa: E402
pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
blob = make_net_blob(np.random.default_rng(1))
for G in (4096, 16384, 65536):
    sp = SelfPlay(G, blob, pool, seed=1, max_updates=512)
    sp.run(16)
    torch.cuda.synchronize()
    t = time.perf_counter()
    games = sp.run(96, collect_every=8)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    print(f"games {G}: {G * 96 / dt / 1e6:.1f} M game-updates/s ({dt / 96 * 1e3:.3f} ms/step, {games} games finished)")
    sp.close()
