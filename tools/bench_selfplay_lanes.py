"""Self-play generation rate of PipelinedSelfPlay (lanes x games) without the frame path: game-updates/s."""
import json, os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oak_b200.selfplay import PipelinedSelfPlay
from oak_ref import make_net_blob
lanes = int(sys.argv[1]) if len(sys.argv) > 1 else 2
games = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
blob = make_net_blob(np.random.default_rng(1234)).tobytes()
psp = PipelinedSelfPlay(games, blob, pool, seed=99, lanes=lanes, max_updates=512)
psp.run(96, lambda st, g: None)
torch.cuda.synchronize()
for rep in range(3):
    t0 = time.perf_counter()
    psp.run(96, lambda st, g: None)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"connections={os.environ.get('CUDA_DEVICE_MAX_CONNECTIONS')} lanes={lanes} x {games}: {lanes * games * 96 / dt / 1e6:.1f} M game-updates/s")
psp.close()
