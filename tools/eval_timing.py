"""Diagnostic: the tensor-core leaf evaluation alone, 64 Ki positions (19 random-playout steps into the synthetic battles,
as bench.py's leaf_eval leg), timed with CUDA events on the engine's stream. With a library built by
OAKB200_NVCC_EXTRA=-DOAKB200_TC_TIMING python -m oak_b200.build --force
CTA 0's MMA issuer and worker 0 also print where their cycles went (then the times include the printf)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oak_b200  # noqa: E402
from oak_ref import make_net_blob, sample_seeds, sample_teams  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
dev = torch.device("cuda:0")
eng = oak_b200.BattleEngine(n, device=0)
stream = torch.cuda.current_stream(dev)
eng.set_stream(stream.cuda_stream)
seeds_np = sample_seeds(0, n)
eng.build_battles(sample_teams(0, n), seeds_np)
hb, hd, hr = eng.download()
b, d, r = (torch.from_numpy(x).to(dev) for x in (hb, hd, hr))
seeds = torch.from_numpy(seeds_np.view(np.int64)).to(dev)
for k in range(19):
    eng.step_random_device(b, d, r, seeds, k)
eng.load_network(make_net_blob(np.random.default_rng(1234)))
v = torch.empty(n, dtype=torch.float32, device=dev)
lg = torch.empty((n, 2, 9), dtype=torch.float32, device=dev)
c = torch.empty((n, 2), dtype=torch.uint8, device=dev)
for _ in range(3):
    eng.evaluate_device(b, d, r, v, lg, c)
torch.cuda.synchronize(dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
for _ in range(iters):
    eng.evaluate_device(b, d, r, v, lg, c)
e1.record(stream)
torch.cuda.synchronize(dev)
ms = e0.elapsed_time(e1) / iters
print(f"eval_tc: {ms:.4f} ms per {n} evaluations = {n / ms / 1e3:.1f} M evals/s (checksum {float(v.sum()):.4f} {float(lg.sum()):.3f})", flush=True)
eng.close()
