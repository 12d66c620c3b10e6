"""CUDA-event timing of the tensor-core leaf evaluation on 64 Ki resident mid-game states (the bench's leaf_eval leg
without the rest of bench.py). OAKB200_EVAL_CLUSTER=1 selects the single-CTA form."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oak_b200  # noqa: E402
from oak_ref import make_net_blob, sample_seeds, sample_teams  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 16
eng = oak_b200.BattleEngine(n)
eng.set_stream(torch.cuda.current_stream().cuda_stream)
eng.build_battles(sample_teams(0, n), sample_seeds(0, n))
eng.rollout(sample_seeds(0, n), max_steps=20, write_state=True, track_chance=True)
eng.load_network(make_net_blob(np.random.default_rng(1234)))
eng.set_eval_mode(2)
pb, pd, pr, _ = eng.resident()


class Raw:
    def __init__(self, p):
        self.p, self.shape = p, (n,)

    def data_ptr(self):
        return self.p


v = torch.empty(n, device="cuda")
lg = torch.empty((n, 2, 9), device="cuda")
cnt = torch.empty((n, 2), dtype=torch.uint8, device="cuda")
for _ in range(5):
    eng.evaluate_device(Raw(pb), Raw(pd), Raw(pr), v, lg, cnt)
torch.cuda.synchronize()
reps = 50
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    eng.evaluate_device(Raw(pb), Raw(pd), Raw(pr), v, lg, cnt)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(json.dumps({"n": n, "ms": ms, "evals_per_s": n / ms * 1e3, "tflops_executed": 2625536 * n / ms / 1e9,
                  "cluster": os.environ.get("OAKB200_EVAL_CLUSTER", "2"), "mean_value": float(v.mean())}))
eng.close()
