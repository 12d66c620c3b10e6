"""A/B timing of playout kernel builds on config C2 (run on a GPU box): the library comes from OAKB200_LIB (a variant
built with oak_b200.build.build(out=..., extra=[...])), the kernel form from OAKB200_ROLLOUT_GROUPED / OAKB200_ROLLOUT_SPLIT.
Device-resident inputs, CUDA events around 6 launches on the engine's stream; prints checksums so variants can be compared."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oak_b200  # noqa: E402
from oak_ref import sample_seeds, sample_teams  # noqa: E402

n = 1 << 20
dev = torch.device("cuda:0")
eng = oak_b200.BattleEngine(n)
seeds_np = sample_seeds(0, n)
eng.build_battles(sample_teams(0, n), seeds_np)
hb, hd, hr = eng.download()
b, d, r = (torch.from_numpy(x).to(dev) for x in (hb, hd, hr))
seeds = torch.from_numpy(seeds_np.view(np.int64)).to(dev)
score = torch.empty(n, dtype=torch.uint8, device=dev)
turns = torch.empty(n, dtype=torch.int16, device=dev)
steps = torch.empty(n, dtype=torch.int32, device=dev)
stream = torch.cuda.Stream(dev)
eng.set_stream(stream.cuda_stream)
with torch.cuda.stream(stream):
    for _ in range(2):
        eng.rollout_device(b, d, r, seeds, score2_out=score, turns_out=turns, steps_out=steps)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    IT = 6
    e0.record(stream)
    for _ in range(IT):
        eng.rollout_device(b, d, r, seeds, score2_out=score, turns_out=turns, steps_out=steps)
    e1.record(stream)
    torch.cuda.synchronize(dev)
ms = e0.elapsed_time(e1) / IT
total = int(steps.to(torch.int64).sum().item())
print(f"lib={os.path.basename(os.environ.get('OAKB200_LIB', 'liboakb200.so'))} {ms:.3f} ms per 2^20 playouts = {total / ms / 1e6:.3f} G battle-steps/s "
      f"(one stream, launches back to back) sum_turns={int(turns.to(torch.int64).sum().item())} score_sum={int(score.to(torch.int64).sum().item())}", flush=True)
eng.close()
