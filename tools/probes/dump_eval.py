"""Debug: print the raw logits buffer of the first rows (debug builds of eval_tc_kernel dump accumulators there)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oak_b200
from oak_ref import Oracle, make_net_blob, oracle_sample_batch
orc = Oracle()
n = 128
b, d, r = oracle_sample_batch(orc, 0, n)
eng = oak_b200.BattleEngine(1 << 16)
eng.load_network(make_net_blob(np.random.default_rng(1234), scale=2.0))
eng.upload(b, d, r)
eng.set_eval_mode(2)
v, lg, cnt = eng.evaluate()
np.set_printoptions(precision=4, suppress=False, linewidth=200)
for row in (0, 1, 5, 77):
    print("row", row, "value", v[row], "raw", lg[row].reshape(-1))
