"""Bring-up helper: tensor-core eval kernel vs the fp32 kernel vs the oracle (run on a GPU box)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oak_b200  # noqa: E402
from oak_ref import NetOracle, Oracle, make_net_blob, oracle_sample_batch, pick_word, sample_seeds  # noqa: E402


def states(orc, n):
    b, d, r = oracle_sample_batch(orc, 0, n)
    seeds = sample_seeds(0, n)
    stops = np.random.default_rng(5).integers(0, 120, n)
    for i in range(n):
        for step in range(int(stops[i])):
            if int(r[i]) & 15:
                break
            r[i], _, _, _ = orc.step(b[i], d[i], int(r[i]), pick_word(int(seeds[i]), step))
    return b, d, r


def main():
    orc = Oracle()
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    b, d, r = states(orc, n)
    eng = oak_b200.BattleEngine(max(n, 1 << 16))
    for act in (1, 2):
        blob = make_net_blob(np.random.default_rng(1234), activation=act, scale=2.0)
        ov, olg, ocnt = NetOracle(orc, blob).eval_batch(b, d, r)
        eng.load_network(blob)
        for m in (1, 2):
            eng.upload(b, d, r)
            eng.set_eval_mode(m)
            t0 = time.time()
            v, lg, cnt = eng.evaluate()
            dt = time.time() - t0
            mask = np.arange(9)[None, None, :] < ocnt[:, :, None]
            dv = np.abs(v - ov)
            dl = np.abs(lg - olg)[mask]
            print(f"act={act} mode={m} n={n} {dt*1e3:.1f} ms counts_equal={np.array_equal(cnt, ocnt)} "
                  f"value max|d|={dv.max():.3e} argmax={dv.argmax()} logits max|d|={dl.max():.3e} rms={np.sqrt((dl**2).mean()):.3e} "
                  f"nan={np.isnan(v).sum()}+{np.isnan(lg).sum()} pad_zero={bool(np.all(lg[~mask] == 0))}", flush=True)
            if m == 2 and dv.max() > 1e-3:
                bad = np.argsort(-dv)[:8]
                print("  worst rows", bad, "got", v[bad], "want", ov[bad])
    # throughput at 64 Ki
    n2 = 1 << 16
    idx = np.arange(n2) % n
    eng.upload(b[idx], d[idx], r[idx])
    for m in (1, 2):
        eng.set_eval_mode(m)
        eng.evaluate()
        t0 = time.time()
        for _ in range(5):
            eng.evaluate()
        dt = (time.time() - t0) / 5
        print(f"mode={m}: {n2/dt/1e6:.2f} M evals/s through the host API (incl. D2H of outputs)")
    eng.close()


if __name__ == "__main__":
    main()
