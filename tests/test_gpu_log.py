"""R14 on the GPU: the LOG build of the step kernel writes the oracle's protocol-log bytes for every update of config C1
(sample battle 0) and of 1 024 lockstep battles x 80 steps; an undersized buffer returns 0b1100 with the oracle's prefix."""
import numpy as np
import pytest

import oak_b200
from oak_ref import Oracle, oracle_sample_batch, pick_word, sample_seeds

pytestmark = pytest.mark.gpu


def _lockstep(n, steps, first=0):
    orc = Oracle()
    b, d, r = oracle_sample_batch(orc, first, n)
    seeds = sample_seeds(first, n)
    eng = oak_b200.BattleEngine(n, device=0)
    eng.upload(b, d, r)
    total = 0
    for s in range(steps):
        c1 = np.zeros(n, np.uint8)
        c2 = np.zeros(n, np.uint8)
        for i in range(n):
            if r[i] & 15:
                continue
            l1 = orc.choices(b[i], 0, (int(r[i]) >> 4) & 3)
            l2 = orc.choices(b[i], 1, (int(r[i]) >> 6) & 3)
            w = pick_word(int(seeds[i]), s)
            c1[i], c2[i] = l1[w % len(l1)], l2[(w >> 32) % len(l2)]
        res, act, logs, lens = eng.update_log(c1, c2)
        for i in range(n):
            if r[i] & 15:
                assert lens[i] == 0 and res[i] == r[i]
                continue
            rr, oa, ol = orc.update_log(b[i], int(c1[i]), int(c2[i]), d[i])
            assert res[i] == rr, (s, i)
            assert lens[i] == len(ol) and logs[i, :lens[i]].tobytes() == ol, (s, i, list(ol), logs[i, :lens[i]].tolist())
            assert np.array_equal(act[i], oa), (s, i)
            r[i] = rr
            total += len(ol)
        gb, gd, gr = eng.download()
        assert np.array_equal(gb, b) and np.array_equal(gd, d) and np.array_equal(gr, r), s
    eng.close()
    return total


def test_c1_trace_log_bytes():
    assert _lockstep(1, 400) > 1500  # battle 0 ends after ~90 updates: 2 048 log bytes


def test_lockstep_1024_x_80_log_bytes():
    assert _lockstep(1024, 80) > 1_000_000


def test_overflow_returns_error_byte():
    orc = Oracle()
    n = 64
    b, d, r = oracle_sample_batch(orc, 0, n)
    eng = oak_b200.BattleEngine(n, device=0)
    eng.upload(b, d, r)
    ch, cnt = eng.choices()
    c1, c2 = ch[np.arange(n), 0, cnt[:, 0] - 1], ch[np.arange(n), 1, cnt[:, 1] - 1]
    res, act, logs, lens = eng.update_log(c1, c2, log_stride=6)
    for i in range(n):
        rr, _, ol = orc.update_log(b[i].copy(), int(c1[i]), int(c2[i]), d[i].copy(), cap=6)
        assert rr == 0x0C and res[i] == 0x0C
        assert logs[i, :lens[i]].tobytes() == ol
    # an errored battle is frozen: the next step leaves it alone
    res2, _, _, lens2 = eng.update_log(c1, c2)
    assert np.all(res2 == 0x0C) and np.all(lens2 == 0)
    eng.close()
