"""Flag cross-check (VERDICT r1 "weak" 2): the reference's known answers run the oracle under libpkmn's default
options {miss, advance, ebc}; the GPU parity target is Oak's build {miss=false, advance=false, ebc=false, key=true}. The same
template code serves both. Here every update of many playouts is run from the SAME pre-state under both flag sets: wherever
the update reached no `advance` site and no hit roll landed on the accuracy itself (the 1/256 miss), battle bytes, durations
and result must be identical, and the actions differ only in the `duration` field (key=true records the constant 2). Where
such a site was reached the two are expected to diverge — and the test checks that they do at least sometimes, so the
instrumentation is not vacuous."""
import ctypes

import numpy as np

from oak_ref import Oracle, pick_word, sample_seed, splitmix64
from test_engine_host_parity import random_team

DURATION_MASK = np.frombuffer(np.array([~(0xF << 20) & 0xFFFFFFFFFFFFFFFF] * 2, np.uint64).tobytes(), np.uint8)


def _update(lib, battle, c1, c2, durations, which):
    opt = np.zeros(40, np.uint8)
    opt[16:24] = durations
    cnt = np.zeros(2, np.uint32)
    lib.oracle_gen1_update_flags.restype = ctypes.c_uint8
    r = lib.oracle_gen1_update_flags(battle.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint8(c1), ctypes.c_uint8(c2),
                                     opt.ctypes.data_as(ctypes.c_void_p), which, cnt.ctypes.data_as(ctypes.c_void_p))
    return int(r), opt[0:16].copy(), opt[16:24].copy(), int(cnt[0]), int(cnt[1])


def _playout(orc, b, d, r, seed, stats, max_steps=600):
    lib = orc.lib
    step = 0
    while (r & 15) == 0 and step < max_steps:
        c1l = orc.choices(b, 0, (r >> 4) & 3)
        c2l = orc.choices(b, 1, (r >> 6) & 3)
        w = pick_word(seed, step)
        c1, c2 = c1l[w % len(c1l)], c2l[(w >> 32) % len(c2l)]
        b_def = b.copy()
        r_oak, a_oak, d_oak, adv_o, edge_o = _update(lib, b, c1, c2, d, 0)
        r_def, a_def, d_def, adv, edge = _update(lib, b_def, c1, c2, d, 1)
        stats["updates"] += 1
        ebc_tie = (r_def & 15) == 3 and (r_oak & 15) == 0  # Endless Battle Clause (ebc) ended it: a third flag, also expected
        if adv == 0 and edge == 0 and edge_o == 0 and not ebc_tie:
            assert r_oak == r_def, (step, c1, c2, r_oak, r_def)
            assert np.array_equal(b, b_def), (step, c1, c2, np.nonzero(b != b_def)[0].tolist())
            assert np.array_equal(d_oak, d_def), (step, c1, c2)
            assert np.array_equal(a_oak & DURATION_MASK, a_def & DURATION_MASK), (step, c1, c2, a_oak.tolist(), a_def.tolist())
            stats["equal"] += 1
        else:
            stats["sites"] += 1
            stats["diverged"] += int(not np.array_equal(b, b_def) or r_oak != r_def)
            stats["ebc"] += int(ebc_tie)
        d[:] = d_oak
        r = r_oak
        step += 1


def test_flag_sets_agree_wherever_they_must():
    orc = Oracle()
    stats = dict(updates=0, equal=0, sites=0, diverged=0, ebc=0)
    for i in range(0, 160):
        b, d, r = orc.sample_battle(i)
        _playout(orc, b, d, r, sample_seed(i), stats)
    rng = np.random.default_rng(4242)
    for k in range(250):
        s1, m1 = random_team(rng)
        s2, m2 = random_team(rng)
        seed = splitmix64(50_000 + k)
        b = orc.battle_init(s1, m1, s2, m2, seed)
        d = np.zeros(8, np.uint8)
        r, _ = orc.update(b, 0, 0, d)
        _playout(orc, b, d, r, seed ^ 0x77, stats)
    assert stats["updates"] > 25_000 and stats["equal"] > 0.9 * stats["updates"], stats
    assert stats["sites"] > 200 and stats["diverged"] > 50, stats  # the sites do occur, and the flags do matter there
