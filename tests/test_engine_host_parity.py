"""Differential test (no GPU needed): the CUDA engine SOURCE, compiled for the host by
tests/host_shim, against the oracle — same seeded battles, same choice stream, every step compared
byte for byte (battle[384], durations[8], actions[16], result, legal choices)."""
import numpy as np
import pytest

from oak_ref import HostShim, Oracle, pick_word, sample_seed, splitmix64


@pytest.fixture(scope="module")
def libs():
    return Oracle(), HostShim()


def playout_compare(orc, shim, battle, durations, result, seed, max_steps=1200, overrides_every=0):
    b1, b2 = battle.copy(), battle.copy()
    d1, d2 = durations.copy(), durations.copy()
    steps = 0
    while (result & 15) == 0 and steps < max_steps:
        ch = []
        for player, req in ((0, (result >> 4) & 3), (1, (result >> 6) & 3)):
            a = orc.choices(b1, player, req)
            b = shim.choices(b2, player, req)
            assert a == b, (steps, player, req, a, b)
            ch.append(a)
        pick = pick_word(seed, steps)
        c1 = ch[0][pick % len(ch[0])]
        c2 = ch[1][(pick >> 32) % len(ch[1])]
        ov = None
        if overrides_every and steps % overrides_every == 0:
            # Oak's clamped damage rolls (mcts.h:569-604): bytes 0 and 8 of the overrides
            ov = np.zeros(16, np.uint8)
            rb = b1[376 + 6], b1[376 + 7]
            ov[0] = 217 + 19 * (int(rb[0]) % 3)
            ov[8] = 217 + 19 * (int(rb[1]) % 3)
        r1, a1 = orc.update(b1, c1, c2, d1, ov)
        r2, a2 = shim.update(b2, c1, c2, d2, ov)
        assert r1 == r2, (steps, c1, c2, r1, r2)
        if not np.array_equal(b1, b2):
            diff = np.nonzero(b1 != b2)[0]
            raise AssertionError(f"step {steps} c1={c1} c2={c2}: battle bytes differ at {diff.tolist()}: "
                                 f"{b1[diff].tolist()} vs {b2[diff].tolist()}")
        assert np.array_equal(d1, d2), (steps, c1, c2, d1, d2)
        assert np.array_equal(a1, a2), (steps, c1, c2, a1.tolist(), a2.tolist())
        result = r1
        steps += 1
    return steps, result


def test_sample_team_playouts(libs):
    orc, shim = libs
    total = 0
    for i in range(0, 256, 3):
        b, d, r = orc.sample_battle(i)
        steps, res = playout_compare(orc, shim, b, d, r, sample_seed(i))
        assert (res & 15) in (1, 2, 3)
        total += steps
    assert total > 5000


def random_team(rng):
    species = rng.integers(1, 152, 6).astype(np.uint8)
    moves = np.zeros(24, np.uint8)
    for k in range(6):
        moves[4 * k:4 * k + 4] = rng.choice(np.arange(1, 165), 4, replace=False)
    n = int(rng.integers(1, 7))
    species[n:] = 0
    moves[4 * n:] = 0
    return species, moves


def test_random_team_playouts_cover_all_moves(libs):
    """Random teams over all 164 selectable moves (fuzz-style, as engine/src/test/fuzz.zig does)."""
    orc, shim = libs
    rng = np.random.default_rng(20261104)
    total = 0
    for k in range(300):
        s1, m1 = random_team(rng)
        s2, m2 = random_team(rng)
        seed = splitmix64(k)
        b = orc.battle_init(s1, m1, s2, m2, seed)
        d = np.zeros(8, np.uint8)
        r, _ = orc.update(b, 0, 0, d)
        steps, res = playout_compare(orc, shim, b, d, r, seed ^ 0xABCDEF, overrides_every=3 if k % 4 == 0 else 0)
        total += steps
    assert total > 10000


def test_nochance_variant_same_battle(libs):
    """The CHANCE=false build (used by the fast rollout kernel) must evolve the battle identically."""
    orc, shim = libs
    for i in range(0, 64, 5):
        b, d, r = orc.sample_battle(i)
        b2 = b.copy()
        seed = sample_seed(i)
        step = 0
        while (r & 15) == 0:
            c1l = orc.choices(b, 0, (r >> 4) & 3)
            c2l = orc.choices(b, 1, (r >> 6) & 3)
            pick = pick_word(seed, step)
            c1, c2 = c1l[pick % len(c1l)], c2l[(pick >> 32) % len(c2l)]
            r, _ = orc.update(b, c1, c2, d)
            r2 = shim.update_nochance(b2, c1, c2)
            assert r == r2 and np.array_equal(b, b2), (i, step)
            step += 1
