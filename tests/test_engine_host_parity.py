"""Differential test (no GPU needed): the CUDA engine SOURCE, compiled for the host by
tests/host_shim, against the oracle — same seeded battles, same choice stream, every step compared
byte for byte (battle[384], durations[8], actions[16], result, legal choices)."""
import numpy as np
import pytest

from oak_ref import HostShim, Oracle, pick_word, sample_seed, splitmix64


@pytest.fixture(scope="module")
def libs():
    return Oracle(), HostShim()


def playout_compare(orc, shim, battle, durations, result, seed, max_steps=1200, overrides_every=0, log=False, log_stats=None):
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
        if log:  # the LOG build of the engine against the oracle's log writer: same bytes, at most MAX_LOGS = 180 of them
            r1, a1, l1 = orc.update_log(b1, c1, c2, d1, ov)
            r2, a2, l2 = shim.update_log(b2, c1, c2, d2, ov)
            assert l1 == l2, (steps, c1, c2, list(l1), list(l2))
            assert 0 < len(l1) <= 180
            if log_stats is not None:
                log_stats["bytes"] += len(l1)
                log_stats["max"] = max(log_stats["max"], len(l1))
                log_stats["ops"].update(_log_ops(l1))
        else:
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


def _log_ops(log):
    """Message opcodes of one update's log (lengths per common/protocol.zig:337-824)."""
    ops, i = [], 0
    fixed = {1: 1, 2: 1, 4: 9, 7: 4, 8: 3, 9: 2, 13: 4, 14: 4, 15: 1, 16: 3, 17: 2, 18: 3, 19: 3, 20: 2, 22: 1, 24: 3, 25: 1,
             26: 2, 27: 2, 28: 2, 29: 3, 30: 3}
    while i < len(log):
        op = log[i]
        ops.append(op)
        if op in fixed:
            i += fixed[op]
        elif op == 3:  # move: src, move, target, None | From move
            i += 5 if log[i + 4] == 0 else 6
        elif op == 5:  # cant: ident, reason [, move] (Disable = 5)
            i += 4 if log[i + 2] == 5 else 3
        elif op == 6:  # faint: ident [, None when done]
            i += 3 if i + 2 < len(log) and log[i + 2] == 0 else 2
        elif op in (10, 11):  # damage / heal: ident, hp, maxhp, status, reason [, of] (RecoilOf = 5, Drain = 2)
            i += 9 if log[i + 7] == (5 if op == 10 else 2) else 8
        elif op == 12:  # status: ident, value, reason [, move] (From = 2)
            i += 5 if log[i + 3] == 2 else 4
        elif op == 21:  # activate: ident, reason
            i += 3
        elif op == 23:  # start: ident, reason [, move | types, target] (TypeChange = 9, Disable = 10, Mimic = 11)
            i += {9: 5, 10: 4, 11: 4}.get(log[i + 2], 3)
        else:
            raise AssertionError(f"unknown opcode {op} at {i} in {list(log)}")
    assert i == len(log), list(log)
    return ops


def test_log_build_equals_oracle_log(libs):
    """R14: the engine's LOG build writes the oracle's protocol-log bytes on every update — config C1 (sample battle 0, the
    single fixed-seed trace), more sample-team battles, and random teams over all 164 moves so that every message kind the
    Generation I engine can emit is produced."""
    import collections
    orc, shim = libs
    stats = {"bytes": 0, "max": 0, "ops": collections.Counter()}
    for i in list(range(0, 48, 1)):
        b, d, r = orc.sample_battle(i)
        playout_compare(orc, shim, b, d, r, sample_seed(i), log=True, log_stats=stats)
    rng = np.random.default_rng(77)
    for k in range(400):
        s1, m1 = random_team(rng)
        s2, m2 = random_team(rng)
        seed = splitmix64(1000 + k)
        b = orc.battle_init(s1, m1, s2, m2, seed)
        d = np.zeros(8, np.uint8)
        r, _, l1 = orc.update_log(b, 0, 0, d)
        assert _log_ops(l1) == [4, 4, 7]  # two switch-ins and turn 1
        playout_compare(orc, shim, b, d, r, seed ^ 0x5555, overrides_every=3 if k % 4 == 0 else 0, log=True, log_stats=stats)
    # |tie|: the 1000-turn limit (mechanics.zig:1650-1653) and a double knock-out of both last Pokémon (Explosion at 1 HP)
    b, d, r = orc.sample_battle(5)
    b[368:370] = np.frombuffer(np.uint16(999).tobytes(), np.uint8)
    playout_compare(orc, shim, b, d, r, 99, max_steps=1, log=True, log_stats=stats)
    one = np.zeros(6, np.uint8), np.zeros(24, np.uint8)
    one[0][0], one[1][0] = 101, 153  # Electrode, Explosion
    b = orc.battle_init(one[0], one[1], one[0], one[1], 7)
    d = np.zeros(8, np.uint8)
    r, _ = orc.update(b, 0, 0, d)
    b[18:20] = b[184 + 18:184 + 20] = (1, 0)  # both sides' only Pokémon at 1 HP: whoever explodes first takes the other along
    steps, res = playout_compare(orc, shim, b, d, r, 5, log=True, log_stats=stats)
    assert (res & 15) == 3
    assert stats["bytes"] > 400000 and stats["max"] <= 180
    missing = set(range(1, 31)) - set(stats["ops"])
    assert not missing, f"message kinds never produced: {sorted(missing)}"


def test_log_overflow_returns_error_byte(libs):
    """bindings/c.zig:6,181: a log buffer that is too small makes update return 0b1100; what did fit is the same prefix."""
    orc, shim = libs
    b, d, r = orc.sample_battle(3)
    c1 = orc.choices(b, 0, (r >> 4) & 3)[-1]
    c2 = orc.choices(b, 1, (r >> 6) & 3)[-1]
    _, _, full = orc.update_log(b.copy(), c1, c2, d.copy())
    for cap in (0, 1, 4, len(full) - 1):
        r1, _, l1 = orc.update_log(b.copy(), c1, c2, d.copy(), cap=cap)
        r2, _, l2 = shim.update_log(b.copy(), c1, c2, d.copy(), cap=cap)
        assert r1 == 0x0C and r2 == 0x0C, (cap, r1, r2)
        assert l1 == l2 and full.startswith(l1), (cap, list(l1), list(l2))
    r1, _, l1 = orc.update_log(b.copy(), c1, c2, d.copy(), cap=len(full))
    r2, _, l2 = shim.update_log(b.copy(), c1, c2, d.copy(), cap=len(full))
    assert r1 == r2 and (r1 & 15) == 0 and l1 == l2 == full


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


class _SplitShim:
    """HostShim whose update / choices go through the SPLIT on-chip layout."""

    def __init__(self, shim):
        self.update = shim.update_split
        self.choices = shim.choices_split


def test_split_layout_same_battle(libs):
    """The SPLIT layout (48 on-chip words; stored stats / move ids / species / types / level read from the record) evolves
    every byte of the battle, the durations and the actions exactly as the oracle does: sample teams, and random teams over
    all 164 moves (Transform, Mimic, Conversion, Haze, boosts under Transform: everything that touches the constant side)."""
    orc, shim = libs
    split = _SplitShim(shim)
    total = 0
    for i in range(0, 256, 5):
        b, d, r = orc.sample_battle(i)
        steps, res = playout_compare(orc, split, b, d, r, sample_seed(i))
        assert (res & 15) in (1, 2, 3)
        total += steps
    rng = np.random.default_rng(8)
    for k in range(300):
        s1, m1 = random_team(rng)
        s2, m2 = random_team(rng)
        seed = splitmix64(7000 + k)
        b = orc.battle_init(s1, m1, s2, m2, seed)
        d = np.zeros(8, np.uint8)
        r, _ = orc.update(b, 0, 0, d)
        steps, res = playout_compare(orc, split, b, d, r, seed ^ 0x1234, overrides_every=3 if k % 4 == 0 else 0)
        total += steps
    assert total > 15000


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
