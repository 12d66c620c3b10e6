"""The reference's known-answer tests that need hand transcription (tests/kat_hand.py), on the oracle and — for the calc
override script — on the engine source compiled for the host. GPU counterpart: tests/test_gpu_kat_hand.py."""
import numpy as np
import pytest

import kat_hand as K
from kat_harness import L, LogExpected, move, swtch
from oak_ref import HostShim, Oracle


def _order(b, p):
    return b[p * 184 + 176:p * 184 + 182].tolist()


def test_switching_order():
    """test.zig:133-170: five double switches with NULL options, then one with a 22-byte log buffer."""
    b = K.battle_init(0x12345678, [dict(species=L.Abra, moves=[L.Teleport])] * 6, [dict(species=L.Gastly, moves=[L.Lick])] * 6)
    b[368:370] = (1, 0)  # battle.turn = 1
    script = [((3, 2), [3, 2, 1, 4, 5, 6], [2, 1, 3, 4, 5, 6]), ((5, 5), [5, 2, 1, 4, 3, 6], [5, 1, 3, 4, 2, 6]),
              ((6, 3), [6, 2, 1, 4, 3, 5], [3, 1, 5, 4, 2, 6]), ((3, 3), [1, 2, 6, 4, 3, 5], [5, 1, 3, 4, 2, 6]),
              ((2, 4), [2, 1, 6, 4, 3, 5], [4, 1, 3, 5, 2, 6])]
    for (s1, s2), o1, o2 in script:
        r, _ = K.update_default(b, swtch(s1), swtch(s2))
        assert r == 0x50 and _order(b, 0) == o1 and _order(b, 1) == o2
    exp = LogExpected()
    shadow = bytearray(b.tobytes())
    from kat_harness import PokemonV
    exp.switched([K.ident(0, 3), PokemonV(shadow, 2 * 24)])        # &p1.pokemon[2]
    exp.switched([K.ident(1, 2), PokemonV(shadow, 184 + 1 * 24)])  # &p2.pokemon[1]
    exp.turn([7])
    assert len(exp.buf) == 22
    r, log = K.update_default(b, swtch(5), swtch(5), log_cap=22)
    assert r == 0x50 and _order(b, 0) == [3, 1, 6, 4, 2, 5] and _order(b, 1) == [2, 1, 3, 5, 4, 6]
    assert log == bytes(exp.buf)


def test_choices():
    """test.zig:1076-1094: a random battle from PSRNG 0x2718281; P1's choice lists for Move / Switch / Pass requests."""
    orc = Oracle()
    b = K.battle_random(K.PSRNG(0x2718281))
    assert all(b[176 + i] == i + 1 for i in range(6))  # six Pokémon each side for this seed
    r, _ = K.update_default(b, 0, 0)
    assert r == 0x50
    assert orc.choices(b, 0, 1) == [swtch(2), swtch(3), swtch(4), swtch(5), swtch(6), move(1), move(2), move(3), move(4)]
    assert orc.choices(b, 0, 2) == [swtch(2), swtch(3), swtch(4), swtch(5), swtch(6)]
    assert orc.choices(b, 0, 0) == [0]


def test_rng_overrides_full_on_the_oracle():
    """test.zig:10131-10327 with every roll answered through the calc overrides (R13: all the Action fields, not only damage)."""
    orc = Oracle()
    hp = K.run_full(lambda b, c1, c2, d, ov: orc.update_log(b, c1, c2, d, ov))
    assert hp == (703 - 52 - 43 - 1 - 43 - 19 - 19 - 43 - 107, 703 - 52 - 32 - 72)


def test_rng_overrides_full_on_the_engine_source():
    shim = HostShim()
    K.run_full(lambda b, c1, c2, d, ov: shim.update_log(b, c1, c2, d, ov))


def test_rng_overrides_partial():
    """test.zig:10329-10387: only the damage roll overridden (217), everything else from PSRNG 0x12345678 under libpkmn's
    default flags: pins the live roll order (speed, hit, crit, secondary) together with a partial override."""
    from kat_harness import PokemonV, build_battle
    import struct
    buf = build_battle([dict(species=L.Zapdos, moves=[L.Thunder])], [dict(species=L.Tauros, moves=[L.BodySlam])])
    struct.pack_into("<Q", buf, 376, 0x12345678)
    b = np.frombuffer(bytes(buf), np.uint8).copy()
    opt = np.zeros(40, np.uint8)
    r, _ = K.update_default(b, 0, 0, opt)
    assert r == 0x50
    shadow = bytearray(b.tobytes())
    p1, p2 = PokemonV(shadow, 0), PokemonV(shadow, 184)
    exp = LogExpected()
    M = K.M
    exp.move([K.ident(1, 1), M["BodySlam"], K.ident(0, 1)])
    p1.hp -= 102
    exp.damage([K.ident(0, 1), p1, L.None_])
    p1.status = K.PAR
    exp.status([K.ident(0, 1), p1.status, L.None_])
    exp.move([K.ident(0, 1), M["Thunder"], K.ident(1, 1)])
    p2.hp -= 191
    exp.damage([K.ident(1, 1), p2, L.None_])
    exp.turn([2])
    opt[:16] = 0
    opt[24:40] = K.overrides16(dict(damage=217), dict(damage=217))
    r, log = K.update_default(b, move(1), move(1), opt, log_cap=180)
    assert r == 0x50 and log == bytes(exp.buf), (list(exp.buf), list(log))
    assert (int(b[18]) | int(b[19]) << 8, int(b[202]) | int(b[203]) << 8) == (p1.hp, p2.hp)
