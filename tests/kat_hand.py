"""Hand-transcribed reference known-answer tests: the three of gen1/test.zig that tools/transcribe_kat.py refuses because
they build their battles with Zig constructs of their own instead of the `Test(rolls)` harness —
  "switching (order)"  test.zig:133-170   (Battle.init with a live PRNG, @splat teams, raw battle.update)
  "choices"            test.zig:1076-1094 (helpers.Battle.random from PSRNG 0x2718281)
  "RNG overrides"      test.zig:10118-10388 (battles whose PRNG type overrides individual rolls)
("transitions", :10390, exercises calc.transitions, which is off Oak's path.) Test infrastructure; every number below is the
reference's. The "RNG overrides" script is data (UPDATES_FULL) so that the oracle, the engine source compiled for the host
and the GPU step kernel can all be run through it."""
import ctypes
import struct

import numpy as np

from kat_harness import (ARG, MOVE_ID, NAMES, SPECIES_ID, LogExpected, PokemonV, build_battle, calc_stat, L, move, oracle_lib, swtch)

P1, P2 = 0, 1


def ident(p, slot):
    return (p << 3) | slot


# ------------------------------------------------------------------ common/rng.zig:34-59 (the harness's own generator)
class PSRNG:
    def __init__(self, seed):
        self.seed = seed & 0xFFFFFFFFFFFFFFFF

    def next(self):
        self.seed = (0x5D588B656C078965 * self.seed + 0x269EC3) & 0xFFFFFFFFFFFFFFFF
        return self.seed >> 32

    def range(self, lo, hi):
        return self.next() * (hi - lo) // 0x100000000 + lo

    def chance(self, n, d):
        return self.range(0, d) < n

    def new_seed(self):
        a, b, c, d = (self.range(0, 0x10000) for _ in range(4))
        return a << 48 | b << 32 | c << 16 | d


def battle_init(seed, p1, p2):
    """helpers.zig:38-49 Battle.init: teams + rng seeded with PSRNG(seed).newSeed() (showdown branch of prng(), :74-79)."""
    buf = build_battle(p1, p2)
    struct.pack_into("<Q", buf, 376, PSRNG(seed).new_seed())
    return np.frombuffer(bytes(buf), np.uint8).copy()


def random_pokemon(rand, buf, off):
    """helpers.zig:194-247 Pokemon.random with Options{cleric = true, block = true, durations = false} (showdown defaults)."""
    species = rand.range(1, 151 + 1)
    sd = NAMES["species_data"][species - 1]
    lvl = rand.range(1, 100) if rand.chance(1, 20) else 100
    dv = [rand.range(1, 16) if rand.chance(1, 5) else 15 for _ in range(4)]  # atk def spe spc (data.zig:605-612)
    dvs = dict(atk=dv[0], spe=dv[2], spc=dv[3])
    dvs["def"] = dv[1]
    dvs["hp"] = (dv[0] & 1) << 3 | (dv[1] & 1) << 2 | (dv[2] & 1) << 1 | (dv[3] & 1)
    pk = PokemonV(buf, off)
    for i, k in enumerate(["hp", "atk", "def", "spe", "spc"]):
        exp = rand.range(0, 0xFFFF + 1) if rand.chance(1, 20) else 0xFFFF
        setattr(pk.stats, k, calc_stat(k == "hp", sd["stats"][i], dvs[k], exp, lvl))
    n = rand.range(1, 4) if rand.chance(1, 100) else 4
    blocked = {MOVE_ID[m] for m in ("Mimic", "Metronome", "MirrorMove", "Transform")}
    chosen = []
    for i in range(n):
        while True:
            m = rand.range(1, 164 + 1)  # Move.size - 1: everything but Struggle
            if m in blocked or m in chosen:
                continue
            break
        chosen.append(m)
        pp = NAMES["pp"][m - 1]
        pk.moves[i].id = m
        pk.moves[i].pp = pp + 3 * min(pp // 5, 7)  # cleric: three PP Ups, full PP
    pk.species = species
    buf[off + 22] = sd["types"]
    pk.level = lvl
    pk.hp = pk.stats.hp
    pk.status = 0


def battle_random(rand):
    """helpers.zig:64-71 Battle.random: field initialisers run in source order: rng, then side 1, then side 2."""
    buf = bytearray(384)
    struct.pack_into("<Q", buf, 376, rand.new_seed())
    for s in range(2):
        n = rand.range(1, 6) if rand.chance(1, 100) else 6
        for i in range(n):
            random_pokemon(rand, buf, s * 184 + i * 24)
            buf[s * 184 + 176 + i] = i + 1
    return np.frombuffer(bytes(buf), np.uint8).copy()


# ------------------------------------------------------------------ oracle calls with libpkmn's default flags
def update_default(battle, c1, c2, opt=None, log_cap=None):
    """-> (result, log bytes or None). opt: the 40-byte Options block (actions 16 | durations 8 | overrides 16)."""
    lib = oracle_lib()
    opt = np.zeros(40, np.uint8) if opt is None else opt
    if log_cap is None:
        lib.oracle_gen1_update_default.restype = ctypes.c_uint8
        return lib.oracle_gen1_update_default(battle.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint8(c1), ctypes.c_uint8(c2),
                                              opt.ctypes.data_as(ctypes.c_void_p)), None
    buf = np.zeros(max(log_cap, 1), np.uint8)
    pos = ctypes.c_size_t(0)
    lib.oracle_gen1_update_default_log.restype = ctypes.c_uint8
    r = lib.oracle_gen1_update_default_log(battle.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint8(c1), ctypes.c_uint8(c2),
                                           opt.ctypes.data_as(ctypes.c_void_p), buf.ctypes.data_as(ctypes.c_void_p),
                                           ctypes.c_size_t(log_cap), ctypes.byref(pos))
    return r, buf[:pos.value].tobytes()


def pokemon_view(battle, p, slot):
    """side.get(slot) of a battle held in a numpy array."""
    buf = bytearray(battle.tobytes())
    idx = buf[p * 184 + 176 + slot - 1] - 1
    return PokemonV(buf, p * 184 + idx * 24)


# ------------------------------------------------------------------ calc override words (chance.zig:160-206)
FIELDS = dict(damage=(0, 8), hit=(8, 2), critical_hit=(10, 2), secondary_chance=(12, 2), speed_tie=(14, 2), confused=(16, 2),
              paralyzed=(18, 2), duration=(20, 4), sleep=(24, 3), confusion=(27, 4), disable=(31, 3), attacking=(34, 2), binding=(36, 2),
              move_slot=(38, 3), multi_hit=(41, 3), pp=(44, 4), psywave=(48, 8), metronome=(56, 8))
FALSE, TRUE = 1, 2  # Optional(bool): None 0, false 1, true 2 (common/optional.zig)


def override_word(**kw):
    w = 0
    for k, v in kw.items():
        s, n = FIELDS[k]
        assert 0 <= v < (1 << n), (k, v)
        w |= v << s
    return w


def overrides16(p1, p2):
    return np.frombuffer(struct.pack("<QQ", override_word(**p1), override_word(**p2)), np.uint8).copy()


# ------------------------------------------------------------------ "RNG overrides", Full (test.zig:10131-10327)
# The test's Random type answers every roll with a constant; the same answers as calc overrides of the roll's Action field:
# speedTie -> P1 (1), criticalHit false, damage 239, hit true, confused false, paralyzed false, secondaryChance true,
# psywave 1 (field holds power + 1), distribution 2 (multi_hit 2), moveSlot 2, and per update the one duration that update
# draws (sleep 1, confusion 2, disable 1, attacking 3) and the Metronome pick (state 0: Psybeam / Sludge, then Psychic / Thrash).
def _base(**extra):
    d = dict(speed_tie=1, critical_hit=FALSE, damage=239, hit=TRUE, confused=FALSE, paralyzed=FALSE, secondary_chance=TRUE,
             psywave=2, multi_hit=2, move_slot=2)
    d.update(extra)
    return d


FULL_TEAMS = ([dict(species=L.Chansey, moves=[L.Spore, L.Metronome, L.ThunderShock, L.Disable])],
              [dict(species=L.Chansey, moves=[L.Metronome, L.Psywave, L.Barrage])])
M = MOVE_ID
SLP1, PSN, PAR = 1, 1 << 3, 1 << 6


def full_script():
    """-> list of (c1, c2, overrides16, hp deltas (p1, p2) list applied in log order, log builder(expected, p1, p2))."""
    def u1(e, p1, p2):
        e.move([ident(P1, 1), M["Spore"], ident(P2, 1)])
        p2.status = SLP1
        e.status([ident(P2, 1), p2.status, L.From, M["Spore"]])
        e.curestatus([ident(P2, 1), p2.status, L.Message])
        p2.status = 0
        e.turn([2])

    def u2(e, p1, p2):
        e.move([ident(P1, 1), M["Metronome"], ident(P1, 1)])
        e.move([ident(P1, 1), M["Psybeam"], ident(P2, 1), M["Metronome"]])
        p2.hp -= 52
        e.damage([ident(P2, 1), p2, L.None_])
        e.start([ident(P2, 1), L.Confusion])
        e.activate([ident(P2, 1), L.Confusion])
        e.move([ident(P2, 1), M["Metronome"], ident(P2, 1)])
        e.move([ident(P2, 1), M["Sludge"], ident(P1, 1), M["Metronome"]])
        p1.hp -= 52
        e.damage([ident(P1, 1), p1, L.None_])
        p1.status = PSN
        e.status([ident(P1, 1), p1.status, L.None_])
        e.turn([3])

    def u3(e, p1, p2):
        e.move([ident(P1, 1), M["ThunderShock"], ident(P2, 1)])
        p2.hp -= 32
        e.damage([ident(P2, 1), p2, L.None_])
        p2.status = PAR
        e.status([ident(P2, 1), p2.status, L.None_])
        p1.hp -= 43
        e.damage([ident(P1, 1), p1, L.Poison])
        e.end([ident(P2, 1), L.Confusion])
        e.move([ident(P2, 1), M["Psywave"], ident(P1, 1)])
        p1.hp -= 1
        e.damage([ident(P1, 1), p1, L.None_])
        e.turn([4])

    def u4(e, p1, p2):
        e.move([ident(P1, 1), M["Disable"], ident(P2, 1)])
        e.start([ident(P2, 1), L.Disable, M["Psywave"]])
        p1.hp -= 43
        e.damage([ident(P1, 1), p1, L.Poison])
        e.end([ident(P2, 1), L.Disable])
        e.move([ident(P2, 1), M["Barrage"], ident(P1, 1)])
        p1.hp -= 19
        e.damage([ident(P1, 1), p1, L.None_])
        p1.hp -= 19
        e.damage([ident(P1, 1), p1, L.None_])
        e.hitcount([ident(P1, 1), 2])
        e.turn([5])

    def u5(e, p1, p2):
        e.move([ident(P1, 1), M["Metronome"], ident(P1, 1)])
        e.move([ident(P1, 1), M["Psychic"], ident(P2, 1), M["Metronome"]])
        p2.hp -= 72
        e.damage([ident(P2, 1), p2, L.None_])
        e.boost([ident(P2, 1), L.SpecialAttack, -1])
        e.boost([ident(P2, 1), L.SpecialDefense, -1])
        p1.hp -= 43
        e.damage([ident(P1, 1), p1, L.Poison])
        e.move([ident(P2, 1), M["Metronome"], ident(P2, 1)])
        e.move([ident(P2, 1), M["Thrash"], ident(P1, 1), M["Metronome"]])
        p1.hp -= 107
        e.damage([ident(P1, 1), p1, L.None_])
        e.turn([6])

    return [
        (move(1), move(1), overrides16(_base(duration=1), _base()), u1),
        (move(2), move(1), overrides16(_base(duration=2, metronome=M["Psybeam"]), _base(metronome=M["Sludge"])), u2),
        (move(3), move(2), overrides16(_base(), _base()), u3),
        (move(4), move(3), overrides16(_base(duration=1), _base()), u4),
        (move(2), move(1), overrides16(_base(metronome=M["Psychic"]), _base(metronome=M["Thrash"], duration=3)), u5),
    ]


def full_start():
    """The battle after update(.{}, .{}) sent the leads out, as a numpy record; rng = 0 (the test's Random has no state)."""
    buf = build_battle(*FULL_TEAMS)
    return np.frombuffer(bytes(buf), np.uint8).copy()


def run_full(update_log):
    """update_log(battle, c1, c2, durations, overrides) -> (result, actions, log bytes), battle / durations updated in place.
    Plays the script and checks result, log bytes and final HP against the reference's expectations."""
    b = full_start()
    d = np.zeros(8, np.uint8)
    r, _, log = update_log(b, 0, 0, d, None)
    assert r == 0x50 and log[0] == ARG["Switch"]
    shadow = bytearray(b.tobytes())  # the test's `copy`: expectations are applied to it
    p1, p2 = PokemonV(shadow, 0), PokemonV(shadow, 184)
    for k, (c1, c2, ov, build) in enumerate(full_script()):
        exp = LogExpected()
        build(exp, p1, p2)
        r, _, log = update_log(b, c1, c2, d, ov)
        assert r == 0x50, (k, r)
        assert log == bytes(exp.buf), (k, list(exp.buf), list(log))
    assert int(b[18]) | int(b[19]) << 8 == p1.hp and int(b[184 + 18]) | int(b[184 + 19]) << 8 == p2.hp
    return p1.hp, p2.hp


__all__ = ["PSRNG", "battle_init", "battle_random", "update_default", "pokemon_view", "overrides16", "run_full", "full_script",
           "full_start", "ident", "P1", "P2", "SPECIES_ID", "swtch"]
