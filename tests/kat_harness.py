"""Python mirror of the reference's Zig test harness (gen1/test.zig:10415-10554, helpers.zig:40-200),
driving the CPU oracle through ctypes.

The known-answer cases in tests/golden/gen1_kat_cases.py are machine-transcribed from the reference's
gen1/test.zig by tools/transcribe_kat.py; they are written against the API below, which keeps the
reference's names (Test(rolls).init(p1, p2), t.update(move(1), move(1)), t.expected.p1.get(1).hp,
t.verify(), ...). Checked after every update, as test.zig:10535-10553 does: the protocol-log bytes the
update wrote against the bytes the case's `t.log.expected.*` calls wrote (the expected writer below is a
separate Python statement of common/protocol.zig:337-824 over the ordinals tools/gen_tables.py read from
that file), HP of all twelve Pokémon, explicit state expectations, results, legal-choice lists and
"scripted roll list fully consumed".
"""
import ctypes
import json
import os
import struct

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = json.load(open(os.path.join(ROOT, "tests", "golden", "gen1_names.json")))
MOVE_ID = {n: i for i, n in enumerate(NAMES["moves"])}
SPECIES_ID = {n: i for i, n in enumerate(NAMES["species"])}
TYPE_ID = {n: i for i, n in enumerate(NAMES["types"])}
MOVE_ID["SKIP_TURN"] = 255

PROTOCOL = {k: {n: i for i, n in enumerate(v)} for k, v in NAMES["protocol"].items()}
ARG = PROTOCOL["ArgType"]
LOG_CAP = 1 << 16

_lib = None


def oracle_lib():
    global _lib
    if _lib is None:
        path = os.path.join(ROOT, "oracle", "liboracle.so")
        if not os.path.exists(path):
            import subprocess
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
        _lib = ctypes.CDLL(path)
        _lib.oracle_gen1_update_kat.restype = ctypes.c_uint8
        _lib.oracle_gen1_update_kat_log.restype = ctypes.c_uint8
        _lib.oracle_gen1_update.restype = ctypes.c_uint8
        _lib.oracle_gen1_choices.restype = ctypes.c_uint8
    return _lib


class SkipCase(Exception):
    pass


class KatFailure(AssertionError):
    pass


# ---------------------------------------------------------------------------- enum literals
class Lit:
    """A Zig enum literal (`.Foo`) whose type is only known from context."""

    def __init__(self, name):
        self.name = name

    def candidates(self):
        out = set()
        for table in (MOVE_ID, SPECIES_ID, TYPE_ID, MISC):
            if self.name in table:
                out.add(table[self.name])
        return out

    def __eq__(self, other):
        if isinstance(other, Lit):
            return self.name == other.name
        return int(other) in self.candidates()

    def __ne__(self, other):
        return not self.__eq__(other)

    def __hash__(self):
        return hash(self.name)

    def __repr__(self):
        return "." + self.name


MISC = {"P1": 0, "P2": 1, "Pass": 0, "Move": 1, "Switch": 2, "None": 0,
        "SLP": 2, "PSN": 3, "BRN": 4, "FRZ": 5, "PAR": 6, "EXT": 7,
        "false": 1, "true": 2, "started": 1, "continuing": 2, "ended": 3, "overwritten": 4}


class _L:
    def __getattr__(self, name):
        return Lit(name.rstrip("_"))


L = _L()


def lit_int(v, table):
    if isinstance(v, Lit):
        return table[v.name]
    return int(v)


class _Enum:
    def __init__(self, table):
        self._t = table

    def __getattr__(self, name):
        return self._t[name.rstrip("_")]


Move = _Enum(MOVE_ID)
Species = _Enum(SPECIES_ID)
Type = _Enum(TYPE_ID)


class _SpeciesNS(_Enum):
    def chance(self, lit):
        return NAMES["species_data"][lit_int(lit, SPECIES_ID) - 1]["chance"]


Species = _SpeciesNS(SPECIES_ID)


class _MoveNS(_Enum):
    def pp(self, lit):
        return NAMES["pp"][lit_int(lit, MOVE_ID) - 1]

    def get(self, lit):
        d = NAMES["move_data"][lit_int(lit, MOVE_ID) - 1]

        class D:
            pass
        o = D()
        o.bp, o.accuracy, o.effect, o.type = d["bp"], d["accuracy"], d["effect"], d["type"]
        return o


Move = _MoveNS(MOVE_ID)


class _Status:
    TOX = 0x88

    @staticmethod
    def init(lit):
        return 1 << lit_int(lit, MISC)

    @staticmethod
    def slp(n):
        return int(n)

    @staticmethod
    def slf(n):
        return 0x80 | int(n)

    @staticmethod
    def is_(num, lit):
        w = lit_int(lit, MISC)
        if w == 2:
            return (num & 7) > 0
        return ((num >> w) & 1) != 0

    @staticmethod
    def any(num):
        return num > 0

    @staticmethod
    def duration(num):
        return num & 7


Status = _Status


class _ResultMeta(type):
    Default = 0x50
    Win = 1
    Lose = 2
    Tie = 3
    Error = 4


class Result(metaclass=_ResultMeta):
    def __new__(cls, type=None, p1=None, p2=None):
        t = {"None": 0, "Win": 1, "Lose": 2, "Tie": 3, "Error": 4}[type.name] if type is not None else 0
        return t | (lit_int(p1, MISC) << 4 if p1 is not None else 0) | (lit_int(p2, MISC) << 6 if p2 is not None else 0)


class Empty:
    """Zig `.{}`: an empty tuple / default-initialised struct depending on context."""

    def __len__(self):
        return 0

    def __iter__(self):
        return iter(())

    def __int__(self):
        return 0

    def __index__(self):
        return 0

    def __eq__(self, o):
        return isinstance(o, Empty) or o == 0


EMPTY = Empty()

MIN = 0
MAX = 0xFFFFFFFF
NOP = MIN
HIT = MIN
CRIT = MIN
MIN_DMG = MIN
MAX_DMG = MAX
P1 = 0
P2 = 1
showdown = True
log = False


class _Player:
    def __init__(self, p):
        self.p = p

    def ident(self, i):
        return (self.p << 3) | i

    def __index__(self):
        return self.p

    def __int__(self):
        return self.p

    def __eq__(self, o):
        return int(o) == self.p if not isinstance(o, Lit) else o.name == ("P1", "P2")[self.p]

    def __hash__(self):
        return self.p


P1 = _Player(0)
P2 = _Player(1)


def ranged(n, d):
    return n * (0x100000000 // d)


def move(n):
    return (int(n) << 2) | 1


def swtch(n):
    return (int(n) << 2) | 2


forced = move(1)


def metronome(m):
    param = lit_int(m, MOVE_ID)
    rng = MOVE_ID["Struggle"] - 2
    mod = 1 if param < MOVE_ID["Metronome"] - 1 else 2
    return ranged((param - mod) + 1, rng) - 1


# ---------------------------------------------------------------------------- state views
class View:
    FIELDS = {}

    def __init__(self, buf, off, root=None):
        object.__setattr__(self, "_b", buf)
        object.__setattr__(self, "_o", off)

    def _rd(self, off, fmt):
        return struct.unpack_from("<" + fmt, self._b, self._o + off)[0]

    def _wr(self, off, fmt, v):
        mask = {"B": 0xFF, "H": 0xFFFF, "I": 0xFFFFFFFF, "Q": (1 << 64) - 1}[fmt]
        struct.pack_into("<" + fmt, self._b, self._o + off, int(v) & mask)

    def __getattr__(self, name):
        name = name.rstrip("_")
        f = type(self).FIELDS.get(name)
        if f is None:
            raise AttributeError(name)
        if isinstance(f[1], str):
            return self._rd(f[0], f[1])
        return f[1](self._b, self._o + f[0])

    def __setattr__(self, name, v):
        name = name.rstrip("_")
        f = type(self).FIELDS.get(name)
        if f is None:
            raise AttributeError(name)
        if isinstance(f[1], str):
            table = {"species": SPECIES_ID, "id": MOVE_ID, "last_selected_move": MOVE_ID, "last_used_move": MOVE_ID}.get(name, MISC)
            self._wr(f[0], f[1], lit_int(v, table))
        elif isinstance(v, View):
            n = v.SIZE
            self._b[self._o + f[0]:self._o + f[0] + n] = v._b[v._o:v._o + n]
        elif isinstance(v, (Empty, dict)):
            size = f[1].SIZE
            self._b[self._o + f[0]:self._o + f[0] + size] = bytes(size)
            if isinstance(v, dict):
                sub = f[1](self._b, self._o + f[0])
                for k, val in v.items():
                    setattr(sub, k, val)
        else:
            raise TypeError(name)

    def __eq__(self, other):
        if isinstance(other, View):
            return bytes(self._b[self._o:self._o + self.SIZE]) == bytes(other._b[other._o:other._o + other.SIZE])
        return NotImplemented


class StatsV(View):
    SIZE = 10
    FIELDS = {"hp": (0, "H"), "atk": (2, "H"), "def": (4, "H"), "spe": (6, "H"), "spc": (8, "H")}


class MoveSlotV(View):
    SIZE = 2
    FIELDS = {"id": (0, "B"), "pp": (1, "B")}


class Array:
    def __init__(self, cls, n):
        self.cls, self.n = cls, n
        self.SIZE = cls.SIZE * n

    def __call__(self, buf, off):
        return ArrayV(self.cls, self.n, buf, off)


class ArrayV:
    def __init__(self, cls, n, buf, off):
        self.cls, self.n, self._b, self._o = cls, n, buf, off
        self.SIZE = cls.SIZE * n

    def __getitem__(self, i):
        if i < 0 or i >= self.n:
            raise IndexError(i)
        return self.cls(self._b, self._o + i * self.cls.SIZE)

    def __len__(self):
        return self.n


class BytesV:
    SIZE = 6

    def __init__(self, buf, off):
        self._b, self._o = buf, off

    def __getitem__(self, i):
        return self._b[self._o + i]

    def __setitem__(self, i, v):
        self._b[self._o + i] = v

    def __eq__(self, o):
        return list(self._b[self._o:self._o + 6]) == list(o)


class PokemonV(View):
    SIZE = 24
    FIELDS = {"stats": (0, StatsV), "moves": (10, Array(MoveSlotV, 4)), "hp": (18, "H"), "status": (20, "B"),
              "species": (21, "B"), "types": (22, "B"), "level": (23, "B")}

    def move(self, i):
        return self.moves[i - 1]


VOL_FLAGS = ["Bide", "Thrashing", "MultiHit", "Flinch", "Charging", "Binding", "Invulnerable", "Confusion", "Mist",
             "FocusEnergy", "Substitute", "Recharging", "Rage", "LeechSeed", "Toxic", "LightScreen", "Reflect", "Transform"]
VOL_FIELDS = {"confusion": (18, 3), "attacks": (21, 3), "state": (24, 16), "substitute": (40, 8), "transform": (48, 4),
              "disable_duration": (52, 4), "disable_move": (56, 3), "toxic": (59, 5)}


class VolatilesV:
    SIZE = 8

    def __init__(self, buf, off):
        object.__setattr__(self, "_b", buf)
        object.__setattr__(self, "_o", off)

    def _get(self):
        return struct.unpack_from("<Q", self._b, self._o)[0]

    def __getattr__(self, name):
        v = self._get()
        if name in VOL_FLAGS:
            return bool((v >> VOL_FLAGS.index(name)) & 1)
        s, n = VOL_FIELDS[name]
        return (v >> s) & ((1 << n) - 1)

    def __setattr__(self, name, val):
        v = self._get()
        if name in VOL_FLAGS:
            s, n = VOL_FLAGS.index(name), 1
        else:
            s, n = VOL_FIELDS[name]
        m = ((1 << n) - 1) << s
        v = (v & ~m) | ((int(val) << s) & m)
        struct.pack_into("<Q", self._b, self._o, v)

    def __eq__(self, o):
        return isinstance(o, VolatilesV) and self._get() == o._get()


BOOST_FIELDS = ["atk", "def", "spe", "spc", "accuracy", "evasion"]


class BoostsV:
    SIZE = 4

    def __init__(self, buf, off):
        object.__setattr__(self, "_b", buf)
        object.__setattr__(self, "_o", off)

    def __getattr__(self, name):
        name = name.rstrip("_")
        v = struct.unpack_from("<I", self._b, self._o)[0]
        x = (v >> (4 * BOOST_FIELDS.index(name))) & 15
        return x - 16 if x >= 8 else x

    def __eq__(self, o):
        if isinstance(o, BoostsV):
            return bytes(self._b[self._o:self._o + 4]) == bytes(o._b[o._o:o._o + 4])
        return NotImplemented

    def __setattr__(self, name, val):
        name = name.rstrip("_")
        v = struct.unpack_from("<I", self._b, self._o)[0]
        s = 4 * BOOST_FIELDS.index(name)
        v = (v & ~(15 << s)) | ((int(val) & 15) << s)
        struct.pack_into("<I", self._b, self._o, v)


class TypesV:
    SIZE = 1

    def __init__(self, buf, off):
        self._b, self._o = buf, off

    @property
    def type1(self):
        return self._b[self._o] & 15

    @property
    def type2(self):
        return self._b[self._o] >> 4

    def __eq__(self, o):
        if isinstance(o, TypesV):
            return self._b[self._o] == o._b[o._o]
        if isinstance(o, dict):
            return self.type1 == lit_int(o["type1"], TYPE_ID) and self.type2 == lit_int(o["type2"], TYPE_ID)
        return self._b[self._o] == int(o)


class ActiveV(View):
    SIZE = 32
    FIELDS = {"stats": (0, StatsV), "species": (10, "B"), "types": (11, TypesV), "boosts": (12, BoostsV),
              "volatiles": (16, VolatilesV), "moves": (24, Array(MoveSlotV, 4))}

    def move(self, i):
        return self.moves[i - 1]


class SideV(View):
    SIZE = 184
    FIELDS = {"pokemon": (0, Array(PokemonV, 6)), "active": (144, ActiveV), "order": (176, BytesV),
              "last_selected_move": (182, "B"), "last_used_move": (183, "B")}

    def get(self, slot):
        return self.pokemon[self.order[slot - 1] - 1]

    def stored(self):
        return self.get(1)


class MoveDetailsV(View):
    SIZE = 2
    FIELDS = {"index": (0, "B"), "counterable": (1, "B")}


class RngV:
    def __init__(self, t):
        self.t = t

    def exhausted(self):
        return self.t.index.value == len(self.t.rolls)


class BattleV(View):
    SIZE = 384
    FIELDS = {"sides": (0, Array(SideV, 2)), "turn": (368, "H"), "last_damage": (370, "H"),
              "last_moves": (372, Array(MoveDetailsV, 2))}

    def __init__(self, buf, off=0, test=None):
        View.__init__(self, buf, off)
        object.__setattr__(self, "_t", test)

    def side(self, p):
        return self.sides[lit_int(p, MISC)]

    def foe(self, p):
        return self.sides[lit_int(p, MISC) ^ 1]

    def lastMove(self, p):
        return self.last_moves[lit_int(p, MISC)]

    def active(self, p):
        p = lit_int(p, MISC)
        return (p << 3) | self.sides[p].order[0]

    @property
    def rng(self):
        return RngV(self._t)

    def choices(self, player, request, out):
        buf = (ctypes.c_uint8 * 16)()
        raw = (ctypes.c_uint8 * 384).from_buffer(self._b)
        n = oracle_lib().oracle_gen1_choices(raw, lit_int(player, MISC), lit_int(request, MISC), buf)
        out[:n] = list(buf[:n])
        return n

    def update(self, c1, c2, options=None):
        return self._t._raw_update(int(c1), int(c2))


# ---------------------------------------------------------------------------- construction (helpers.zig:126-200)
def calc_stat(is_hp, base, dv, exp, level):
    import math
    evs = min(255, int(math.ceil(math.sqrt(exp))))
    core = 2 * (base + dv) + evs // 4
    factor = level + 10 if is_hp else 5
    return core * level // 100 + factor


def build_pokemon(buf, off, p):
    species = lit_int(p["species"], SPECIES_ID)
    sd = NAMES["species_data"][species - 1]
    level = int(p.get("level", 100))
    dvs = {"atk": 15, "def": 15, "spe": 15, "spc": 15}
    if isinstance(p.get("dvs"), dict):
        dvs.update({k.rstrip("_"): int(v) for k, v in p["dvs"].items()})
    dvs["hp"] = (dvs["atk"] & 1) << 3 | (dvs["def"] & 1) << 2 | (dvs["spe"] & 1) << 1 | (dvs["spc"] & 1)
    exp = {"hp": 0xFFFF, "atk": 0xFFFF, "def": 0xFFFF, "spe": 0xFFFF, "spc": 0xFFFF}
    st = p.get("stats")
    if isinstance(st, Empty):
        exp = dict.fromkeys(exp, 0)
    elif isinstance(st, dict):
        exp = dict.fromkeys(exp, 0)
        exp.update({k.rstrip("_"): int(v) for k, v in st.items()})
    pk = PokemonV(buf, off)
    for i, k in enumerate(["hp", "atk", "def", "spe", "spc"]):
        setattr(pk.stats, k, calc_stat(k == "hp", sd["stats"][i], dvs[k], exp[k], level))
    moves = p["moves"]
    if not isinstance(moves, (list, tuple)):
        moves = [moves]
    for j, m in enumerate(moves):
        mid = lit_int(m, MOVE_ID)
        pk.moves[j].id = mid
        pk.moves[j].pp = min(NAMES["pp"][mid - 1] // 5 * 8, 61)
    pk.hp = int(p["hp"]) if p.get("hp") is not None else pk.stats.hp
    pk.status = int(p.get("status", 0))
    pk.species = species
    buf[off + 22] = sd["types"]
    pk.level = level


def build_battle(p1, p2):
    buf = bytearray(384)
    for s, team in enumerate((p1, p2)):
        if isinstance(team, dict):
            team = [team]
        for i, p in enumerate(team):
            build_pokemon(buf, s * 184 + i * 24, p)
            buf[s * 184 + 176 + i] = i + 1
    return buf


class _Pair:
    pass


def ID():
    """Zig `ID{}`: no target."""
    return 0


def Types(type1, type2):
    return dict(type1=type1, type2=type2)


class LogExpected:
    """`t.log.expected` of the Zig harness: each method appends what common/protocol.zig's `Log` method of the same
    name writes (:337-824). Enum literals resolve against the argument's own reason enum, as Zig's result-location typing
    does."""

    def __init__(self):
        self.buf = bytearray()

    @staticmethod
    def _r(v, enum):
        if isinstance(v, Lit):
            return PROTOCOL[enum][v.name]
        return int(v)

    @staticmethod
    def _mv(v):
        return lit_int(v, MOVE_ID)

    def _pk(self, pk):
        self.buf += struct.pack("<HH", pk.hp, pk.stats.hp)

    def move(self, a):
        self.buf += bytes([ARG["Move"], int(a[0]), self._mv(a[1]), int(a[2])])
        if len(a) < 4 or self._mv(a[3]) == 0:
            self.buf.append(0)
        else:
            self.buf += bytes([1, self._mv(a[3])])

    def switched(self, a):
        pk = a[1]
        self.buf += bytes([ARG["Switch"], int(a[0]), pk.species, pk.level])
        self._pk(pk)
        self.buf.append(pk.status)

    def cant(self, a):
        self.buf += bytes([ARG["Cant"], int(a[0]), self._r(a[1], "Cant")])
        if len(a) == 3:
            self.buf.append(self._mv(a[2]))

    def faint(self, a):
        self.buf += bytes([ARG["Faint"], int(a[0])])
        if a[1]:
            self.buf.append(0)

    def turn(self, a):
        self.buf += bytes([ARG["Turn"]]) + struct.pack("<H", int(a[0])) + b"\0"

    def win(self, a):
        self.buf += bytes([ARG["Win"], lit_int(a[0], MISC), 0])

    def tie(self, a):
        self.buf += bytes([ARG["Tie"], 0])

    def _hp(self, op, a, enum):
        self.buf += bytes([ARG[op], int(a[0])])
        self._pk(a[1])
        self.buf += bytes([a[1].status, self._r(a[2], enum)])
        if len(a) == 4:
            self.buf.append(int(a[3]))

    def damage(self, a):
        self._hp("Damage", a, "Damage")

    def heal(self, a):
        self._hp("Heal", a, "Heal")

    def status(self, a):
        self.buf += bytes([ARG["Status"], int(a[0]), int(a[1]), self._r(a[2], "Status")])
        if len(a) == 4:
            self.buf.append(self._mv(a[3]))

    def curestatus(self, a):
        self.buf += bytes([ARG["CureStatus"], int(a[0]), int(a[1]), self._r(a[2], "CureStatus")])

    def boost(self, a):
        self.buf += bytes([ARG["Boost"], int(a[0]), self._r(a[1], "Boost"), int(a[2]) + 6])

    def clearallboost(self, a):
        self.buf.append(ARG["ClearAllBoost"])

    def fail(self, a):
        self.buf += bytes([ARG["Fail"], int(a[0]), self._r(a[1], "Fail")])

    def miss(self, a):
        self.buf += bytes([ARG["Miss"], int(a[0])])

    def hitcount(self, a):
        self.buf += bytes([ARG["HitCount"], int(a[0]), int(a[1])])

    def prepare(self, a):
        self.buf += bytes([ARG["Prepare"], int(a[0]), self._mv(a[1])])

    def mustrecharge(self, a):
        self.buf += bytes([ARG["MustRecharge"], int(a[0])])

    def activate(self, a):
        self.buf += bytes([ARG["Activate"], int(a[0]), self._r(a[1], "Activate")])

    def fieldactivate(self, a):
        self.buf.append(ARG["FieldActivate"])

    def start(self, a):
        self.buf += bytes([ARG["Start"], int(a[0]), self._r(a[1], "Start")])
        if len(a) == 3:
            self.buf.append(self._mv(a[2]))
        elif len(a) == 4:
            t = a[2]
            types = t._b[t._o] if isinstance(t, TypesV) else (lit_int(t["type1"], TYPE_ID) | lit_int(t["type2"], TYPE_ID) << 4)
            self.buf += bytes([types, int(a[3])])

    def end(self, a):
        self.buf += bytes([ARG["End"], int(a[0]), self._r(a[1], "End")])

    def ohko(self, a):
        self.buf.append(ARG["OHKO"])

    def crit(self, a):
        self.buf += bytes([ARG["Crit"], int(a[0])])

    def supereffective(self, a):
        self.buf += bytes([ARG["SuperEffective"], int(a[0])])

    def resisted(self, a):
        self.buf += bytes([ARG["Resisted"], int(a[0])])

    def immune(self, a):
        self.buf += bytes([ARG["Immune"], int(a[0]), self._r(a[1], "Immune")])

    def transform(self, a):
        self.buf += bytes([ARG["Transform"], int(a[0]), int(a[1])])

    def laststill(self, a):
        self.buf.append(ARG["LastStill"])

    def lastmiss(self, a):
        self.buf.append(ARG["LastMiss"])


def format_log(b):
    """Best-effort split of a log stream into messages, for failure output."""
    inv = {v: k for k, v in ARG.items()}
    return " ".join(f"{inv.get(x, '')}:{x}" if x in inv and 0 < x < 31 else str(x) for x in b)


class OptionsV:
    """options.chance / options.calc as seen by the few tests that poke them."""

    def __init__(self, t):
        self.t = t


class TestCase:
    def __init__(self, rolls, p1, p2):
        rolls = list(rolls)
        self.rolls = [int(r) & 0xFFFFFFFF for r in rolls]
        self._rolls_c = (ctypes.c_uint32 * max(1, len(self.rolls)))(*self.rolls)
        self.index = ctypes.c_size_t(0)
        self.overrun = ctypes.c_int(0)
        init = build_battle(p1, p2)
        self.battle = _Pair()
        self.battle.expected = BattleV(bytearray(init), 0, self)
        self.battle.actual = BattleV(bytearray(init), 0, self)
        self.expected = _Pair()
        self.actual = _Pair()
        self.expected.p1 = self.battle.expected.side(0)
        self.expected.p2 = self.battle.expected.side(1)
        self.actual.p1 = self.battle.actual.side(0)
        self.actual.p2 = self.battle.actual.side(1)
        # Options: actions[2] u64, durations[2] u32, overrides[2] u64
        self.opt = (ctypes.c_uint8 * 40)()
        self.options = OptionsV(self)
        self.updates = 0
        self.log = _Pair()
        self.log.expected = LogExpected()
        self.log_actual = (ctypes.c_uint8 * LOG_CAP)()
        self.log_pos = ctypes.c_size_t(0)
        self.log_offset = 0

    def deinit(self):
        pass

    def _raw_update(self, c1, c2, fresh_options=False):
        lib = oracle_lib()
        raw = (ctypes.c_uint8 * 384).from_buffer(self.battle.actual._b)
        opt = (ctypes.c_uint8 * 40)() if fresh_options else self.opt
        if not fresh_options:
            # calc.update: options.chance.reset() keeps durations (calc.zig:425-426)
            for i in range(16):
                opt[i] = 0
        if fresh_options:  # Test.start() logs into its own scratch buffer (test.zig:10474-10492)
            scratch = (ctypes.c_uint8 * 64)()
            pos = ctypes.c_size_t(0)
            r = lib.oracle_gen1_update_kat_log(raw, c1, c2, opt, self._rolls_c, len(self.rolls), ctypes.byref(self.index),
                                               ctypes.byref(self.overrun), scratch, 64, ctypes.byref(pos))
            exp = LogExpected()
            exp.switched([P1.ident(1), self.actual.p1.get(1)])
            exp.switched([P2.ident(1), self.actual.p2.get(1)])
            exp.turn([1])
            if bytes(scratch[:pos.value]) != bytes(exp.buf):
                raise KatFailure(f"start log: expected {format_log(exp.buf)} actual {format_log(scratch[:pos.value])}")
        else:
            r = lib.oracle_gen1_update_kat_log(raw, c1, c2, opt, self._rolls_c, len(self.rolls), ctypes.byref(self.index),
                                               ctypes.byref(self.overrun), self.log_actual, LOG_CAP, ctypes.byref(self.log_pos))
        if self.overrun.value:
            raise KatFailure("Insufficient number of rolls provided")
        self.updates += 1
        return r

    def start(self):
        r = self._raw_update(0, 0, fresh_options=True)
        if r != Result.Default:
            raise KatFailure(f"start: result {r:#x}")

    def update(self, c1, c2):
        if self.battle.actual.turn == 0:
            self.start()
        r = self._raw_update(int(c1), int(c2))
        self.validate()
        return r

    def validate(self):
        exp = bytes(self.log.expected.buf)
        act = bytes(self.log_actual[:self.log_pos.value])
        if exp != act:
            o = self.log_offset
            raise KatFailure(f"protocol log after update #{self.updates}:\n  expected {format_log(exp[o:])}\n  actual   {format_log(act[o:])}")
        self.log_offset = len(exp)
        for s, (e, a) in enumerate(((self.expected.p1, self.actual.p1), (self.expected.p2, self.actual.p2))):
            for i in range(6):
                if e.pokemon[i].hp != a.pokemon[i].hp:
                    raise KatFailure(f"P{s + 1} pokemon[{i}] hp expected {e.pokemon[i].hp} actual {a.pokemon[i].hp} "
                                     f"(after update #{self.updates})")

    def verify(self):
        self.validate()
        if self.index.value != len(self.rolls):
            raise KatFailure(f"rolls not exhausted: used {self.index.value} of {len(self.rolls)}")

    def expectProbability(self, p, q):
        pass

    @property
    def durations(self):
        return struct.unpack_from("<II", bytes(self.opt), 16)

    @property
    def actions(self):
        return struct.unpack_from("<QQ", bytes(self.opt), 0)


class Test:
    def __init__(self, rolls):
        self.rolls = rolls

    def init(self, p1, p2):
        return TestCase(self.rolls, p1, p2)


def expectEqual(a, b):
    if isinstance(a, Lit):
        ok = a == b
    elif isinstance(b, Lit):
        ok = b == a
    else:
        ok = a == b
    if not ok:
        raise KatFailure(f"expected {a!r}, found {b!r}")


def expect(v):
    if not v:
        raise KatFailure("expect(false)")


def expectStatsEqualExceptHp(a, b):
    for f in ("atk", "def_", "spe", "spc"):
        if getattr(a, f) != getattr(b, f):
            raise KatFailure(f"stats.{f}: expected {getattr(a, f)}, found {getattr(b, f)}")


def expectEqualSlices(_ty, a, b):
    if [int(x) for x in a] != [int(x) for x in b]:
        raise KatFailure(f"expected {[int(x) for x in a]}, found {[int(x) for x in b]}")


def namespace():
    g = dict(Test=Test, move=move, swtch=swtch, forced=forced, Result=Result, Status=Status, Move=Move,
             Species=Species, ranged=ranged, metronome=metronome, MIN=MIN, MAX=MAX, NOP=NOP, HIT=HIT, CRIT=CRIT,
             MIN_DMG=MIN_DMG, MAX_DMG=MAX_DMG, P1=P1, P2=P2, L=L, ID=ID, Types=Types, EMPTY=EMPTY, expectEqual=expectEqual, expect=expect,
             expectEqualSlices=expectEqualSlices, expectStatsEqualExceptHp=expectStatsEqualExceptHp, showdown=True, SkipCase=SkipCase, Choice=None,
             choices=[0] * 16)
    return g

NULL = None
CHOICES_SIZE = 16
choices = [0] * 16
Choice = None
EXP = 0xFFFF
def repeat(xs, n):
    return list(xs) * n


class Battle:
    @staticmethod
    def fixed(rolls, p1, p2):
        return TestCase(rolls, p1, p2).battle.actual

options = None


class _ZeroVolatiles:
    def __eq__(self, o):
        return o._get() == 0


class _DataNS:
    @staticmethod
    def Volatiles():
        return _ZeroVolatiles()

    @staticmethod
    def Boosts():
        class Z:
            def __eq__(self, o):
                return struct.unpack_from("<I", o._b, o._o)[0] == 0
        return Z()


data = _DataNS
