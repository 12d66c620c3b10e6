"""ctypes bindings to the ORACLE (oracle/liboracle.so) and to the test-only host build of the CUDA
engine source (tests/host_shim). Test infrastructure only."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MASK64 = (1 << 64) - 1


def splitmix64(x):
    z = (x + 0x9E3779B97F4A7C15) & MASK64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
    return z ^ (z >> 31)


def sample_seed(i):
    return ((i + 1) * 0x9E3779B97F4A7C15) & MASK64


def pick_word(seed, step):
    return splitmix64(seed ^ ((0xD1B54A32D192ED03 * step) & MASK64))


class Oracle:
    def __init__(self):
        path = os.path.join(ROOT, "oracle", "liboracle.so")
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
        self.lib = lib = ctypes.CDLL(path)
        lib.oracle_gen1_update.restype = ctypes.c_uint8
        lib.oracle_gen1_choices.restype = ctypes.c_uint8
        lib.oracle_oak_step.restype = ctypes.c_uint8
        lib.oracle_oak_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint64,
                                        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.oracle_oak_sample_battle.argtypes = [ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.oracle_oak_rollout.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                                           ctypes.c_uint32, ctypes.c_void_p]
        lib.oracle_oak_battle_init.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_uint64, ctypes.c_void_p]
        lib.oracle_oak_randomize_hidden.argtypes = [ctypes.c_void_p, ctypes.c_void_p]

    @staticmethod
    def _p(a):
        return a.ctypes.data_as(ctypes.c_void_p)

    def sample_battle(self, i):
        b = np.zeros(384, np.uint8)
        d = np.zeros(8, np.uint8)
        r = np.zeros(1, np.uint8)
        self.lib.oracle_oak_sample_battle(i, self._p(b), self._p(d), self._p(r))
        return b, d, int(r[0])

    def battle_init(self, p1_species, p1_moves, p2_species, p2_moves, seed):
        b = np.zeros(384, np.uint8)
        a = [np.ascontiguousarray(x, np.uint8) for x in (p1_species, p1_moves, p2_species, p2_moves)]
        self.lib.oracle_oak_battle_init(*[self._p(x) for x in a], seed, self._p(b))
        return b

    def choices(self, battle, player, request):
        out = np.zeros(16, np.uint8)
        n = self.lib.oracle_gen1_choices(self._p(battle), player, request, self._p(out))
        return [int(x) for x in out[:n]]

    def update(self, battle, c1, c2, durations, overrides=None):
        """options_set(null[, calc]) + update. battle/durations updated in place. Returns (result, actions[16])."""
        opt = np.zeros(40, np.uint8)
        opt[16:24] = durations
        if overrides is not None:
            opt[24:40] = overrides
        r = self.lib.oracle_gen1_update(self._p(battle), c1, c2, self._p(opt))
        durations[:] = opt[16:24]
        return int(r), opt[0:16].copy()

    def update_log(self, battle, c1, c2, durations, overrides=None, cap=180):
        """update with the binary protocol log -> (result, actions[16], log bytes); result 0x0C if the log overflowed cap."""
        opt = np.zeros(40, np.uint8)
        opt[16:24] = durations
        if overrides is not None:
            opt[24:40] = overrides
        log = np.zeros(max(cap, 1), np.uint8)
        pos = ctypes.c_size_t(0)
        self.lib.oracle_gen1_update_log.restype = ctypes.c_uint8
        r = self.lib.oracle_gen1_update_log(self._p(battle), ctypes.c_uint8(c1), ctypes.c_uint8(c2), self._p(opt), self._p(log),
                                            ctypes.c_size_t(cap), ctypes.byref(pos))
        durations[:] = opt[16:24]
        return int(r), opt[0:16].copy(), log[:pos.value].tobytes()

    def step(self, battle, durations, result, pick):
        act = np.zeros(16, np.uint8)
        c = np.zeros(2, np.uint8)
        r = self.lib.oracle_oak_step(self._p(battle), self._p(durations), result, pick, self._p(act),
                                     self._p(c[0:1]), self._p(c[1:2]))
        return int(r), act, int(c[0]), int(c[1])

    def rollout(self, battle, durations, result, seed, max_steps=100000):
        r = np.array([result], np.uint8)
        steps = np.zeros(1, np.uint32)
        self.lib.oracle_oak_rollout(self._p(battle), self._p(durations), self._p(r), seed, max_steps, self._p(steps))
        return int(r[0]), int(steps[0])

    def randomize_hidden(self, battle, durations):
        self.lib.oracle_oak_randomize_hidden(self._p(battle), self._p(durations))


class HostShim:
    """The CUDA engine source compiled for the host (tests only)."""

    def __init__(self):
        d = os.path.join(ROOT, "tests", "host_shim")
        so = os.path.join(d, "libengine_host.so")
        src = os.path.join(d, "engine_host_shim.cpp")
        deps = [src] + [os.path.join(ROOT, "oak_b200", "csrc", f) for f in os.listdir(os.path.join(ROOT, "oak_b200", "csrc"))]
        if not os.path.exists(so) or any(os.path.getmtime(x) > os.path.getmtime(so) for x in deps):
            subprocess.check_call(["g++", "-O2", "-fno-strict-aliasing", "-std=c++17", "-fPIC", "-shared", "-o", so, src])
        self.lib = lib = ctypes.CDLL(so)
        lib.shim_update.restype = ctypes.c_uint8
        lib.shim_update.argtypes = [ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint8, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.shim_update_nochance.restype = ctypes.c_uint8
        lib.shim_update_nochance.argtypes = [ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint8]
        lib.shim_choices.restype = ctypes.c_uint8
        lib.shim_choices.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]

    def choices(self, battle, player, request):
        out = np.zeros(16, np.uint8)
        n = self.lib.shim_choices(battle.ctypes.data, player, request, out.ctypes.data)
        return [int(x) for x in out[:n]]

    def update(self, battle, c1, c2, durations, overrides=None):
        act = np.zeros(16, np.uint8)
        ov = None if overrides is None else np.ascontiguousarray(overrides, np.uint8).ctypes.data
        r = self.lib.shim_update(battle.ctypes.data, c1, c2, durations.ctypes.data, act.ctypes.data, ov)
        return int(r), act

    def update_split(self, battle, c1, c2, durations, overrides=None):
        """The SPLIT on-chip layout (constant Pokémon fields left in the record): same contract as update()."""
        act = np.zeros(16, np.uint8)
        ov = None if overrides is None else np.ascontiguousarray(overrides, np.uint8).ctypes.data
        self.lib.shim_update_split.restype = ctypes.c_uint8
        self.lib.shim_update_split.argtypes = self.lib.shim_update.argtypes
        r = self.lib.shim_update_split(battle.ctypes.data, c1, c2, durations.ctypes.data, act.ctypes.data, ov)
        return int(r), act

    def choices_split(self, battle, player, request):
        out = np.zeros(16, np.uint8)
        self.lib.shim_choices_split.restype = ctypes.c_uint8
        self.lib.shim_choices_split.argtypes = self.lib.shim_choices.argtypes
        n = self.lib.shim_choices_split(battle.ctypes.data, player, request, out.ctypes.data)
        return [int(x) for x in out[:n]]

    def update_nochance(self, battle, c1, c2):
        return int(self.lib.shim_update_nochance(battle.ctypes.data, c1, c2))

    def update_log(self, battle, c1, c2, durations, overrides=None, cap=180):
        """LOG build of the engine -> (result, actions, log bytes)."""
        act = np.zeros(16, np.uint8)
        log = np.zeros(max(cap, 1), np.uint8)
        n = ctypes.c_uint32(0)
        ov = None if overrides is None else np.ascontiguousarray(overrides, np.uint8).ctypes.data
        self.lib.shim_update_log.restype = ctypes.c_uint8
        self.lib.shim_update_log.argtypes = [ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint8, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p]
        r = self.lib.shim_update_log(battle.ctypes.data, c1, c2, durations.ctypes.data, act.ctypes.data, ov, log.ctypes.data, cap,
                                     ctypes.byref(n))
        return int(r), act, log[:n.value].tobytes()


def oracle_rollout_batch(orc, battles, durations, results, seeds, max_steps=100000, threads=None):
    """In-place terminal playouts of a whole batch on the host. -> (score2, turns, steps)."""
    n = battles.shape[0]
    threads = threads or os.cpu_count() or 1
    score = np.empty(n, np.uint8)
    turns = np.empty(n, np.uint16)
    steps = np.empty(n, np.uint32)
    f = orc.lib.oracle_oak_rollout_batch
    f.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int] + [ctypes.c_void_p] * 3
    f.restype = None
    f(battles.ctypes.data, durations.ctypes.data, results.ctypes.data, seeds.ctypes.data, n, max_steps, threads,
      score.ctypes.data, turns.ctypes.data, steps.ctypes.data)
    return score, turns, steps


def oracle_sample_batch(orc, first, n, threads=None):
    threads = threads or os.cpu_count() or 1
    b = np.empty((n, 384), np.uint8)
    d = np.empty((n, 8), np.uint8)
    r = np.empty((n,), np.uint8)
    f = orc.lib.oracle_oak_sample_batch
    f.argtypes = [ctypes.c_uint64, ctypes.c_size_t, ctypes.c_int] + [ctypes.c_void_p] * 3
    f.restype = None
    f(first, n, threads, b.ctypes.data, d.ctypes.data, r.ctypes.data)
    return b, d, r


def sample_seeds(first, n):
    i = np.arange(first + 1, first + n + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return i * np.uint64(0x9E3779B97F4A7C15)


def sample_teams(first, n):
    """teams[n,2,6,5] for synthetic battles first..first+n: teams[i & 15] vs teams[(i >> 4) & 15]."""
    import json
    t = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
    i = np.arange(first, first + n)
    return np.stack([t[i & 15], t[(i >> 4) & 15]], axis=1)


class NetOracle:
    """Oracle leaf evaluation (oracle/nn_oracle.hpp) on a parameter blob in the reference's file format."""

    def __init__(self, orc, blob):
        self.lib = orc.lib
        blob = np.ascontiguousarray(blob, np.uint8)
        self.lib.oracle_net_create.restype = ctypes.c_void_p
        self.lib.oracle_net_create.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
        self.h = self.lib.oracle_net_create(blob.ctypes.data, blob.size)
        if not self.h:
            raise ValueError("oracle_net_create: bad parameter blob")
        dims = np.zeros(7, np.uint32)
        self.lib.oracle_net_dims.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        self.lib.oracle_net_dims(self.h, dims.ctypes.data)
        self.phd, self.pod, self.ahd, self.aod, self.hd, self.vhd, self.pohd = (int(x) for x in dims)
        self.side_dim = (1 + self.aod) + 5 * (1 + self.pod)

    def eval_batch(self, battles, durations, results, threads=None):
        n = battles.shape[0]
        v = np.zeros(n, np.float32)
        lg = np.zeros((n, 2, 9), np.float32)
        cnt = np.zeros((n, 2), np.uint8)
        f = self.lib.oracle_net_eval_batch
        f.restype = None
        f.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_size_t, ctypes.c_int] + [ctypes.c_void_p] * 3
        f(self.h, battles.ctypes.data, durations.ctypes.data, results.ctypes.data, n, threads or os.cpu_count() or 1,
          v.ctypes.data, lg.ctypes.data, cnt.ctypes.data)
        return v, lg, cnt

    def eval_batch_cached(self, battles, durations, results, group, threads=None, repeat=1):
        """The reference's memoised inference path (PokemonCache / ActivePokemonCache); positions sorted by `group`
        (the battle they come from: one fill_cache per group)."""
        n = battles.shape[0]
        v = np.zeros(n, np.float32)
        lg = np.zeros((n, 2, 9), np.float32)
        cnt = np.zeros((n, 2), np.uint8)
        group = np.ascontiguousarray(group, np.uint32)
        f = self.lib.oracle_net_eval_cached_batch
        f.restype = None
        f.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_size_t, ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 3
        f(self.h, battles.ctypes.data, durations.ctypes.data, results.ctypes.data, group.ctypes.data, n, threads or os.cpu_count() or 1,
          int(repeat), v.ctypes.data, lg.ctypes.data, cnt.ctypes.data)
        return v, lg, cnt

    def eval_one(self, battle, durations, c1, c2):
        l1 = np.zeros(9, np.float32)
        l2 = np.zeros(9, np.float32)
        emb = np.zeros(2 * self.side_dim, np.float32)
        a1 = np.array(c1, np.uint8)
        a2 = np.array(c2, np.uint8)
        f = self.lib.oracle_net_eval
        f.restype = ctypes.c_float
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 5
        v = f(self.h, battle.ctypes.data, durations.ctypes.data, len(c1), len(c2), a1.ctypes.data, a2.ctypes.data, l1.ctypes.data,
              l2.ctypes.data, emb.ctypes.data)
        return float(v), l1[:len(c1)], l2[:len(c2)], emb


def make_net_blob(rng, phd=128, pod=59, ahd=128, aod=83, hd=64, vhd=32, pohd=64, activation=1, scale=1.0):
    """Random parameters in the reference's file format (affine.h:34-69, network.h:52-70, search.cc:127-147):
    U(-1/sqrt(in), 1/sqrt(in)) like Affine::initialize (affine.h:104-114)."""
    import struct
    side = (1 + aod) + 5 * (1 + pod)
    shapes = [(198, phd), (phd, pod), (427, ahd), (ahd, aod), (2 * side, hd), (hd, hd), (hd, vhd), (vhd, 1),
              (hd, pohd), (pohd, 315), (hd, pohd), (pohd, 315)]
    out = [struct.pack("<Q", activation - 1)]
    for i, o in shapes:
        k = scale / np.sqrt(i)
        out.append(struct.pack("<II", i, o))
        out.append(rng.uniform(-k, k, o).astype("<f4").tobytes())
        out.append(rng.uniform(-k, k, (o, i)).astype("<f4").tobytes())
    return np.frombuffer(b"".join(out), np.uint8).copy()


# ---------------------------------------------------------------------------------------------------------------
# training frames (oracle/frames_oracle.hpp; the reference's own serialiser through oracle/_ref when it is built)
def _frames_fn(lib, name):
    f = getattr(lib, name)
    f.restype = ctypes.c_uint32
    f.argtypes = [ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint32] + [ctypes.c_void_p] * 8
    return f


def frames_write_game(lib, name, battle, result, k, choice, iterations, ev, nv, emp, nash):
    """One game -> bytes. k/choice [n,2] u8, iterations [n] u32, ev/nv [n] f64, emp/nash [n,2,9] f64."""
    n = k.shape[0]
    a = [np.ascontiguousarray(battle, np.uint8), np.ascontiguousarray(k, np.uint8), np.ascontiguousarray(choice, np.uint8),
         np.ascontiguousarray(iterations, np.uint32), np.ascontiguousarray(ev, np.float64), np.ascontiguousarray(nv, np.float64),
         np.ascontiguousarray(emp, np.float64), np.ascontiguousarray(nash, np.float64)]
    out = np.zeros(391 + 83 * n + 16, np.uint8)
    nb = _frames_fn(lib, name)(a[0].ctypes.data, int(result), n, *[x.ctypes.data for x in a[1:]], out.ctypes.data)
    return out[:nb].copy()


def ref_library():
    path = os.path.join(ROOT, "oracle", "_ref", "libref_oak.so")
    return ctypes.CDLL(path) if os.path.exists(path) else None


def random_frame_games(rng, games, max_len):
    """Ragged synthetic games in lockstep layout: dict of [T,G,...] arrays + lengths[G] + battles[G,384] + final results[G]."""
    L = rng.integers(1, max_len + 1, games)
    T = int(L.max())
    k = rng.integers(1, 10, (T, games, 2)).astype(np.uint8)
    choice = rng.integers(0, 256, (T, games, 2)).astype(np.uint8)
    iters = rng.integers(0, 1 << 32, (T, games), dtype=np.uint64).astype(np.uint32)
    ev = rng.random((T, games)).astype(np.float32)
    nv = rng.random((T, games)).astype(np.float32)
    emp = rng.random((T, games, 2, 9)).astype(np.float32)
    nash = rng.random((T, games, 2, 9)).astype(np.float32)
    emp /= emp.sum(-1, keepdims=True)
    nash /= nash.sum(-1, keepdims=True)
    ev[0, :3], nv[0, :3] = 1.0, 0.0  # the ends of the u16 range
    battles = rng.integers(0, 256, (games, 384)).astype(np.uint8)
    final = rng.integers(1, 4, games).astype(np.uint8) | 0x50
    return dict(L=L, T=T, k=k, choice=choice, iters=iters, ev=ev, nv=nv, emp=emp, nash=nash, battles=battles, final=final)


def frames_expected_game(lib, name, G, g):
    n = int(G["L"][g])
    return frames_write_game(lib, name, G["battles"][g], G["final"][g], G["k"][:n, g], G["choice"][:n, g], G["iters"][:n, g],
                             G["ev"][:n, g].astype(np.float64), G["nv"][:n, g].astype(np.float64),
                             G["emp"][:n, g].astype(np.float64), G["nash"][:n, g].astype(np.float64))


def parse_battle_data(stream):
    """Split a .battle.data byte stream into games (the reader of cpp/src/pyoak.cc:43-71 + compressed-frame.h:201-224).
    -> list of dict(battle, result, updates=[dict(m, n, c1, c2, iterations, ev, nv, p1_emp, p1_nash, p2_emp, p2_nash)])."""
    stream = bytes(stream)
    games, at = [], 0
    while at < len(stream):
        nb = int.from_bytes(stream[at:at + 4], "little")
        nu = int.from_bytes(stream[at + 4:at + 6], "little")
        g = dict(battle=np.frombuffer(stream[at + 6:at + 390], np.uint8).copy(), result=stream[at + 390], updates=[])
        p, end = at + 391, at + nb
        while p < end:
            m, n = (stream[p] & 15) + 1, (stream[p] >> 4) + 1
            u = dict(m=m, n=n, c1=stream[p + 1], c2=stream[p + 2], iterations=int.from_bytes(stream[p + 3:p + 7], "little"),
                     ev=int.from_bytes(stream[p + 7:p + 9], "little"), nv=int.from_bytes(stream[p + 9:p + 11], "little"))
            q = p + 11
            for name, cnt in (("p1_emp", m), ("p1_nash", m), ("p2_emp", n), ("p2_nash", n)):
                u[name] = np.frombuffer(stream[q:q + 2 * cnt], np.uint16).copy()
                q += 2 * cnt
            g["updates"].append(u)
            p = q
        assert p == end and len(g["updates"]) == nu, "corrupt .battle.data stream"
        games.append(g)
        at = end
    return games


# ---------------------------------------------------------------------------------------------------------------
# training rows (oracle/train_rows_oracle.hpp; pinned to the reference's EncodedFrames::from_bytes by the golden file)
ROW_FIELDS = ["pokemon", "active", "hp", "choice_indices", "k", "choice", "iterations", "empirical_policies", "nash_policies",
              "empirical_value", "nash_value", "score"]
ROW_SHAPES = dict(pokemon=((2, 6, 198), np.float32), active=((2, 1, 229), np.float32), hp=((2, 6, 1), np.float32),
                  choice_indices=((2, 9), np.int64), k=((2, 1), np.uint8), choice=((2, 1), np.uint8), iterations=((1,), np.uint32),
                  empirical_policies=((2, 9), np.float32), nash_policies=((2, 9), np.float32), empirical_value=((1,), np.float32),
                  nash_value=((1,), np.float32), score=((1,), np.float32))


def play_frame_games(orc, first, games, seed, max_updates=2000):
    """`games` sample battles played to the end on the oracle with playout picks, stored as one .battle.data stream."""
    f = orc.lib.oracle_play_frame_game
    f.restype = ctypes.c_uint32
    f.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_void_p]
    out = np.zeros(391 + 83 * max_updates + 16, np.uint8)
    parts = []
    for g in range(games):
        nb = f(first + g, splitmix64(seed ^ (first + g)), max_updates, out.ctypes.data)
        if nb:
            parts.append(out[:nb].copy())
    return np.concatenate(parts)


def split_games(stream):
    b, at, out = bytes(stream), 0, []
    while at < len(b):
        nb = int.from_bytes(b[at:at + 4], "little")
        out.append(np.frombuffer(b[at:at + nb], np.uint8))
        at += nb
    return out


def count_rows(stream):
    return sum(int.from_bytes(bytes(g[4:6]), "little") for g in split_games(stream))


def empty_rows(n):
    return {name: np.zeros((n,) + shape, dtype) for name, (shape, dtype) in ROW_SHAPES.items()}


def oracle_training_rows(orc, stream, threads=0):
    """-> dict of the twelve arrays for every stored update of every game in the stream (oracle_training_rows)."""
    f = orc.lib.oracle_training_rows
    f.restype = ctypes.c_int64
    f.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int] + [ctypes.c_void_p] * 12
    stream = np.ascontiguousarray(stream, np.uint8)
    n = count_rows(stream)
    R = empty_rows(n)
    got = f(stream.ctypes.data, stream.size, threads, *[R[x].ctypes.data for x in ROW_FIELDS])
    assert got == n, f"oracle_training_rows -> {got}, expected {n}"
    return R


def oracle_battle_frames(orc, stream):
    """-> dict battle[n,384], durations[n,8], result[n,1], legal_choices[n,2,9] for every stored update (oracle_battle_frames)."""
    f = orc.lib.oracle_battle_frames
    f.restype = ctypes.c_int64
    f.argtypes = [ctypes.c_void_p, ctypes.c_uint64] + [ctypes.c_void_p] * 4
    stream = np.ascontiguousarray(stream, np.uint8)
    n = count_rows(stream)
    R = dict(battle=np.zeros((n, 384), np.uint8), durations=np.zeros((n, 8), np.uint8), result=np.zeros((n, 1), np.uint8),
             legal_choices=np.zeros((n, 2, 9), np.uint8))
    got = f(stream.ctypes.data, stream.size, *[R[x].ctypes.data for x in ("battle", "durations", "result", "legal_choices")])
    assert got == n, f"oracle_battle_frames -> {got}, expected {n}"
    return R


# ---------------------------------------------------------------------------------------------------------------
# "discrete" evaluation (oracle/nn_oracle.hpp: eval_discrete; integer main net pinned to the reference's quantized layers)
def oracle_eval_discrete_batch(net, battles, durations, results, threads=None, want_embeddings=False):
    """NetOracle -> (values, logits[n,2,9], counts[n,2][, u8 embeddings[n, 2*side_dim]])"""
    n = battles.shape[0]
    v, lg, cnt = np.zeros(n, np.float32), np.zeros((n, 2, 9), np.float32), np.zeros((n, 2), np.uint8)
    emb = np.zeros((n, 2 * net.side_dim), np.uint8) if want_embeddings else None
    f = net.lib.oracle_net_eval_discrete_batch
    f.restype = None
    f.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_size_t, ctypes.c_int] + [ctypes.c_void_p] * 4
    f(net.h, battles.ctypes.data, durations.ctypes.data, results.ctypes.data, n, threads or os.cpu_count() or 1, v.ctypes.data,
      lg.ctypes.data, cnt.ctypes.data, emb.ctypes.data if want_embeddings else None)
    return (v, lg, cnt, emb) if want_embeddings else (v, lg, cnt)


def oracle_main_discrete(net, q, idx1, idx2):
    """Integer main net alone on a u8 battle embedding -> (value before the sigmoid, l1[len(idx1)], l2[len(idx2)])."""
    f = net.lib.oracle_net_main_discrete
    f.restype = ctypes.c_float
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 4
    q = np.ascontiguousarray(q, np.uint8)
    i1, i2 = np.ascontiguousarray(idx1, np.uint16), np.ascontiguousarray(idx2, np.uint16)
    l1, l2 = np.zeros(9, np.float32), np.zeros(9, np.float32)
    v = f(net.h, q.ctypes.data, len(i1), len(i2), i1.ctypes.data, i2.ctypes.data, l1.ctypes.data, l2.ctypes.data)
    return np.float32(v), l1[:len(i1)].copy(), l2[:len(i2)].copy()


def blob_layers(blob):
    """Parameter image -> 12 (w[out, in], b[out]) pairs in file order (pk0 pk1 ac0 ac1 fc0 fc1 v2 v3 p1a p1b p2a p2b)."""
    blob = np.ascontiguousarray(blob, np.uint8)
    p, layers = 8, []
    for _ in range(12):
        i, o = (int(x) for x in np.frombuffer(blob[p:p + 8].tobytes(), np.uint32))
        p += 8
        b = np.frombuffer(blob[p:p + 4 * o].tobytes(), np.float32).copy()
        p += 4 * o
        w = np.frombuffer(blob[p:p + 4 * o * i].tobytes(), np.float32).copy().reshape(o, i)
        p += 4 * o * i
        layers.append((w, b))
    return layers


class RefQuantizedMain:
    """The reference's own int8 layers (oracle/_ref: ref_quantized_main), default shape only; parameters prepared once."""

    def __init__(self, ref, blob):
        main = blob_layers(blob)[4:]

        def pad(w, b):
            return (np.vstack([w, np.zeros((320 - w.shape[0], w.shape[1]), np.float32)]), np.r_[b, np.zeros(320 - b.size, np.float32)])
        wb = [main[0], main[1], main[2], main[3], main[4], pad(*main[5]), main[6], pad(*main[7])]
        self.ws = [np.ascontiguousarray(x[0], np.float32) for x in wb]
        self.bs = [np.ascontiguousarray(x[1], np.float32) for x in wb]
        self.PW = (ctypes.c_void_p * 8)(*[x.ctypes.data for x in self.ws])
        self.PB = (ctypes.c_void_p * 8)(*[x.ctypes.data for x in self.bs])
        self.f = ref.ref_quantized_main
        self.f.restype = ctypes.c_float
        self.f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 4

    def __call__(self, q, idx1, idx2):
        q = np.ascontiguousarray(q, np.uint8)
        i1, i2 = np.ascontiguousarray(idx1, np.uint16), np.ascontiguousarray(idx2, np.uint16)
        l1, l2 = np.zeros(9, np.float32), np.zeros(9, np.float32)
        v = self.f(self.PW, self.PB, q.ctypes.data, len(i1), len(i2), i1.ctypes.data, i2.ctypes.data, l1.ctypes.data, l2.ctypes.data)
        return np.float32(v), l1[:len(i1)].copy(), l2[:len(i2)].copy()


def layers_to_blob(header8, layers):
    """Inverse of blob_layers: 8 header bytes + 12 Affine blocks {u32 in, u32 out, f32 bias[out], f32 w[out][in]}."""
    parts = [np.frombuffer(bytes(header8), np.uint8)]
    for w, b in layers:
        parts.append(np.array([w.shape[1], w.shape[0]], np.uint32).view(np.uint8))
        parts.append(np.ascontiguousarray(b, np.float32).view(np.uint8))
        parts.append(np.ascontiguousarray(w, np.float32).reshape(-1).view(np.uint8))
    return np.concatenate(parts)


def dyadic_clamp_blob(rng, scale=3.0):
    """A clamp network whose EMBEDDING nets compute exactly in any summation order: weights and biases are multiples of
    1/256 and the real-valued input columns (stats, boosts, volatile fields) have zero weight, so every float sum is a sum
    of exactly representable dyadic numbers. Used to show that the discrete path is bit-exact once its u8 inputs are."""
    blob = make_net_blob(rng, activation=2, scale=scale)
    layers = blob_layers(blob)
    out = []
    for i, (w, b) in enumerate(layers):
        if i < 4:
            w, b = np.round(w * 256) / 256, np.round(b * 256) / 256
            if i == 0:
                w[:, 0:5] = 0
            if i == 2:
                for lo, hi in ((0, 5), (20, 26), (42, 45), (229, 234)):
                    w[:, lo:hi] = 0
        out.append((w.astype(np.float32), b.astype(np.float32)))
    return layers_to_blob(blob[:8].tobytes(), out)
