"""The battle-path names of the reference's `oak` Python module (pyoak, reference cpp/src/pyoak.cc:428-716), served by
liboakb200 on the GPU, so that the reference's lab / learner scripts (src/oak/lab.py:42-70, src/oak/battle.py:320-376)
find what they call:

    read_battle_data, SampleIndexer, sample, EncodedBattleFrames, BattleFrames, OutputBuffer, cpp_inference,
    solve_matrix, Input, update, Agent, Heap, Output, search, the network / encoding dimensions and the constant tables
    (species_move_list, species_move_table, species_names, move_names, *_dim_labels, build_* dims; pyoak.cc:586-636),
    generated from the reference's headers by tools/gen_oak_constants.py into oak_constants.py.

Differences a caller sees, all deliberate:
  * EncodedBattleFrames / BattleFrames hold torch tensors in DEVICE memory (same names, shapes, dtypes) — the learner
    moves them there anyway (battle.py:327) — and `from_bytes` accepts any number of games;
  * `threads` of `sample` is accepted and ignored: the replay + encode of the drawn games is two kernel launches;
  * `cpp_inference` evaluates every frame in one tensor-core launch; `budget` only costs time in the reference (it
    reports the root's network value and priors, pyoak.cc:351-356), so it is parsed and otherwise unused;
  * `search` runs the library's batched tree search on one root: bandits "ucb-<c>", "ucb1-<c>", "pucb-<c>",
    "exp3-<gamma>[-<alpha>]", "pexp3-<gamma>[-<alpha>]" (cpp/src/search.cc:236-300) with MonteCarlo or network leaves
    (pucb / pexp3 need a network: their priors are its policy). Matrix-UCB, the transposition table and time budgets are
    not built and raise.
  * the reference's src/oak/torch.py wraps these objects with torch.from_numpy; device tensors want `.host()` first
    (EncodedBattleFrames.host() -> the same names as numpy arrays), or torch.as_tensor in place of torch.from_numpy.
Not here (host text utilities and the team-builder, off the battle path): parse_battle, battle_string, format;
BuildTrajectories / read_build_trajectories exist as names only and raise.
There is no CPU path: everything below that touches a battle needs a CUDA device and the built library.
"""
import os

import numpy as np

from .engine import BattleEngine, OakB200Error, rows_index_host  # noqa: F401
from .search import TreeSearch, solve_matrix  # noqa: F401
from .training_rows import BattleFrames, EncodedBattleFrames  # noqa: F401

# Encode::Battle::{Pokemon, ActivePokemon}::n_dim, NN::Battle::Default::*, Encode::Battle::Policy::n_dim, the team-builder
# tables and the name tables (pyoak.cc:586-636). active_in_dim is the ACTIVE NET's input (229 active features + the 198 of
# the Pokémon itself = 427); the `active` tensor of EncodedBattleFrames holds the first 229 (encoded-frames.h:30).
from .oak_constants import *  # noqa: F401,F403,E402
from .oak_constants import active_out_dim, pokemon_out_dim, policy_out_dim  # noqa: E402


def read_battle_data(path):
    """-> [(bytes of one stored game, its frame count), ...] (pyoak.cc:43-71). Raises on a file that does not split into
    whole games."""
    try:
        data = np.fromfile(path, np.uint8)
    except OSError as e:
        raise RuntimeError(f"read_battle_data: Failed to open file: {path}") from e
    try:
        off, base, rows = rows_index_host(data)
    except Exception as e:
        raise RuntimeError("read_battle_data: bad battle buffer read") from e
    out = []
    for o in off:
        o = int(o)
        nb = int.from_bytes(data[o:o + 4].tobytes(), "little")
        out.append((data[o:o + nb].tobytes(), int.from_bytes(data[o + 4:o + 6].tobytes(), "little")))
    return out


class SampleIndexer:
    """path -> [(byte offset of a game in the file, its frame count), ...], cached (pyoak.cc:73-109)."""

    def __init__(self):
        self.data = {}

    def size(self):
        return len(self.data)

    def get(self, path):
        if path not in self.data:
            data = np.fromfile(path, np.uint8)
            off, _, _ = rows_index_host(data)
            self.data[path] = [(int(o), int.from_bytes(data[int(o) + 4:int(o) + 6].tobytes(), "little")) for o in off]
        return self.data[path]

    def prune(self, paths):
        keep = set(paths)
        for p in list(self.data):
            if p not in keep:
                del self.data[p]


def sample(encoded_frames, indexer, threads=1, max_battle_length=1 << 30, min_iterations=0, rng=None):
    """Fill `encoded_frames` with one row per drawn game (pyoak.cc:111-260): games are drawn uniformly over all indexed
    files (with replacement), a game longer than max_battle_length or without an update of at least min_iterations search
    iterations is drawn again, and the row is the one built before a uniformly drawn such update. -> rows written
    (encoded_frames.size), or 0 on an error, as the reference."""
    rng = rng or np.random.default_rng()
    paths = list(indexer.data)
    counts = np.array([len(indexer.data[p]) for p in paths], np.int64)
    total = int(counts.sum())
    if total == 0:
        return 0
    starts = np.concatenate([[0], np.cumsum(counts)])
    want = encoded_frames.size
    maps = {}
    games, picks = [], []
    attempts = 0
    try:
        while len(games) < want:
            attempts += 1
            if attempts > 64 * want + 1024:  # nothing in the files qualifies: the reference would spin forever
                return 0
            g = int(rng.integers(0, total))
            pi = int(np.searchsorted(starts, g, side="right") - 1)
            off, n_frames = indexer.data[paths[pi]][g - int(starts[pi])]
            if n_frames > max_battle_length or n_frames == 0:
                continue
            if pi not in maps:
                maps[pi] = np.memmap(paths[pi], np.uint8, mode="r")
            m = maps[pi]
            nb = int.from_bytes(m[off:off + 4].tobytes(), "little")
            if nb > 200000 or nb < 384 or off + nb > m.size:  # pyoak.cc:187-190
                return 0
            game = np.asarray(m[off:off + nb])
            p, valid = 391, []
            for u in range(n_frames):
                if p + 11 > nb:  # a corrupt record: the stored (m, n) walked out of the game
                    raise ValueError(f"sample: update {u} of the game at byte {off} of {paths[pi]} lies outside its record")
                hdr = int(game[p])
                if int.from_bytes(game[p + 3:p + 7].tobytes(), "little") >= min_iterations:
                    valid.append(u)
                p += 11 + 4 * ((hdr & 15) + (hdr >> 4) + 2)
            if p != nb:
                raise ValueError(f"sample: the game at byte {off} of {paths[pi]} does not end on its last update")
            if not valid:
                continue
            games.append(game)
            picks.append(valid[int(rng.integers(0, len(valid)))])
        return encoded_frames.fill(np.concatenate(games), only_update=np.asarray(picks, np.int32))
    except (OSError, ValueError, OakB200Error) as e:
        import sys
        print(e, file=sys.stderr)
        return 0


class OutputBuffer:
    """Py::Battle::Output (cpp/include/py/battle/output.h:13-64): host arrays, as the reference."""

    def __init__(self, size, pokemon_out_dim=pokemon_out_dim, active_out_dim=active_out_dim):
        self.size, self.pokemon_out_dim, self.active_out_dim = int(size), int(pokemon_out_dim), int(active_out_dim)
        self.side_out_dim = (1 + self.active_out_dim) + 5 * (1 + self.pokemon_out_dim)
        n = self.size
        self.pokemon = np.zeros((n, 2, 5, self.pokemon_out_dim), np.float32)
        self.active_pokemon = np.zeros((n, 2, 1, self.active_out_dim), np.float32)
        self.sides = np.zeros((n, 2, 1, self.side_out_dim), np.float32)
        self.value = np.zeros((n, 1), np.float32)
        self.logit = np.zeros((n, 2, policy_out_dim + 1), np.float32)  # last entry -inf: where illegal actions point
        self.policy_logit = np.zeros((n, 2, 9), np.float32)
        self.policy = np.zeros((n, 2, 9), np.float32)
        self.clear()

    def clear(self):
        for a in (self.pokemon, self.active_pokemon, self.sides, self.value, self.logit, self.policy_logit, self.policy):
            a.fill(0.0)
        self.logit[:, :, -1] = -np.inf


class BuildTrajectories:
    """The team-builder's trajectory container (pyoak.cc:697-707): off the battle path (SURVEY.md §8 out of scope). The name
    exists so that the reference's src/oak/torch.py, which annotates with it, imports; constructing one raises."""

    def __init__(self, size):
        raise NotImplementedError("the team-builder (BuildTrajectories, read_build_trajectories) is outside the battle path")


def read_build_trajectories(*args, **kwargs):
    raise NotImplementedError("the team-builder (BuildTrajectories, read_build_trajectories) is outside the battle path")


def cpp_inference(battle_frames, network_path, discrete=False, budget="0"):
    """The network's value, the logits of the legal choices and their softmax for every frame (pyoak.cc:329-392) ->
    OutputBuffer (value, policy_logit, policy filled; the rest stays cleared, as there). One tensor-core launch."""
    _parse_budget(budget)
    torch = battle_frames.torch
    eng = battle_frames.engine
    eng.load_network(network_path, discrete=bool(discrete))
    n = battle_frames.rows if battle_frames.rows else battle_frames.size
    out = OutputBuffer(battle_frames.size)
    if n == 0:
        return out
    dev = battle_frames.device
    value = torch.empty(n, dtype=torch.float32, device=dev)
    logits = torch.empty((n, 2, 9), dtype=torch.float32, device=dev)
    counts = torch.empty((n, 2), dtype=torch.uint8, device=dev)
    eng.evaluate_device(battle_frames.battle[:n], battle_frames.durations[:n], battle_frames.result[:n].reshape(n), value, logits, counts)
    eng.synchronize()
    k = counts.to(torch.int64).unsqueeze(-1)
    legal = torch.arange(9, device=dev).view(1, 1, 9) < k
    prior = torch.softmax(logits.masked_fill(~legal, float("-inf")), dim=-1).masked_fill(~legal, 0.0)
    out.value[:n, 0] = value.cpu().numpy()
    out.policy_logit[:n] = logits.masked_fill(~legal, 0.0).cpu().numpy()
    out.policy[:n] = prior.cpu().numpy()
    return out


def _parse_budget(budget):
    """'1024' iterations, '100ms', '8s' (cpp/src/search.cc:292-310) -> (number, unit)"""
    s = str(budget)
    digits = len(s) - len(s.lstrip("0123456789"))
    if digits == 0:
        raise RuntimeError("Invalid search duration specification: " + s)
    number, unit = int(s[:digits]), s[digits:]
    if unit in ("", "ms", "millisec", "milliseconds", "s", "sec", "seconds"):
        return number, unit
    raise RuntimeError("Invalid search duration specification: " + s)


class Input:
    """MCTS::Input (cpp/include/search/mcts.h:62-66): one battle, its durations and the last result."""

    def __init__(self, battle=None, durations=None, result=0x50):
        self.battle = np.zeros(384, np.uint8) if battle is None else np.array(battle, np.uint8).reshape(384)
        self.durations = np.zeros(8, np.uint8) if durations is None else np.array(durations, np.uint8).reshape(8)
        self.result = int(result)


def _scalar_lib():
    import ctypes
    from .engine import load_library
    lib = load_library()
    vp = ctypes.c_void_p
    lib.pkmn_gen1_battle_update.restype = ctypes.c_uint8
    lib.pkmn_gen1_battle_update.argtypes = [vp, ctypes.c_uint8, ctypes.c_uint8, vp]
    lib.pkmn_gen1_battle_options_set.restype = None
    lib.pkmn_gen1_battle_options_set.argtypes = [vp, vp, vp, vp]
    lib.pkmn_gen1_battle_options_chance_durations.restype = vp
    lib.pkmn_gen1_battle_options_chance_durations.argtypes = [vp]
    return lib


def update(input, c1, c2):
    """oak.update (pyoak.cc:468-478): advance `input` in place by one update with its durations carried over. Runs on the
    GPU through the library's libpkmn-compatible scalar symbols (include/oakb200_pkmn.h)."""
    import ctypes
    lib = _scalar_lib()
    options = np.zeros(128, np.uint8)
    chance = np.zeros(40, np.uint8)  # probability[16] | actions[16] | durations[8]
    chance[32:40] = input.durations
    lib.pkmn_gen1_battle_options_set(options.ctypes.data, None, chance.ctypes.data, None)
    input.result = int(lib.pkmn_gen1_battle_update(input.battle.ctypes.data, int(c1), int(c2), options.ctypes.data))
    d = lib.pkmn_gen1_battle_options_chance_durations(options.ctypes.data)
    input.durations = np.ctypeslib.as_array(ctypes.cast(d, ctypes.POINTER(ctypes.c_uint8)), (8,)).copy()


class Agent:
    """RuntimeSearch::AgentParams (cpp/include/util/search.h:34-43)."""

    def __init__(self, budget="1024", bandit="ucb-1.0", eval="mc", matrix_ucb="", discrete=False, table=False):
        self.budget, self.bandit, self.eval, self.matrix_ucb, self.discrete, self.table = budget, bandit, eval, matrix_ucb, discrete, table

    def is_monte_carlo(self):
        return self.eval in ("", "mc", "montecarlo", "monte-carlo")


class Heap:
    """RuntimeSearch::Heap: the reference keeps the tree here between calls. This library's trees live inside the search
    handle and are rebuilt per call, so the object only carries the handle's engine."""

    def __init__(self):
        self._engine = None

    def empty(self):
        return True

    def type(self):
        return "oakb200"


class Output:
    """MCTS::Output as pyoak exposes it (pyoak.cc:494-574)."""

    def __init__(self):
        self.iterations = 0
        self.empirical_value = self.nash_value = 0.0
        self.duration_ms = 0
        self.visit_matrix = np.zeros((9, 9), np.uint64)
        self.value_matrix = np.zeros((9, 9), np.float64)
        for p in ("p1", "p2"):
            for f in ("prior", "empirical", "nash"):
                setattr(self, f"{p}_{f}", np.zeros(9, np.float64))


def search(input, heap, agent, output=None, seed=None, particles=64):
    """oak.search (pyoak.cc:575-584): one search from `input` under `agent` -> Output. The root is one row of the
    library's batched search (oakb200_search_*), `particles` descents in flight."""
    import time
    number, unit = _parse_budget(agent.budget)
    if unit:
        raise NotImplementedError("time budgets are not built: give an iteration count")
    parts = str(agent.bandit).split("-")
    if len(parts) < 2:
        raise RuntimeError("Could not parse bandit string: " + str(agent.bandit))
    if parts[0] not in TreeSearch.BANDITS:
        raise RuntimeError("Could not parse bandit string: " + str(agent.bandit))
    if agent.matrix_ucb or agent.table:
        raise NotImplementedError("matrix-UCB and the transposition table are not built")
    if parts[0] in ("pucb", "pexp3") and agent.is_monte_carlo():
        raise RuntimeError(f"bandit '{parts[0]}' takes its priors from a network: set agent.eval to a network path")
    if heap._engine is None:
        heap._engine = BattleEngine(max(int(particles), 1))
    eng = heap._engine
    ts = TreeSearch(eng, 1, particles=int(particles))
    try:
        if not agent.is_monte_carlo():
            eng.load_network(agent.eval, discrete=bool(agent.discrete))
            ts.set_leaf(1)
        seed = int(np.random.default_rng().integers(0, 1 << 63)) if seed is None else int(seed)
        ts.set_roots(input.battle.reshape(1, 384), input.durations.reshape(1, 8), np.array([input.result], np.uint8),
                     np.array([seed], np.uint64))
        t0 = time.perf_counter()
        ts.set_bandit(str(agent.bandit))
        ts.run(int(number))
        o, (nash, nash_value) = ts.output(), ts.nash()
        out = Output()
        out.duration_ms = int((time.perf_counter() - t0) * 1e3)
        out.iterations = int(o["iterations"][0])
        out.empirical_value, out.nash_value = float(o["empirical_value"][0]), float(nash_value[0])
        out.visit_matrix[:] = o["visit_matrix"][0]
        out.value_matrix[:] = o["value_matrix"][0]
        out.p1_empirical[:], out.p2_empirical[:] = o["empirical"][0, 0], o["empirical"][0, 1]
        out.p1_nash[:], out.p2_nash[:] = nash[0, 0], nash[0, 1]
        out.k = (int(o["k"][0, 0]), int(o["k"][0, 1]))
        if parts[0] in ("pucb", "pexp3"):  # the root's priors: softmax of the network's legal logits (mcts.h:197-210)
            value, logits, _ = eng.evaluate_host(input.battle.reshape(1, 384), input.durations.reshape(1, 8),
                                                 np.array([input.result], np.uint8))
            out.initial_value = float(value[0])
            for side, name in enumerate(("p1_prior", "p2_prior")):
                lg = logits[0, side, :out.k[side]].astype(np.float64)
                e = np.exp(lg - lg.max())
                getattr(out, name)[:out.k[side]] = e / e.sum()
        out.choices = o["choices"][0].copy()
        return out
    finally:
        ts.close()
