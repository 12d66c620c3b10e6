"""Stored games -> the learner's dense tensors, on the GPU.

Host-side mirror of pyoak's `EncodedBattleFrames` (reference cpp/src/pyoak.cc:655-673, cpp/include/py/battle/
encoded-frames.h:24-138, cpp/include/py/battle/target.h:17-70): same attribute names, shapes and dtypes, but the arrays are
torch tensors in device memory (where the learner consumes them) and `from_bytes` takes a whole .battle.data stream — any
number of games — instead of one game. `sample` mirrors pyoak.sample (pyoak.cc:111-260): one row per drawn game, taken
before a uniformly drawn stored update with at least `min_iterations` search iterations.
All of the work is two kernels of liboakb200 (replay + encode); there is no CPU path.
"""
import numpy as np

from .engine import BattleEngine, rows_index_host

POKEMON_DIM, ACTIVE_DIM, POLICY_DIM = 198, 229, 315


def _shapes(torch):
    return dict(pokemon=((2, 6, POKEMON_DIM), torch.float32), active=((2, 1, ACTIVE_DIM), torch.float32), hp=((2, 6, 1), torch.float32),
                choice_indices=((2, 9), torch.int64), k=((2, 1), torch.uint8), choice=((2, 1), torch.uint8),
                iterations=((1,), torch.int32), empirical_policies=((2, 9), torch.float32), nash_policies=((2, 9), torch.float32),
                empirical_value=((1,), torch.float32), nash_value=((1,), torch.float32), score=((1,), torch.float32))


def _frame_shapes(torch):
    t = _shapes(torch)
    for name in ("pokemon", "active", "hp", "choice_indices"):
        del t[name]
    t.update(battle=((384,), torch.uint8), durations=((8,), torch.uint8), result=((1,), torch.uint8), choices=((2, 9), torch.uint8),
             legal_choices=((2, 9), torch.uint8))
    return t


class EncodedBattleFrames:
    """`size` rows of training data in device memory. `iterations` holds the u32 of the reference bit-for-bit (torch has no
    uint32 arithmetic; view it with .view(torch.uint32) where needed)."""

    def __init__(self, size, device=0, engine=None):
        import torch
        self.torch = torch
        self.size = int(size)
        self.device = torch.device("cuda", device)
        self.engine = engine or BattleEngine(1, device=device)
        self._own_engine = engine is None
        self.engine.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        for name, (shape, dtype) in self._fields(torch).items():
            setattr(self, name, torch.zeros((self.size,) + shape, dtype=dtype, device=self.device))
        self.rows = 0
        self.games = 0

    _fields = staticmethod(_shapes)

    def close(self):
        if self._own_engine:
            self.engine.close()

    def _out(self):
        return {name: getattr(self, name, None) for name in BattleEngine.ROW_FIELDS}

    def clear(self):
        for name in self._fields(self.torch):
            getattr(self, name).zero_()

    def host(self):
        """The same attribute names as numpy arrays in host memory (plus `size`) — what the reference's wrappers that call
        torch.from_numpy on each field (src/oak/torch.py:16-28) take. `iterations` comes back as the reference's uint32."""
        import types
        h = types.SimpleNamespace(size=self.size, rows=self.rows, games=self.games)
        for name in self._fields(self.torch):
            a = getattr(self, name).cpu().numpy()
            setattr(h, name, a.view(np.uint32) if name == "iterations" else a)
        return h

    def fill(self, data, only_update=None):
        """data: bytes / uint8 array (pinned tensors upload fastest) holding whole games. -> rows written.
        Raises if a game's replay does not end on its stored result (the reference asserts that)."""
        if isinstance(data, (bytes, bytearray, memoryview)):
            data = np.frombuffer(data, np.uint8)
        games, rows, bad = self.engine.rows_build_host(data, self._out(), self.size, only_update)
        self.games, self.rows = games, rows
        if bad >= 0:
            raise ValueError(f"game {bad} of the stream does not replay to its stored result")
        return rows

    @classmethod
    def from_bytes(cls, data, size=None, device=0, engine=None):
        """from_bytes(data, size): every stored update of every game in `data` becomes a row."""
        if isinstance(data, (bytes, bytearray, memoryview)):
            data = np.frombuffer(data, np.uint8)
        if size is None:
            size = rows_index_host(data if isinstance(data, np.ndarray) else data.numpy())[2]
        f = cls(size, device=device, engine=engine)
        f.fill(data)
        return f

    def sample(self, data, min_iterations=0, max_battle_length=None, rng=None):
        """pyoak.sample over an in-memory stream: for every game (up to `size`, longer-than-max games skipped) draw one
        stored update with iterations >= min_iterations and write the row built before it. -> rows written."""
        rng = rng or np.random.default_rng()
        data = np.frombuffer(data, np.uint8) if isinstance(data, (bytes, bytearray, memoryview)) else np.ascontiguousarray(data, np.uint8)
        off, _, _ = rows_index_host(data)
        keep, picks = [], []
        for o in off:
            o = int(o)
            nb = int.from_bytes(data[o:o + 4].tobytes(), "little")
            nu = int.from_bytes(data[o + 4:o + 6].tobytes(), "little")
            if nu == 0 or (max_battle_length is not None and nu > max_battle_length):
                continue
            p, valid = o + 391, []
            for u in range(nu):
                if p + 11 > o + nb:  # a corrupt record: the stored (m, n) walked out of the game
                    raise ValueError(f"sample: update {u} of the game at byte {o} lies outside its record")
                hdr = int(data[p])
                if int.from_bytes(data[p + 3:p + 7].tobytes(), "little") >= min_iterations:
                    valid.append(u)
                p += 11 + 4 * ((hdr & 15) + (hdr >> 4) + 2)
            if p != o + nb:
                raise ValueError(f"sample: the game at byte {o} does not end on its last update")
            if valid and len(keep) < self.size:
                keep.append((o, nb))
                picks.append(valid[int(rng.integers(0, len(valid)))])
        if not keep:
            self.rows = self.games = 0
            return 0
        sub = np.concatenate([data[o:o + nb] for o, nb in keep])
        return self.fill(sub, only_update=np.asarray(picks, np.int32))


class BattleFrames(EncodedBattleFrames):
    """pyoak's `BattleFrames` (reference cpp/include/py/battle/frames.h:17-80, pyoak.cc:638-653): the same replay, but each
    row keeps the raw state — battle[size,384], durations[size,8], result[size,1] u8 — next to the targets instead of the
    dense encodings. `choices` stays zero as in the reference (frames.h:47-49 computes the legal choices and does not store
    them); `legal_choices` is this library's extra: the list `choices` was meant to hold, zero past k."""
    _fields = staticmethod(_frame_shapes)
