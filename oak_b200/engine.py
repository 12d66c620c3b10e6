"""Host-side mirror of the reference's battle interface over the oakb200 C ABI.

Reference interface being mirrored (same argument meaning and error behaviour):
  PKMN::choices / pkmn_gen1_battle_choices   cpp/include/libpkmn/pkmn.h:141-156, engine pkmn.h:283-288
  PKMN::update  / pkmn_gen1_battle_update    cpp/include/libpkmn/pkmn.h:106-139, engine pkmn.h:264-268
  MCTS rollout                                cpp/include/search/mcts.h:448-496
  PKMN::battle(p1, p2, seed) + update(0, 0)  cpp/include/libpkmn/pkmn.h:50-57, cpp/src/generate.cc:238-240
Batches are numpy uint8 arrays in the reference's AoS record formats ([n,384] battles, [n,8]
durations, [n] results), or torch CUDA tensors for the *_device calls.
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboakb200.so")

ROLLOUT_WRITE_STATE = 1
ROLLOUT_TRACK_CHANCE = 2

_lib = None


class OakB200Error(RuntimeError):
    pass


def load_library():
    """dlopen oak_b200/liboakb200.so (built in-tree by oak_b200.build). Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("OAKB200_LIB", LIB_PATH)
    if not os.path.exists(path):
        raise OakB200Error(f"{path} is missing: run `python -m oak_b200.build` (needs nvcc). "
                           "oak_b200 has no CPU fallback.")
    lib = ctypes.CDLL(path)
    vp, sz, u32, i32, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int, ctypes.c_uint64
    sig = {
        "oakb200_create": (i32, [i32, sz, ctypes.POINTER(vp)]),
        "oakb200_destroy": (None, [vp]),
        "oakb200_last_error": (ctypes.c_char_p, [vp]),
        "oakb200_device": (i32, [vp]),
        "oakb200_set_stream": (i32, [vp, vp]),
        "oakb200_synchronize": (i32, [vp]),
        "oakb200_launch_count": (ctypes.c_uint64, [vp]),
        "oakb200_upload": (i32, [vp, sz, vp, vp, vp]),
        "oakb200_download": (i32, [vp, sz, vp, vp, vp]),
        "oakb200_resident": (i32, [vp, ctypes.POINTER(vp), ctypes.POINTER(vp), ctypes.POINTER(vp), ctypes.POINTER(sz)]),
        "oakb200_choices": (i32, [vp, vp, vp]),
        "oakb200_choices_device": (i32, [vp, sz, vp, vp, vp, vp]),
        "oakb200_step": (i32, [vp, vp, vp, vp, vp, vp]),
        "oakb200_step_device": (i32, [vp, sz, vp, vp, vp, vp, vp, vp, vp]),
        "oakb200_step_random_device": (i32, [vp, sz, vp, vp, vp, vp, u32, vp, vp]),
        "oakb200_rollout": (i32, [vp, vp, u32, u32, vp, vp, vp]),
        "oakb200_rollout_device": (i32, [vp, sz, vp, vp, vp, vp, u32, u32, vp, vp, vp]),
        "oakb200_rollout_host": (i32, [vp, sz, vp, vp, vp, vp, u32, vp, vp, vp]),
        "oakb200_total_steps": (i32, [vp, ctypes.POINTER(ctypes.c_uint64)]),
        "oakb200_randomize_hidden_device": (i32, [vp, sz, vp, vp]),
        "oakb200_build_battles": (i32, [vp, sz, vp, vp]),
        "oakb200_net_load": (i32, [vp, vp, sz]),
        "oakb200_net_load_discrete": (i32, [vp, vp, sz]),
        "oakb200_eval_set_mode": (i32, [vp, i32]),
        "oakb200_respawn_device": (i32, [vp, sz, vp, vp, vp, vp, vp]),
        "oakb200_policy_sample_device": (i32, [vp, sz, vp, vp, vp, vp, u32, ctypes.c_double, ctypes.c_double, vp, vp, vp]),
        "oakb200_frames_create": (i32, [vp, sz, u32]),
        "oakb200_frames_begin_device": (i32, [vp, sz, vp, vp]),
        "oakb200_frames_record_device": (i32, [vp, sz, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
        "oakb200_frames_finish_device": (i32, [vp, sz, vp]),
        "oakb200_selfplay_step_device": (i32, [vp, sz, vp, vp, vp, vp, u32, ctypes.c_double, ctypes.c_double] + [vp] * 9),
        "oakb200_frames_pending": (i32, [vp, ctypes.POINTER(u32), ctypes.POINTER(u32)]),
        "oakb200_frames_drain": (i32, [vp, vp, sz, ctypes.POINTER(sz), ctypes.POINTER(u32)]),
        "oakb200_rows_index_host": (i32, [vp, u64, vp, vp, u64, ctypes.POINTER(u64), ctypes.POINTER(u64)]),
        "oakb200_rows_build_device": (i32, [vp, vp, u64, vp, vp, vp, u64, u64, vp]),
        "oakb200_rows_build_static": (i32, [vp, vp, u64, vp, vp, vp, u64, u64, vp, vp]),
        "oakb200_rows_status": (i32, [vp, ctypes.POINTER(ctypes.c_int64)]),
        "oakb200_rows_build_host": (i32, [vp, vp, u64, vp, vp, u64, ctypes.POINTER(u64), ctypes.POINTER(u64),
                                         ctypes.POINTER(ctypes.c_int64)]),
        "oakb200_step_log": (i32, [vp, vp, vp, vp, vp, vp, vp, u32, vp]),
        "oakb200_step_log_device": (i32, [vp, sz, vp, vp, vp, vp, vp, vp, vp, vp, u32, vp]),
        "oakb200_eval": (i32, [vp, vp, vp, vp]),
        "oakb200_eval_device": (i32, [vp, sz, vp, vp, vp, vp, vp, vp]),
        "oakb200_eval_host": (i32, [vp, sz, vp, vp, vp, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _np(a, dtype, shape=None):
    a = np.ascontiguousarray(a, dtype=dtype)
    if shape is not None and tuple(a.shape) != tuple(shape):
        raise ValueError(f"expected shape {shape}, got {a.shape}")
    return a


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


class BattleEngine:
    """One context per GPU holding a resident batch of battles (include/oakb200.h)."""

    def __init__(self, n_max, device=0):
        self.lib = load_library()
        h = ctypes.c_void_p()
        rc = self.lib.oakb200_create(int(device), int(n_max), ctypes.byref(h))
        self._h = h if h.value else None
        if rc != 0:
            msg = self.lib.oakb200_last_error(h).decode() if h.value else "no CUDA device (oak_b200 has no CPU fallback)"
            if h.value:
                self.lib.oakb200_destroy(h)
                self._h = None
            raise OakB200Error(f"oakb200_create failed ({rc}): {msg}")
        self.n_max = int(n_max)
        self.n = 0
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            self.lib.oakb200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise OakB200Error(f"oakb200 error {rc}: {self.lib.oakb200_last_error(self._h).decode()}")

    # ---------------------------------------------------------------- plumbing
    def set_stream(self, cuda_stream_ptr):
        self._check(self.lib.oakb200_set_stream(self._h, ctypes.c_void_p(cuda_stream_ptr)))

    def synchronize(self):
        self._check(self.lib.oakb200_synchronize(self._h))

    @property
    def launch_count(self):
        return int(self.lib.oakb200_launch_count(self._h))

    def total_steps(self):
        v = ctypes.c_uint64()
        self._check(self.lib.oakb200_total_steps(self._h, ctypes.byref(v)))
        return v.value

    # ---------------------------------------------------------------- resident batch (host arrays)
    def upload(self, battles, durations=None, results=None):
        battles = _np(battles, np.uint8)
        n = battles.shape[0]
        battles = _np(battles, np.uint8, (n, 384))
        durations = None if durations is None else _np(durations, np.uint8, (n, 8))
        results = None if results is None else _np(results, np.uint8, (n,))
        self._check(self.lib.oakb200_upload(self._h, n, _ptr(battles), _ptr(durations), _ptr(results)))
        self.n = n

    def download(self):
        n = self.n
        b = np.empty((n, 384), np.uint8)
        d = np.empty((n, 8), np.uint8)
        r = np.empty((n,), np.uint8)
        self._check(self.lib.oakb200_download(self._h, n, _ptr(b), _ptr(d), _ptr(r)))
        return b, d, r

    def build_battles(self, teams, seeds):
        """teams[n,2,6,5] = {species, move x4}; seeds[n] u64. PKMN::battle + update(0,0) on the device."""
        teams = _np(teams, np.uint8)
        n = teams.shape[0]
        teams = _np(teams, np.uint8, (n, 2, 6, 5))
        seeds = _np(seeds, np.uint64, (n,))
        self._check(self.lib.oakb200_build_battles(self._h, n, _ptr(teams), _ptr(seeds)))
        self.n = n

    def choices(self):
        """-> (choices[n,2,9] u8, counts[n,2] u8): pkmn_gen1_battle_choices for both players."""
        out = np.empty((self.n, 2, 9), np.uint8)
        cnt = np.empty((self.n, 2), np.uint8)
        self._check(self.lib.oakb200_choices(self._h, _ptr(out), _ptr(cnt)))
        return out, cnt

    def update(self, c1, c2, overrides=None, want_actions=False):
        """options_set(null[, calc overrides]) + pkmn_gen1_battle_update for every resident battle."""
        n = self.n
        c1 = _np(c1, np.uint8, (n,))
        c2 = _np(c2, np.uint8, (n,))
        overrides = None if overrides is None else _np(overrides, np.uint8, (n, 16))
        res = np.empty((n,), np.uint8)
        act = np.empty((n, 16), np.uint8) if want_actions else None
        self._check(self.lib.oakb200_step(self._h, _ptr(c1), _ptr(c2), _ptr(overrides), _ptr(res), _ptr(act)))
        return (res, act) if want_actions else res

    def update_log(self, c1, c2, overrides=None, log_stride=180):
        """update() with libpkmn's binary protocol log -> (results[n], actions[n,16], logs[n,log_stride] u8, lens[n] u32).
        A battle whose log does not fit log_stride bytes gets the result byte 0x0C (pkmn_error)."""
        n = self.n
        c1 = _np(c1, np.uint8, (n,))
        c2 = _np(c2, np.uint8, (n,))
        overrides = None if overrides is None else _np(overrides, np.uint8, (n, 16))
        res = np.empty((n,), np.uint8)
        act = np.empty((n, 16), np.uint8)
        logs = np.zeros((n, max(int(log_stride), 1)), np.uint8)
        lens = np.empty((n,), np.uint32)
        self._check(self.lib.oakb200_step_log(self._h, _ptr(c1), _ptr(c2), _ptr(overrides), _ptr(res), _ptr(act), _ptr(logs),
                                              int(log_stride), _ptr(lens)))
        return res, act, logs, lens

    def rollout(self, seeds, max_steps=100000, write_state=False, track_chance=False):
        """Random playouts to a terminal result -> (score2[n] u8, turns[n] u16, steps[n] u32)."""
        n = self.n
        seeds = _np(seeds, np.uint64, (n,))
        score = np.empty((n,), np.uint8)
        turns = np.empty((n,), np.uint16)
        steps = np.empty((n,), np.uint32)
        flags = (ROLLOUT_WRITE_STATE if write_state else 0) | (ROLLOUT_TRACK_CHANCE if track_chance else 0)
        self._check(self.lib.oakb200_rollout(self._h, _ptr(seeds), int(max_steps), flags, _ptr(score), _ptr(turns), _ptr(steps)))
        return score, turns, steps

    def rollout_host(self, battles, durations, results, seeds, max_steps=100000):
        """upload + rollout with the transfers pipelined behind the playouts -> (score2, turns, steps)."""
        battles = _np(battles, np.uint8)
        n = battles.shape[0]
        battles = _np(battles, np.uint8, (n, 384))
        durations = None if durations is None else _np(durations, np.uint8, (n, 8))
        results = None if results is None else _np(results, np.uint8, (n,))
        seeds = _np(seeds, np.uint64, (n,))
        score, turns, steps = np.empty((n,), np.uint8), np.empty((n,), np.uint16), np.empty((n,), np.uint32)
        self._check(self.lib.oakb200_rollout_host(self._h, n, _ptr(battles), _ptr(durations), _ptr(results), _ptr(seeds),
                                                  int(max_steps), _ptr(score), _ptr(turns), _ptr(steps)))
        self.n = n
        return score, turns, steps

    # ---------------------------------------------------------------- leaf evaluation
    def load_network(self, blob, discrete=False):
        """Parameter file image in the reference's format (bytes / uint8 array / path). discrete=True loads it as the
        reference's `.discrete` agent: float embedding nets, u8 embeddings, int8 main net."""
        if isinstance(blob, (str, os.PathLike)):
            blob = np.fromfile(blob, np.uint8)
        blob = np.frombuffer(bytes(blob), np.uint8) if isinstance(blob, (bytes, bytearray)) else _np(blob, np.uint8)
        fn = self.lib.oakb200_net_load_discrete if discrete else self.lib.oakb200_net_load
        self._check(fn(self._h, _ptr(blob), blob.size))

    def set_eval_mode(self, mode):
        """0 auto (tensor cores when the layer shapes fit), 1 fp32 CUDA-core kernel, 2 tensor cores or error."""
        self._check(self.lib.oakb200_eval_set_mode(self._h, int(mode)))

    def evaluate(self):
        """value_policy_inference for every resident battle -> (value[n] f32, logits[n,2,9] f32, counts[n,2] u8)."""
        n = self.n
        v = np.empty((n,), np.float32)
        lg = np.empty((n, 2, 9), np.float32)
        cnt = np.empty((n, 2), np.uint8)
        self._check(self.lib.oakb200_eval(self._h, _ptr(v), _ptr(lg), _ptr(cnt)))
        return v, lg, cnt

    def evaluate_host(self, battles, durations, results, value_out=None, logits_out=None, counts_out=None):
        """Host records in, value / logits / counts out, transfers pipelined behind the kernels (pinned arrays are fastest)."""
        n = battles.shape[0]
        value_out = np.empty(n, np.float32) if value_out is None else value_out
        logits_out = np.empty((n, 2, 9), np.float32) if logits_out is None else logits_out
        counts_out = np.empty((n, 2), np.uint8) if counts_out is None else counts_out
        self._check(self.lib.oakb200_eval_host(self._h, n, _ptr(battles), _ptr(durations), _ptr(results), _ptr(value_out),
                                               _ptr(logits_out), _ptr(counts_out)))
        self.n = n
        return value_out, logits_out, counts_out

    def evaluate_device(self, battles, durations, results, value_out, logits_out, counts_out):
        self._check(self.lib.oakb200_eval_device(self._h, battles.shape[0], _ptr(battles), _ptr(durations), _ptr(results),
                                                 _ptr(value_out), _ptr(logits_out), _ptr(counts_out)))

    # ---------------------------------------------------------------- device-pointer calls (torch CUDA tensors)
    def resident(self):
        b, d, r, n = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_size_t()
        self._check(self.lib.oakb200_resident(self._h, ctypes.byref(b), ctypes.byref(d), ctypes.byref(r), ctypes.byref(n)))
        return b.value, d.value, r.value, n.value

    def choices_device(self, battles, results, out, counts):
        self._check(self.lib.oakb200_choices_device(self._h, battles.shape[0], _ptr(battles), _ptr(results), _ptr(out), _ptr(counts)))

    def step_device(self, battles, durations, results, c1, c2, overrides=None, actions_out=None):
        self._check(self.lib.oakb200_step_device(self._h, battles.shape[0], _ptr(battles), _ptr(durations), _ptr(results),
                                                 _ptr(c1), _ptr(c2), _ptr(overrides), _ptr(actions_out)))

    def step_random_device(self, battles, durations, results, seeds, step_index, actions_out=None, chosen_out=None):
        self._check(self.lib.oakb200_step_random_device(self._h, battles.shape[0], _ptr(battles), _ptr(durations), _ptr(results),
                                                        _ptr(seeds), int(step_index), _ptr(actions_out), _ptr(chosen_out)))

    def rollout_device(self, battles, durations, results, seeds, max_steps=100000, flags=0, score2_out=None, turns_out=None,
                       steps_out=None):
        self._check(self.lib.oakb200_rollout_device(self._h, battles.shape[0], _ptr(battles), _ptr(durations), _ptr(results),
                                                    _ptr(seeds), int(max_steps), int(flags), _ptr(score2_out), _ptr(turns_out),
                                                    _ptr(steps_out)))

    def randomize_hidden_device(self, battles, durations):
        self._check(self.lib.oakb200_randomize_hidden_device(self._h, battles.shape[0], _ptr(battles), _ptr(durations)))

    def respawn_device(self, battles, durations, results, teams, seeds):
        """Rebuild in place every battle whose result says it has ended (teams[n,2,6,5], seeds[n] on the device)."""
        self._check(self.lib.oakb200_respawn_device(self._h, battles.shape[0], _ptr(battles), _ptr(durations), _ptr(results),
                                                    _ptr(teams), _ptr(seeds)))

    # ---------------------------------------------------------------- policy sampling + training frames (device pointers)
    def policy_sample_device(self, choices, counts, logits, seeds, step_index, probs_out, c1_out, c2_out, temperature=1.0,
                             min_prob=0.0):
        self._check(self.lib.oakb200_policy_sample_device(self._h, counts.shape[0], _ptr(choices), _ptr(counts), _ptr(logits),
                                                          _ptr(seeds), int(step_index), float(temperature), float(min_prob),
                                                          _ptr(probs_out), _ptr(c1_out), _ptr(c2_out)))

    def frames_create(self, games, max_updates=1024):
        self._check(self.lib.oakb200_frames_create(self._h, int(games), int(max_updates)))

    def frames_begin_device(self, battles, results):
        self._check(self.lib.oakb200_frames_begin_device(self._h, results.shape[0], _ptr(battles), _ptr(results)))

    def frames_record_device(self, results, c1, c2, counts, empirical_value, nash_value, empirical, nash, iterations=None):
        self._check(self.lib.oakb200_frames_record_device(self._h, results.shape[0], _ptr(results), _ptr(c1), _ptr(c2), _ptr(counts),
                                                          _ptr(iterations), _ptr(empirical_value), _ptr(nash_value),
                                                          _ptr(empirical), _ptr(nash)))

    def frames_finish_device(self, results):
        self._check(self.lib.oakb200_frames_finish_device(self._h, results.shape[0], _ptr(results)))

    def selfplay_step_device(self, battles, durations, results, seeds, step_index, temperature, min_prob, choices, counts, value,
                             logits, eval_counts, probs, c1, c2, before):
        self._check(self.lib.oakb200_selfplay_step_device(self._h, results.shape[0], _ptr(battles), _ptr(durations), _ptr(results),
                                                          _ptr(seeds), int(step_index), float(temperature), float(min_prob),
                                                          _ptr(choices), _ptr(counts), _ptr(value), _ptr(logits), _ptr(eval_counts),
                                                          _ptr(probs), _ptr(c1), _ptr(c2), _ptr(before)))

    def frames_pending(self):
        a, b = ctypes.c_uint32(), ctypes.c_uint32()
        self._check(self.lib.oakb200_frames_pending(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def frames_drain(self, capacity=None, out=None):
        """-> (stream bytes as a uint8 array, games in it): finished games in slot order, the reference's .battle.data format.
        `out` (a torch uint8 CUDA tensor) keeps the stream on the device — for a gather over NCCL — and is returned sliced;
        ECAPACITY if it is too small."""
        need, games = ctypes.c_size_t(), ctypes.c_uint32()
        if out is not None:
            self._check(self.lib.oakb200_frames_drain(self._h, _ptr(out), out.numel(), ctypes.byref(need), ctypes.byref(games)))
            return out[:need.value], games.value
        if capacity is None:
            rc = self.lib.oakb200_frames_drain(self._h, None, 0, ctypes.byref(need), ctypes.byref(games))
            if need.value == 0:
                self._check(rc)
                return np.empty(0, np.uint8), 0
            capacity = need.value
        # device -> page-locked staging (cached, grown geometrically) -> the caller's own array: a copy straight into
        # pageable memory runs at a fraction of the PCIe rate, and steady-state self-play drains ~4 MB per step
        stage = getattr(self, "_drain_stage", None)
        if stage is None or stage.numel() < int(capacity):
            import torch
            stage = self._drain_stage = torch.empty(max(2 * int(capacity), 1 << 20), dtype=torch.uint8).pin_memory()
        self._check(self.lib.oakb200_frames_drain(self._h, ctypes.c_void_p(stage.data_ptr()), int(capacity), ctypes.byref(need),
                                                  ctypes.byref(games)))
        return stage.numpy()[:need.value].copy(), games.value

    # ---------------------------------------------------------------- training rows (include/oakb200.h, train_rows.cu)
    ROW_FIELDS = ("pokemon", "active", "hp", "choice_indices", "k", "choice", "iterations", "empirical_policies", "nash_policies",
                  "empirical_value", "nash_value", "score", "battle", "durations", "result", "legal_choices")

    def _rows_struct(self, out):
        """dict name -> device tensor (or None) -> the oakb200_training_rows struct as an array of 16 pointers"""
        arr = (ctypes.c_void_p * 16)()
        for i, name in enumerate(self.ROW_FIELDS):
            t = out.get(name)
            arr[i] = _ptr(t) if t is not None else None
        return arr

    def rows_build_host(self, stream, out, capacity_rows, only_update=None):
        """Host .battle.data bytes -> rows in the device tensors of `out`. -> (games, rows, first_bad_game or -1)."""
        stream = np.ascontiguousarray(stream, np.uint8) if isinstance(stream, np.ndarray) else stream
        nbytes = stream.size if isinstance(stream, np.ndarray) else stream.numel()
        only = None if only_update is None else _np(only_update, np.int32)
        g, r, bad = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_int64()
        st = self._rows_struct(out)
        self._check(self.lib.oakb200_rows_build_host(self._h, _ptr(stream), nbytes, _ptr(only), ctypes.addressof(st), int(capacity_rows),
                                                     ctypes.byref(g), ctypes.byref(r), ctypes.byref(bad)))
        return g.value, r.value, bad.value

    def rows_build_device(self, stream, n_bytes, game_off, row_base, n_games, n_rows, out, only_update=None):
        st = self._rows_struct(out)
        self._check(self.lib.oakb200_rows_build_device(self._h, _ptr(stream), int(n_bytes), _ptr(game_off), _ptr(row_base),
                                                       _ptr(only_update), int(n_games), int(n_rows), ctypes.addressof(st)))

    def rows_build_static(self, stream, n_bytes, game_off, row_base, n_games, n_rows, out, only_update=None, first_bad_out=None):
        """oakb200_rows_build_static: the same build with the library's static device scratch (no allocation; n_rows <= 16384)."""
        st = self._rows_struct(out)
        self._check(self.lib.oakb200_rows_build_static(self._h, _ptr(stream), int(n_bytes), _ptr(game_off), _ptr(row_base),
                                                       _ptr(only_update), int(n_games), int(n_rows), ctypes.addressof(st),
                                                       _ptr(first_bad_out)))

    def rows_status(self):
        bad = ctypes.c_int64()
        self._check(self.lib.oakb200_rows_status(self._h, ctypes.byref(bad)))
        return bad.value


def rows_index_host(stream):
    """Header walk of a .battle.data stream -> (game_off u64[G], row_base u32[G], rows). No GPU needed."""
    lib = load_library()
    stream = np.ascontiguousarray(stream, np.uint8)
    g, r = ctypes.c_uint64(), ctypes.c_uint64()
    rc = lib.oakb200_rows_index_host(_ptr(stream), stream.size, None, None, 0, ctypes.byref(g), ctypes.byref(r))
    if rc != 0:
        raise OakB200Error(f"oakb200_rows_index_host: error {rc} (the stream does not split into whole games)")
    off, base = np.zeros(g.value, np.uint64), np.zeros(g.value, np.uint32)
    rc = lib.oakb200_rows_index_host(_ptr(stream), stream.size, _ptr(off), _ptr(base), g.value, None, None)
    if rc != 0:
        raise OakB200Error(f"oakb200_rows_index_host: error {rc}")
    return off, base, r.value
