#!/usr/bin/env python3
"""Golden training rows produced by THE REFERENCE'S OWN row builder (py/battle/encoded-frames.h compiled in place into
oracle/_ref/ref_pyframes*.so by oracle/ref_pyframes.cpp; the libpkmn calls inside it are served by the oracle engine).
Writes tests/golden/training_rows_golden.npz: a .battle.data stream of 8 played games and, for every stored update, the
twelve arrays EncodedBattleFrames.from_bytes returns; and tests/golden/battle_frames_golden.npz: for the same stream, the
battle / durations / result / choices arrays of BattleFrames.from_bytes (py/battle/frames.h, same module). Run in the build
container only (needs /root/reference)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import ref_pyframes  # noqa: E402
from oak_ref import Oracle, ROW_FIELDS, play_frame_games, split_games  # noqa: E402

orc = Oracle()
stream = play_frame_games(orc, first=40_000, games=8, seed=77)
parts = {name: [] for name in ROW_FIELDS}
for g in split_games(stream):
    n = int.from_bytes(bytes(g[4:6]), "little")
    ref = ref_pyframes.from_bytes(bytes(g), n)
    for name in ROW_FIELDS:
        parts[name].append(np.asarray(ref[name]))
out = {name: np.concatenate(v) for name, v in parts.items()}
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"), stream=stream, **out)
print("wrote", stream.size, "stream bytes,", out["k"].shape[0], "rows")

FRAME_FIELDS = ("battle", "durations", "result", "choices")
fparts = {name: [] for name in FRAME_FIELDS}
for g in split_games(stream):
    n = int.from_bytes(bytes(g[4:6]), "little")
    ref = ref_pyframes.frames_from_bytes(bytes(g), n)
    for name in FRAME_FIELDS:
        fparts[name].append(np.asarray(ref[name]))
    for name in ("k", "choice", "iterations", "empirical_policies", "nash_policies", "empirical_value", "nash_value", "score"):
        assert np.array_equal(np.asarray(ref[name]), np.asarray(ref_pyframes.from_bytes(bytes(g), n)[name])), name
fout = {name: np.concatenate(v) for name, v in fparts.items()}
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "battle_frames_golden.npz"), **fout)
print("wrote", fout["battle"].shape[0], "uncompressed frames")
