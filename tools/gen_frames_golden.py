#!/usr/bin/env python3
"""Golden .battle.data stream produced by THE REFERENCE'S OWN serialiser (train/battle/compressed-frame.h compiled in
place into oracle/_ref/libref_oak.so by oracle/ref_shim.cpp). Writes tests/golden/frames_golden.npz: the expected
stream for random_frame_games(default_rng(2024), 24, 40). Run in the build container only."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oak_ref import frames_expected_game, random_frame_games, ref_library  # noqa: E402

ref = ref_library()
assert ref is not None, "build oracle/_ref first (make -C oracle)"
G = random_frame_games(np.random.default_rng(2024), 24, 40)
stream = np.concatenate([frames_expected_game(ref, "ref_frames_write_game", G, g) for g in range(24)])
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "frames_golden.npz"), stream=stream)
print("wrote", stream.size, "bytes for 24 games")
