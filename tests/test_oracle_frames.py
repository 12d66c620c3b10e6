"""The frame-serialisation oracle (oracle/frames_oracle.hpp) against the reference's own serialiser: the committed golden
stream (tools/gen_frames_golden.py, made with train/battle/compressed-frame.h compiled in place) and, where oracle/_ref is
built, the live reference library on fresh random games, including a round trip through the reference's reader."""
import ctypes
import os

import numpy as np
import pytest

from oak_ref import ROOT, Oracle, frames_expected_game, parse_battle_data, random_frame_games, ref_library


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def test_oracle_matches_reference_golden_stream(orc):
    want = np.load(os.path.join(ROOT, "tests", "golden", "frames_golden.npz"))["stream"]
    G = random_frame_games(np.random.default_rng(2024), 24, 40)
    got = np.concatenate([frames_expected_game(orc.lib, "oracle_frames_write_game", G, g) for g in range(24)])
    assert np.array_equal(got, want)


def test_python_reader_round_trip(orc):
    G = random_frame_games(np.random.default_rng(5), 10, 30)
    stream = np.concatenate([frames_expected_game(orc.lib, "oracle_frames_write_game", G, g) for g in range(10)])
    games = parse_battle_data(stream)
    assert len(games) == 10
    for g, game in enumerate(games):
        n = int(G["L"][g])
        assert np.array_equal(game["battle"], G["battles"][g]) and game["result"] == G["final"][g] and len(game["updates"]) == n
        for t, u in enumerate(game["updates"]):
            assert (u["m"], u["n"], u["c1"], u["c2"]) == (G["k"][t, g, 0], G["k"][t, g, 1], G["choice"][t, g, 0], G["choice"][t, g, 1])
            assert u["iterations"] == G["iters"][t, g]
            assert u["ev"] == int(np.float64(G["ev"][t, g]) * 65535.0)
            assert np.array_equal(u["p2_nash"], (G["nash"][t, g, 1, :u["n"]].astype(np.float64) * 65535.0).astype(np.uint16))


def test_oracle_matches_live_reference(orc):
    ref = ref_library()
    if ref is None:
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    ref.ref_frames_read_game.restype = ctypes.c_uint32
    G = random_frame_games(np.random.default_rng(99), 40, 120)
    for g in range(40):
        a = frames_expected_game(orc.lib, "oracle_frames_write_game", G, g)
        b = frames_expected_game(ref, "ref_frames_write_game", G, g)
        assert np.array_equal(a, b)
        battle, result = np.zeros(384, np.uint8), np.zeros(1, np.uint8)
        n = ref.ref_frames_read_game(a.ctypes.data_as(ctypes.c_void_p), battle.ctypes.data_as(ctypes.c_void_p),
                                     result.ctypes.data_as(ctypes.c_void_p))
        assert n == G["L"][g] and result[0] == G["final"][g] and np.array_equal(battle, G["battles"][g])


def test_policy_sample_oracle_basics(orc):
    f = orc.lib.oracle_policy_sample
    f.restype = ctypes.c_int
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_void_p]
    lg = np.array([0.0, np.log(3.0), 0, 0, 0, 0, 0, 0, 0], np.float32)
    pr = np.zeros(9, np.float32)
    assert f(lg.ctypes.data, 2, 1.0, 0.0, 0.2, pr.ctypes.data) == 0
    np.testing.assert_allclose(pr[:2], [0.25, 0.75], rtol=1e-6)
    assert f(lg.ctypes.data, 2, 1.0, 0.0, 0.26, pr.ctypes.data) == 1
    assert f(lg.ctypes.data, 2, 1.0, 0.3, 0.01, pr.ctypes.data) == 1  # the 0.25 entry falls under the floor
    assert f(lg.ctypes.data, 2, 3.0, 0.0, 0.03, pr.ctypes.data) == 0  # 1/28 after p^3
    assert f(lg.ctypes.data, 2, 3.0, 0.0, 0.04, pr.ctypes.data) == 1
