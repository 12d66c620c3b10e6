"""Training rows on the GPU (oakb200_rows_*: replay_kernel + rows_encode_kernel) against the oracle restatement and the
golden rows made by the reference's own EncodedFrames::from_bytes. Bit-exact, floats included."""
import os

import numpy as np
import pytest

from oak_ref import ROOT, ROW_FIELDS, Oracle, count_rows, oracle_training_rows, play_frame_games, split_games

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def _host(f, name):
    a = getattr(f, name).cpu().numpy()
    return a.view(np.uint32) if name == "iterations" else a


def _check(f, want, n):
    for name in ROW_FIELDS:
        got = _host(f, name)[:n]
        assert got.shape == want[name].shape and got.dtype == want[name].dtype, name
        assert np.array_equal(got, want[name]), name


def test_golden_rows_from_the_reference():
    from oak_b200.training_rows import EncodedBattleFrames
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    f = EncodedBattleFrames.from_bytes(G["stream"].tobytes())
    assert f.rows == 1107 and f.games == 8
    _check(f, G, 1107)
    f.close()


def test_many_games_match_the_oracle(orc):
    import torch
    from oak_b200.training_rows import EncodedBattleFrames
    stream = play_frame_games(orc, first=700_000, games=1500, seed=99)
    want = oracle_training_rows(orc, stream)
    n = want["k"].shape[0]
    f = EncodedBattleFrames(n + 7)
    pinned = torch.from_numpy(stream).pin_memory()
    assert f.fill(pinned) == n
    _check(f, want, n)
    # a second fill into the same (dirty) tensors gives the same rows: nothing relies on a cleared output
    for name in ROW_FIELDS:
        getattr(f, name).fill_(1)
    assert f.fill(pinned) == n
    _check(f, want, n)
    f.close()


def test_sample_mode_builds_the_drawn_row(orc):
    from oak_b200.engine import BattleEngine, rows_index_host
    from oak_b200.training_rows import EncodedBattleFrames
    stream = play_frame_games(orc, first=31_000, games=64, seed=3)
    want = oracle_training_rows(orc, stream)
    off, base, rows = rows_index_host(stream)
    lens = np.diff(np.append(base, rows))
    rng = np.random.default_rng(5)
    picks = (rng.random(64) * lens).astype(np.int32)
    picks[0], picks[1] = 0, lens[1] - 1
    f = EncodedBattleFrames(64)
    g, r, bad = f.engine.rows_build_host(stream, f._out(), 64, picks)
    assert (g, r, bad) == (64, 64, -1)
    sel = base + picks
    _check(f, {name: want[name][sel] for name in ROW_FIELDS}, 64)
    # the Python-level sampler draws valid updates only
    f2 = EncodedBattleFrames(64, engine=f.engine)
    n = f2.sample(stream, min_iterations=1 << 20, rng=np.random.default_rng(1))
    it = _host(f2, "iterations")[:n, 0]
    assert n > 0 and (it >= (1 << 20)).all()
    f.close()


def test_partial_outputs_and_bad_games(orc):
    import torch
    from oak_b200.engine import BattleEngine, OakB200Error
    from oak_b200.training_rows import EncodedBattleFrames
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    eng = BattleEngine(1)
    n = 1107
    hp = torch.full((n, 2, 6, 1), -1.0, device="cuda")
    ci = torch.zeros((n, 2, 9), dtype=torch.int64, device="cuda")
    g, r, bad = eng.rows_build_host(G["stream"], dict(hp=hp, choice_indices=ci), n)
    assert (g, r, bad) == (8, n, -1)
    assert np.array_equal(hp.cpu().numpy(), G["hp"]) and np.array_equal(ci.cpu().numpy(), G["choice_indices"])
    # capacity
    with pytest.raises(OakB200Error):
        eng.rows_build_host(G["stream"], dict(hp=hp), n - 1)
    # a game whose stored result is not what its replay ends on
    stream = G["stream"].copy()
    second = int.from_bytes(stream[0:4].tobytes(), "little")
    stream[second + 390] ^= 3
    g, r, bad = eng.rows_build_host(stream, dict(hp=hp), n)
    assert bad == 1
    with pytest.raises(ValueError):
        EncodedBattleFrames.from_bytes(stream.tobytes(), engine=eng)
    # a corrupt update header that claims more bytes than its game has: reported, and nothing is read past the game
    from oak_ref import parse_battle_data
    stream = G["stream"].copy()
    games = parse_battle_data(stream)
    last = games[-1]["updates"][-1]
    tail = 11 + 4 * (last["m"] + last["n"])
    assert last["m"] + last["n"] < 18
    stream[stream.size - tail] = 0x88  # m = n = 9
    g, r, bad = eng.rows_build_host(stream, dict(hp=hp), n)
    assert bad == len(games) - 1
    # truncated stream / empty stream
    with pytest.raises(OakB200Error):
        eng.rows_build_host(G["stream"][:-5], dict(hp=hp), n)
    g, r, bad = eng.rows_build_host(np.zeros(0, np.uint8), dict(hp=hp), n)
    assert (g, r, bad) == (0, 0, -1)
    eng.close()


def test_round_trip_from_gpu_self_play(orc):
    """Frames recorded by the GPU self-play loop replay into rows: same rows as the oracle builds from that stream, and
    the row targets are the recorded ones."""
    import json
    from oak_ref import make_net_blob, parse_battle_data
    from oak_b200.selfplay import SelfPlay
    from oak_b200.training_rows import EncodedBattleFrames
    pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
    blob = make_net_blob(np.random.default_rng(7), scale=3.0)
    sp = SelfPlay(512, blob, pool, seed=11, max_updates=400)
    chunks = []
    for _ in range(40):
        for _ in range(6):
            sp.step()
        s, g = sp.collect()
        if g:
            chunks.append(np.array(s, copy=True))
    sp.close()
    stream = np.concatenate(chunks)
    games = parse_battle_data(stream)
    assert len(games) > 100
    f = EncodedBattleFrames.from_bytes(stream.tobytes())
    assert f.rows == sum(len(g["updates"]) for g in games) == count_rows(stream)
    _check(f, oracle_training_rows(orc, stream), f.rows)
    k = _host(f, "k")[:f.rows, :, 0]
    ch = _host(f, "choice")[:f.rows, :, 0]
    at = 0
    for g in games:
        for u in g["updates"]:
            assert (k[at, 0], k[at, 1], ch[at, 0], ch[at, 1]) == (u["m"], u["n"], u["c1"], u["c2"])
            at += 1
    ci = _host(f, "choice_indices")[:f.rows]
    assert ((ci < 315).sum(-1) == k).all()  # exactly k real policy indices per side
    f.close()
