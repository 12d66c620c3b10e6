"""Host side of oak_b200.pyoak (the reference's `oak` module names, cpp/src/pyoak.cc:43-109,394-425,675-689): stream
reader, sample index, budget strings, output buffer. No GPU."""
import os

import numpy as np
import pytest

from oak_ref import ROOT, split_games

G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))


@pytest.fixture()
def data_file(tmp_path):
    p = tmp_path / "a.battle.data"
    G["stream"].tofile(p)
    return str(p)


def test_read_battle_data_splits_games(data_file):
    import oak_b200.pyoak as oak
    games = oak.read_battle_data(data_file)
    want = split_games(G["stream"])
    assert len(games) == len(want) == 8
    for (buf, n), g in zip(games, want):
        assert isinstance(buf, bytes) and buf == bytes(g)
        assert n == int.from_bytes(buf[4:6], "little") and len(buf) == int.from_bytes(buf[:4], "little")
    assert sum(n for _, n in games) == 1107
    with pytest.raises(RuntimeError):
        oak.read_battle_data(data_file + ".missing")


def test_read_battle_data_rejects_a_torn_file(tmp_path):
    import oak_b200.pyoak as oak
    p = tmp_path / "torn.battle.data"
    G["stream"][:-5].tofile(p)
    with pytest.raises(RuntimeError):
        oak.read_battle_data(str(p))


def test_sample_indexer(data_file, tmp_path):
    import oak_b200.pyoak as oak
    ix = oak.SampleIndexer()
    assert ix.size() == 0
    lst = ix.get(data_file)
    assert ix.size() == 1 and ix.get(data_file) is lst
    off = 0
    for (o, n), g in zip(lst, split_games(G["stream"])):
        assert o == off and n == int.from_bytes(bytes(g[4:6]), "little")
        off += len(g)
    other = tmp_path / "b.battle.data"
    G["stream"].tofile(other)
    ix.get(str(other))
    assert ix.size() == 2
    ix.prune([str(other)])
    assert ix.size() == 1 and data_file not in ix.data


def test_budget_strings_and_output_buffer():
    import oak_b200.pyoak as oak
    assert oak._parse_budget("1024") == (1024, "") and oak._parse_budget("100ms") == (100, "ms") and oak._parse_budget("8s") == (8, "s")
    for bad in ("fast", "10h", ""):
        with pytest.raises(RuntimeError):
            oak._parse_budget(bad)
    o = oak.OutputBuffer(5)
    assert o.pokemon.shape == (5, 2, 5, 59) and o.active_pokemon.shape == (5, 2, 1, 83) and o.sides.shape == (5, 2, 1, 384)
    assert o.logit.shape == (5, 2, 316) and np.isneginf(o.logit[:, :, -1]).all() and not o.logit[:, :, :-1].any()
    o.value[:] = 3
    o.clear()
    assert not o.value.any() and o.policy.shape == o.policy_logit.shape == (5, 2, 9)
    assert (oak.pokemon_in_dim, oak.active_in_dim, oak.policy_out_dim) == (198, 427, 315)  # pyoak.cc:587: the active net takes 229 + 198
    p1, p2, v = oak.solve_matrix(np.array([[1.0, 0.0], [0.0, 1.0]], np.float32))
    assert abs(v - 0.5) < 1e-6 and np.allclose(p1, 0.5) and np.allclose(p2, 0.5)


def test_sample_returns_zero_on_a_corrupt_update_header(tmp_path, capsys):
    """pyoak.sample's '0 on error' contract (pyoak.cc:187-190,254-259) also holds for a record whose stored (m, n) nibbles
    walk the header scan out of the game: that is reported and 0 returned before any device work."""
    import types
    import oak_b200.pyoak as oak
    game = np.array(split_games(G["stream"])[0], copy=True)
    at = 391
    for _ in range(int.from_bytes(bytes(game[4:6]), "little") - 1):
        at += 11 + 4 * ((int(game[at]) & 15) + (int(game[at]) >> 4) + 2)
    game[at] = 0xFF  # the last update claims m = n = 15: 139 bytes, past the end of the game
    path = tmp_path / "bad.battle.data"
    game.tofile(path)
    ix = oak.SampleIndexer()
    ix.get(str(path))
    filled = []
    frames = types.SimpleNamespace(size=4, fill=lambda *a, **k: filled.append(1) or 4)
    assert oak.sample(frames, ix, rng=np.random.default_rng(1)) == 0 and not filled
    assert "does not end on its last update" in capsys.readouterr().err
