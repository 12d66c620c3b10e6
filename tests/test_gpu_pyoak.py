"""oak_b200.pyoak — the reference's `oak` module names (cpp/src/pyoak.cc:428-716) served by the GPU library: uncompressed
frames against the reference's own Frames::from_bytes (golden) and the oracle, `sample` over files, `cpp_inference`
against the oracle network, `update` and `search` on a single Input."""
import os

import numpy as np
import pytest

from oak_ref import (ROOT, ROW_FIELDS, NetOracle, Oracle, make_net_blob, oracle_battle_frames, oracle_training_rows,
                     play_frame_games, split_games)

pytestmark = pytest.mark.gpu
TARGETS = ("k", "choice", "iterations", "empirical_policies", "nash_policies", "empirical_value", "nash_value", "score")


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def _host(f, name):
    a = getattr(f, name).cpu().numpy()
    return a.view(np.uint32) if name == "iterations" else a


def test_battle_frames_match_the_reference_golden(orc):
    import oak_b200.pyoak as oak
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    F = np.load(os.path.join(ROOT, "tests", "golden", "battle_frames_golden.npz"))
    f = oak.BattleFrames.from_bytes(G["stream"].tobytes())
    assert f.rows == 1107 and f.games == 8 and not hasattr(f, "pokemon")
    for name in ("battle", "durations", "result", "choices"):  # the reference's own Frames::from_bytes
        got = _host(f, name)
        assert got.shape == F[name].shape and got.dtype == F[name].dtype and np.array_equal(got, F[name]), name
    for name in TARGETS:
        assert np.array_equal(_host(f, name), G[name]), name
    assert np.array_equal(_host(f, "legal_choices"), oracle_battle_frames(orc, G["stream"])["legal_choices"])
    f.close()


def test_battle_frames_many_games_match_the_oracle(orc):
    import oak_b200.pyoak as oak
    stream = play_frame_games(orc, first=910_000, games=400, seed=21)
    want, rows = oracle_battle_frames(orc, stream), oracle_training_rows(orc, stream)
    n = want["battle"].shape[0]
    f = oak.BattleFrames(n + 3)
    for name in ("battle", "legal_choices"):
        getattr(f, name).fill_(7)  # nothing relies on a cleared output
    assert f.fill(stream) == n
    for name in ("battle", "durations", "result", "legal_choices"):
        assert np.array_equal(_host(f, name)[:n], want[name]), name
    for name in TARGETS:
        assert np.array_equal(_host(f, name)[:n], rows[name]), name
    assert not _host(f, "choices").any()
    f.close()


def test_sample_over_files_draws_valid_rows(orc, tmp_path):
    import oak_b200.pyoak as oak
    files, table, its = [], {}, []
    for j, (first, games) in enumerate(((11_000, 40), (12_000, 25))):
        stream = play_frame_games(orc, first=first, games=games, seed=50 + j)
        p = tmp_path / f"{j}.battle.data"
        stream.tofile(p)
        files.append(str(p))
        R = oracle_training_rows(orc, stream)
        its.append(R["iterations"].reshape(-1))
        for i in range(R["k"].shape[0]):
            table[R["pokemon"][i].tobytes() + R["active"][i].tobytes() + R["iterations"][i].tobytes() + R["choice"][i].tobytes()] = \
                {name: R[name][i] for name in ROW_FIELDS}
    ix = oak.SampleIndexer()
    for p in files:
        ix.get(p)
    enc = oak.EncodedBattleFrames(96)
    min_it = int(np.median(np.concatenate(its)))  # half of the stored updates qualify
    got = oak.sample(enc, ix, threads=4, max_battle_length=10_000, min_iterations=min_it, rng=np.random.default_rng(8))
    assert got == 96
    H = {name: _host(enc, name) for name in ROW_FIELDS}
    distinct = set()
    for i in range(96):
        key = H["pokemon"][i].tobytes() + H["active"][i].tobytes() + H["iterations"][i].tobytes() + H["choice"][i].tobytes()
        assert key in table, i  # every sampled row is a row the oracle builds from one of the files' games
        for name in ROW_FIELDS:
            assert np.array_equal(H[name][i], table[key][name]), (i, name)
        assert int(H["iterations"][i, 0]) >= min_it
        distinct.add(key)
    assert len(distinct) > 48
    # nothing qualifies -> 0, as the reference's error return
    assert oak.sample(enc, ix, 1, max_battle_length=0, min_iterations=0) == 0
    assert oak.sample(oak.EncodedBattleFrames(4, engine=enc.engine), oak.SampleIndexer(), 1, 100, 0) == 0
    enc.close()


def test_cpp_inference_matches_the_oracle_network(orc, tmp_path):
    import oak_b200.pyoak as oak
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    F = np.load(os.path.join(ROOT, "tests", "golden", "battle_frames_golden.npz"))
    blob = make_net_blob(np.random.default_rng(99), activation=1, scale=2.0)
    path = tmp_path / "net.battle.net"
    np.asarray(blob, np.uint8).tofile(path)
    frames = oak.BattleFrames.from_bytes(G["stream"].tobytes())
    out = oak.cpp_inference(frames, str(path), False, "0")
    ov, olg, ocnt = NetOracle(orc, blob).eval_batch(np.ascontiguousarray(F["battle"]), np.ascontiguousarray(F["durations"]),
                                                    np.ascontiguousarray(F["result"].reshape(-1)))
    assert out.value.shape == (1107, 1) and out.policy_logit.shape == (1107, 2, 9)
    np.testing.assert_allclose(out.value[:, 0], ov, rtol=1e-4, atol=2e-5)
    mask = np.arange(9)[None, None, :] < ocnt[:, :, None]
    np.testing.assert_allclose(out.policy_logit[mask], olg[mask], rtol=1e-4, atol=2e-5)
    assert not out.policy_logit[~mask].any() and not out.policy[~mask].any()
    e = np.where(mask, np.exp(olg - np.where(mask, olg, -np.inf).max(-1, keepdims=True)), 0.0)
    np.testing.assert_allclose(out.policy, e / e.sum(-1, keepdims=True), rtol=1e-4, atol=1e-5)
    assert not out.pokemon.any() and np.isneginf(out.logit[:, :, -1]).all()
    with pytest.raises(RuntimeError):
        oak.cpp_inference(frames, str(path), False, "soon")
    frames.close()


def test_update_and_search_on_one_input(orc, tmp_path):
    import oak_b200.pyoak as oak
    ob, od, r = orc.sample_battle(3)
    inp = oak.Input(ob.copy(), od.copy(), r)
    for step in range(60):
        if r & 15:
            break
        c = [orc.choices(ob, p, (r >> (4 + 2 * p)) & 3) for p in (0, 1)]
        c1, c2 = c[0][step % len(c[0])], c[1][(step * 5) % len(c[1])]
        oak.update(inp, c1, c2)
        r, _ = orc.update(ob, c1, c2, od)
        assert inp.result == r and np.array_equal(inp.battle, ob) and np.array_equal(inp.durations, od)
    ob, od, r = orc.sample_battle(9)
    inp = oak.Input(ob, od, r)
    agent = oak.Agent(budget="512", bandit="ucb-1.0", eval="mc")
    heap = oak.Heap()
    out = oak.search(inp, heap, agent, seed=4)
    assert out.iterations == 512 and int(out.visit_matrix.sum()) == 512
    m, n = out.k
    assert out.visit_matrix[m:, :].sum() == 0 and out.visit_matrix[:, n:].sum() == 0
    for v in (out.p1_empirical, out.p2_empirical, out.p1_nash, out.p2_nash):
        assert abs(v.sum() - 1.0) < 1e-3 and (v >= 0).all()
    assert 0.0 <= out.empirical_value <= 1.0 and 0.0 <= out.nash_value <= 1.0
    assert np.array_equal(out.choices[0, :m], orc.choices(ob, 0, (r >> 4) & 3))
    for bad in (oak.Agent(budget="1s"), oak.Agent(table=True), oak.Agent(matrix_ucb="2-1.0")):
        with pytest.raises(NotImplementedError):
            oak.search(inp, heap, bad)
    for bad in (oak.Agent(bandit="thompson-1"), oak.Agent(bandit="pucb-1.0", eval="mc")):
        with pytest.raises(RuntimeError):
            oak.search(inp, heap, bad)
    # the other bandit strings of cpp/src/search.cc:236-300; the contextual ones report the root's network priors
    net_path = str(tmp_path / "net.battle.net")
    make_net_blob(np.random.default_rng(5)).tofile(net_path)
    for bandit, ev in (("ucb1-2.0", "mc"), ("exp3-0.1", "mc"), ("exp3-0.1-0.02", net_path), ("pucb-1.5", net_path), ("pexp3-0.1", net_path)):
        out = oak.search(inp, heap, oak.Agent(budget="300", bandit=bandit, eval=ev), seed=11)
        assert out.iterations == 300 and int(out.visit_matrix.sum()) == 300, bandit
        assert abs(out.p1_empirical.sum() - 1.0) < 1e-3 and abs(out.p1_nash.sum() - 1.0) < 1e-3, bandit
        if bandit.startswith("p"):
            assert abs(out.p1_prior[:m].sum() - 1.0) < 1e-6 and abs(out.p2_prior[:n].sum() - 1.0) < 1e-6 and not out.p1_prior[m:].any()
            assert 0.0 < out.initial_value < 1.0
        else:
            assert not out.p1_prior.any()
    heap._engine.close()
