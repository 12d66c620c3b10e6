"""Self-play with a tree search per turn (generate.cc:286-333 on the GPU): the recorded stream must replay exactly on
the oracle, carry the search's outputs (iterations, empirical + Nash policies and values), and turn into training rows."""
import json
import os

import numpy as np
import pytest

from oak_ref import ROOT, Oracle, oracle_training_rows, parse_battle_data

pytestmark = pytest.mark.gpu


def test_search_selfplay_stream():
    from oak_b200.selfplay import SearchSelfPlay
    from oak_b200.training_rows import EncodedBattleFrames
    orc = Oracle()
    pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
    sp = SearchSelfPlay(48, pool, iterations=64, particles=8, policy_mode="e0.5-n0.5", seed=5, max_updates=600, min_prob=0.02)
    chunks, games = [], 0
    for s_ in range(700):
        sp.step()
        if (s_ + 1) % 10 == 0:
            st, g = sp.collect()
            if g:
                chunks.append(np.array(st, copy=True))
                games += g
        if games >= 24:
            break
    sp.close()
    assert games >= 24 and sp.searched > 0
    stream = np.concatenate(chunks)
    parsed = parse_battle_data(stream)
    assert len(parsed) == games
    n_updates = 0
    for g in parsed:
        b, d, r = g["battle"].copy(), np.zeros(8, np.uint8), 0x50
        for ui, u in enumerate(g["updates"]):
            ch1, ch2 = orc.choices(b, 0, (r >> 4) & 3), orc.choices(b, 1, (r >> 6) & 3)
            assert (u["m"], u["n"]) == (len(ch1), len(ch2)) and u["c1"] in ch1 and u["c2"] in ch2
            assert u["iterations"] == 64  # rounded up to a multiple of the 8 descents in flight
            for emp, nash, cnt in ((u["p1_emp"], u["p1_nash"], u["m"]), (u["p2_emp"], u["p2_nash"], u["n"])):
                assert abs(int(emp.astype(np.int64).sum()) - 65535) <= cnt + 1  # visit fractions, each truncated to u16
                assert abs(int(nash.astype(np.int64).sum()) - 65535) <= cnt + 1
            # the chosen action had probability >= the floor under the mixed policy
            mix = 0.5 * u["p1_emp"][ch1.index(u["c1"])] + 0.5 * u["p1_nash"][ch1.index(u["c1"])]
            assert mix / 65535.0 >= 0.02 * 0.9
            assert 0 <= u["ev"] <= 65535 and 0 <= u["nv"] <= 65535
            r, _ = orc.update(b, u["c1"], u["c2"], d)
            n_updates += 1
        assert r == g["result"] and (r & 15) != 0
    # searches see further than one random playout: decided positions get decided values near the end of a game
    last = [g["updates"][-1] for g in parsed if (g["result"] & 15) in (1, 2)]
    agree = sum(((u["ev"] > 32767) == ((g["result"] & 15) == 1)) for u, g in zip(last, [g for g in parsed if (g["result"] & 15) in (1, 2)]))
    assert agree >= 0.8 * len(last)
    # and the stream feeds the row builder
    f = EncodedBattleFrames.from_bytes(stream.tobytes())
    assert f.rows == n_updates
    want = oracle_training_rows(orc, stream)
    for name in ("pokemon", "active", "hp", "choice_indices", "empirical_policies", "nash_policies", "nash_value", "score"):
        assert np.array_equal(getattr(f, name).cpu().numpy(), want[name]), name
    f.close()


def test_search_selfplay_with_network_leaves():
    """The same loop with the network evaluating the leaves (the reference's usual generate configuration)."""
    from oak_ref import make_net_blob
    from oak_b200.selfplay import SearchSelfPlay
    orc = Oracle()
    pool = np.array(json.load(open(os.path.join(ROOT, "tests", "golden", "oak_teams.json")))["ou_sample_teams"], np.uint8)
    blob = make_net_blob(np.random.default_rng(3), scale=3.0)
    sp = SearchSelfPlay(32, pool, iterations=32, particles=8, net_blob=blob, seed=9, max_updates=700)
    assert sp.search.lib is not None
    chunks, games = [], 0
    for s_ in range(900):
        sp.step()
        if (s_ + 1) % 20 == 0:
            st, g = sp.collect()
            if g:
                chunks.append(np.array(st, copy=True))
                games += g
        if games >= 8:
            break
    sp.close()
    assert games >= 8
    for g in parse_battle_data(np.concatenate(chunks)):
        b, d, r = g["battle"].copy(), np.zeros(8, np.uint8), 0x50
        for u in g["updates"]:
            assert u["iterations"] == 32
            r, _ = orc.update(b, u["c1"], u["c2"], d)
        assert r == g["result"] and (r & 15) != 0
