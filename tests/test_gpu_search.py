"""Batched-leaf tree search (BASELINE configs[3]) through the C ABI.
  * particles = 1: visit and value matrices (and node counts) are IDENTICAL to the oracle's sequential restatement of the
    reference search (oracle/mcts_oracle.hpp) for the same random streams — every selection, transition, observation
    key and playout matched;
  * particles > 1 (virtual visits): conservation + agreement of the root value with the sequential search;
  * two of the reference's own statistical known-answer positions (cpp/src/search-test.cc:91-109, tolerance 0.03 there)."""
import ctypes
import json
import os

import numpy as np
import pytest

from oak_ref import ROOT, Oracle, oracle_sample_batch, pick_word, sample_seeds

pytestmark = pytest.mark.gpu
NAMES = json.load(open(os.path.join(ROOT, "tests", "golden", "gen1_names.json")))


def oracle_search(orc, b, d, r, seed, iters, c, rr, orl):
    f = orc.lib.oracle_mcts_run
    f.restype = None
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_float, ctypes.c_uint32,
                  ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    visit, value, nodes = np.zeros(81, np.uint32), np.zeros(81, np.float64), np.zeros(1, np.uint64)
    f(b.ctypes.data, d.ctypes.data, int(r), int(seed), iters, c, rr, orl, visit.ctypes.data, value.ctypes.data, nodes.ctypes.data)
    return visit.reshape(9, 9), value.reshape(9, 9), int(nodes[0])


def midgame(orc, n, first=0, lo=0, hi=60):
    b, d, r = oracle_sample_batch(orc, first, n)
    seeds = sample_seeds(first, n)
    stops = np.random.default_rng(3).integers(lo, hi, n)
    for i in range(n):
        for step in range(int(stops[i])):
            r2, _, _, _ = orc.step(b[i].copy(), d[i].copy(), int(r[i]), pick_word(int(seeds[i]), step))
            if r2 & 15:
                break  # keep the root non-terminal
            r[i], _, _, _ = orc.step(b[i], d[i], int(r[i]), pick_word(int(seeds[i]), step))
    return b, d, r


@pytest.mark.parametrize("rolls", [(3, 1), (39, 39), (20, 2)])
def test_single_particle_search_equals_sequential_oracle(rolls):
    import oak_b200
    from oak_b200.search import TreeSearch
    orc = Oracle()
    n, iters = 6, 400
    b, d, r = midgame(orc, n, first=100)
    seeds = np.arange(1, n + 1, dtype=np.uint64) * np.uint64(0x1234567)
    eng = oak_b200.BattleEngine(64)
    ts = TreeSearch(eng, n, particles=1)
    ts.set_roots(b, d, r, seeds)
    ts.run(iters // 2, c=1.0, root_rolls=rolls[0], other_rolls=rolls[1])
    ts.run(iters - iters // 2, c=1.0, root_rolls=rolls[0], other_rolls=rolls[1])  # trees persist across calls
    out = ts.output()
    for i in range(n):
        visit, value, nodes = oracle_search(orc, b[i], d[i], r[i], seeds[i], iters, 1.0, rolls[0], rolls[1])
        assert np.array_equal(out["visit_matrix"][i], visit), i
        assert np.array_equal(out["value_matrix"][i], value), i
        assert out["nodes"][i] == nodes and out["iterations"][i] == iters
        assert abs(out["empirical_value"][i] - value.sum() / iters) < 1e-12
    ts.close()
    eng.close()


def test_batched_particles_conserve_and_agree():
    import oak_b200
    from oak_b200.search import TreeSearch
    orc = Oracle()
    n, iters, P = 48, 1024, 32
    b, d, r = midgame(orc, n, first=500)
    seeds = sample_seeds(7000, n)
    eng = oak_b200.BattleEngine(64)
    seq = TreeSearch(eng, n, particles=1)
    seq.set_roots(b, d, r, seeds)
    seq.run(iters)
    a = seq.output()
    seq.close()
    par = TreeSearch(eng, n, particles=P)
    par.set_roots(b, d, r, seeds)
    par.run(iters)
    o = par.output()
    waves, levels = par.stats()
    par.close()
    eng.close()
    assert waves == iters // P and levels > waves
    assert np.all(o["iterations"] == iters) and np.all(o["visit_matrix"].reshape(n, -1).sum(1) == iters)
    assert np.all((o["value_matrix"] >= 0) & (o["value_matrix"] <= o["visit_matrix"]))
    np.testing.assert_allclose(o["empirical"].sum(-1), 1.0, atol=1e-9)
    # the same position searched sequentially and in batches of 32 descents: same value estimate up to search noise
    diff = np.abs(o["empirical_value"] - a["empirical_value"])
    assert diff.mean() < 0.04 and diff.max() < 0.15, (diff.mean(), diff.max())


def kat_position(orc, p1, p2):
    """One Pokémon per side: (species, move, hp, sleeps or None) as cpp/include/util/parse.h builds `name move Nhp slpK`."""
    sp = lambda name: NAMES["species"].index(name)
    mv = lambda name: NAMES["moves"].index(name)
    sets = []
    for name, move, _, _ in (p1, p2):
        sets.append((np.array([sp(name), 0, 0, 0, 0, 0], np.uint8), np.array([[mv(move), 0, 0, 0]] + [[0, 0, 0, 0]] * 5, np.uint8)))
    b = orc.battle_init(sets[0][0], sets[0][1], sets[1][0], sets[1][1], 0x123456)
    d = np.zeros(8, np.uint8)
    r, _ = orc.update(b, 0, 0, d)  # send out the leads
    for side, (_, _, hp, sleeps) in enumerate((p1, p2)):
        o = 184 * side
        b[o + 18], b[o + 19] = hp & 255, hp >> 8
        if sleeps is not None:
            b[o + 20] = 7                      # Status::Sleep7
            d[4 * side] |= (sleeps + 1) & 7    # Duration.sleeps[0] = turns slept + 1 (parse.h:79, init.h:156-165)
    return b, d, r


@pytest.mark.parametrize("slept,expected", [(3, 1 / 4), (5, 1 / 2), (6, 1.0)])
def test_reference_sleep_known_answers(slept, expected):
    """`starmie seismictoss 101hp slpK | snorlax seismictoss 1hp` -> 1 / (7 - K) (search-test.cc:91-109)."""
    import oak_b200
    from oak_b200.search import TreeSearch
    orc = Oracle()
    b, d, r = kat_position(orc, ("Starmie", "SeismicToss", 101, slept), ("Snorlax", "SeismicToss", 1, None))
    eng = oak_b200.BattleEngine(64)
    ts = TreeSearch(eng, 1, particles=64)
    ts.set_roots(b[None], d[None], np.array([r], np.uint8), np.array([42], np.uint64))
    ts.run(1 << 14)
    out = ts.output()
    ts.close()
    eng.close()
    assert abs(out["empirical_value"][0] - expected) <= (0 if expected == 1.0 else 0.03)


def oracle_search_net(orc, net, discrete, b, d, r, seed, iters, c, rr, orl):
    f = orc.lib.oracle_mcts_run_net
    f.restype = None
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint64, ctypes.c_uint32,
                  ctypes.c_float, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    visit, value, nodes = np.zeros(81, np.uint32), np.zeros(81, np.float64), np.zeros(1, np.uint64)
    f(net.h, int(discrete), b.ctypes.data, d.ctypes.data, int(r), int(seed), iters, c, rr, orl, visit.ctypes.data, value.ctypes.data,
      nodes.ctypes.data)
    return visit.reshape(9, 9), value.reshape(9, 9), int(nodes[0])


@pytest.mark.parametrize("discrete", [False, True])
def test_network_leaves_follow_the_sequential_oracle(discrete):
    """Leaf = the network's value (mcts.h:409-424) instead of a playout, one descent in flight. The leaf values are floats
    computed in a different order on the GPU (tensor cores, ~1e-6 apart from the oracle's), and UCB by design drives the
    arms' indices together, so near-ties — where such a difference decides the selection — are common: some roots follow
    the oracle selection for selection for all 300 iterations, the others part ways at a near-tie and stay statistically
    the same search. Required: at least half identical (value matrices then equal to 1e-4), every root's value within
    0.02 and root policies within 0.2 (L1) of the oracle's."""
    import oak_b200
    from oak_b200.search import TreeSearch
    from oak_ref import NetOracle, dyadic_clamp_blob, make_net_blob
    orc = Oracle()
    n, iters = 8, 300
    b, d, r = midgame(orc, n, first=900)
    seeds = np.arange(1, n + 1, dtype=np.uint64) * np.uint64(0x7654321)
    blob = dyadic_clamp_blob(np.random.default_rng(4)) if discrete else make_net_blob(np.random.default_rng(4), scale=3.0)
    net = NetOracle(orc, blob)
    eng = oak_b200.BattleEngine(64)
    eng.load_network(blob, discrete=discrete)
    ts = TreeSearch(eng, n, particles=1)
    ts.set_leaf(1)
    ts.set_roots(b, d, r, seeds)
    ts.run(iters, c=1.0, root_rolls=3, other_rolls=1)
    out = ts.output()
    same = 0
    for i in range(n):
        visit, value, nodes = oracle_search_net(orc, net, discrete, b[i], d[i], r[i], seeds[i], iters, 1.0, 3, 1)
        assert out["iterations"][i] == iters and out["visit_matrix"][i].sum() == iters
        assert abs(out["empirical_value"][i] - value.sum() / iters) < 0.02
        for axis in (0, 1):
            assert np.abs(out["visit_matrix"][i].sum(axis) - visit.sum(axis)).sum() / iters < 0.2
        if np.array_equal(out["visit_matrix"][i], visit) and out["nodes"][i] == nodes:
            same += 1
            assert np.abs(out["value_matrix"][i] - value).max() < 1e-4 * iters
    assert same >= n // 2, same
    # the values really are the network's: a search of playouts on the same roots gives different statistics
    ts.set_leaf(0)
    ts.set_roots(b, d, r, seeds)
    ts.run(iters)
    assert not np.array_equal(ts.output()["value_matrix"], out["value_matrix"])
    ts.close()
    eng.close()


def test_network_leaves_need_a_network():
    import oak_b200
    from oak_b200.engine import OakB200Error
    from oak_b200.search import TreeSearch
    eng = oak_b200.BattleEngine(8)
    ts = TreeSearch(eng, 4, particles=2)
    with pytest.raises(OakB200Error):
        ts.set_leaf(1)
    with pytest.raises(OakB200Error):
        ts.set_leaf(7)
    ts.close()
    eng.close()
