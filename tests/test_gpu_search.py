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


SEARCH_KATS = [  # cpp/src/search-test.cc:50-109: (p1 = (species, move, hp, slept, conf), p2, expected value, tolerance)
    (("Starmie", "SeismicToss", 1, None, 5), ("Snorlax", "BodySlam", 1, None, None), 1.0, 0.0),
    (("Starmie", "SeismicToss", 1, None, 4), ("Snorlax", "BodySlam", 1, None, None), .5 + .5 / 2, 0.03),
    (("Starmie", "SeismicToss", 1, None, 3), ("Snorlax", "BodySlam", 1, None, None), .33 + .66 / 2, 0.03),
    (("Starmie", "SeismicToss", 1, None, 2), ("Snorlax", "BodySlam", 1, None, None), .25 + .75 / 2, 0.03),
    (("Starmie", "SeismicToss", 1, None, 1), ("Snorlax", "BodySlam", 1, None, None), .5, 0.03),
    (("Starmie", "SeismicToss", 1, 6, None), ("Snorlax", "SeismicToss", 1, None, None), 0.0, 0.0),
] + [(("Starmie", "SeismicToss", 101, k, None), ("Snorlax", "SeismicToss", 1, None, None), 1.0 / (7 - k), 0.0 if k == 6 else 0.03)
     for k in range(7)]


def kat_position(orc, p1, p2):
    """One Pokémon per side: (species, move, hp, sleeps or None[, observed confusion duration]) as cpp/include/util/parse.h
    builds `name move Nhp slpK (conf:C)`: slpK = Status::Sleep7 + Duration.sleeps[0] = K + 1 (parse.h:71-79, init.h:156-165),
    (conf:C) = the Confusion volatile + Duration.confusion = C (parse.h:190-198); the hidden counters are whatever — every
    search iteration redraws them from the durations (durations.h:25-97)."""
    p1, p2 = tuple(p1) + (None,) * (5 - len(p1)), tuple(p2) + (None,) * (5 - len(p2))
    sp = lambda name: NAMES["species"].index(name)
    mv = lambda name: NAMES["moves"].index(name)
    sets = []
    for name, move, _, _, _ in (p1, p2):
        sets.append((np.array([sp(name), 0, 0, 0, 0, 0], np.uint8), np.array([[mv(move), 0, 0, 0]] + [[0, 0, 0, 0]] * 5, np.uint8)))
    b = orc.battle_init(sets[0][0], sets[0][1], sets[1][0], sets[1][1], 0x123456)
    d = np.zeros(8, np.uint8)
    r, _ = orc.update(b, 0, 0, d)  # send out the leads
    for side, (_, _, hp, sleeps, conf) in enumerate((p1, p2)):
        o = 184 * side
        b[o + 18], b[o + 19] = hp & 255, hp >> 8
        if sleeps is not None:
            b[o + 20] = 7                      # Status::Sleep7
            d[4 * side] |= (sleeps + 1) & 7    # Duration.sleeps[0] = turns slept + 1 (parse.h:79, init.h:156-165)
        if conf is not None:
            b[o + 144 + 16] |= 1 << 7          # Volatiles.Confusion (data.zig:325-383, bit 7)
            dw = int.from_bytes(d[4 * side:4 * side + 4].tobytes(), "little") | (conf << 18)  # Duration.confusion (chance.zig:279-287)
            d[4 * side:4 * side + 4] = np.frombuffer(dw.to_bytes(4, "little"), np.uint8)
    return b, d, r


def test_reference_search_known_answers():
    """All 13 statistical known answers of the reference (cpp/src/search-test.cc:50-109: five observed-confusion positions,
    eight observed-sleep positions; tolerance 0.03, exact where the reference says error = 0) as 13 roots of one batched search."""
    import oak_b200
    from oak_b200.search import TreeSearch
    orc = Oracle()
    pos = [kat_position(orc, p1, p2) for p1, p2, _, _ in SEARCH_KATS]
    b, d = np.stack([x[0] for x in pos]), np.stack([x[1] for x in pos])
    r = np.array([x[2] for x in pos], np.uint8)
    eng = oak_b200.BattleEngine(13 * 64)
    ts = TreeSearch(eng, 13, particles=64)
    ts.set_roots(b, d, r, np.arange(42, 55, dtype=np.uint64))
    ts.run(1 << 14)
    out = ts.output()
    ts.close()
    eng.close()
    for i, (p1, p2, expected, tol) in enumerate(SEARCH_KATS):
        assert abs(out["empirical_value"][i] - expected) <= tol + 1e-12, (p1, out["empirical_value"][i], expected)


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


def test_config_c4_scale_4096_roots():
    """BASELINE configs[3] at the size bench.py runs: 4 096 roots. (a) x 64 descents in flight, 4 waves: every root's visits
    are conserved, values lie in [0, visits], policies sum to 1, and the 262 144 resident particles really are distinct
    playouts (value matrices differ between roots). (b) the same 4 096 roots with ONE descent in flight for 24 iterations:
    every root's matrices and node count equal the sequential oracle's — the root sharding over host threads and stream
    groups changes nothing at scale."""
    import oak_b200
    from oak_b200.search import TreeSearch
    from test_gpu_eval import device_midgame_states
    orc = Oracle()
    n, P = 4096, 64
    b, d, r = device_midgame_states(n, 19)
    live = (r & 15) == 0
    b, d, r = b[live], d[live], r[live]
    n = b.shape[0]
    assert n > 4000
    seeds = sample_seeds(90_000, n)
    eng = oak_b200.BattleEngine(n * P)
    ts = TreeSearch(eng, n, particles=P)
    ts.set_roots(b, d, r, seeds)
    ts.run(4 * P)
    o = ts.output()
    waves, levels = ts.stats()
    ts.close()
    assert waves == 4 and levels >= waves
    assert np.all(o["iterations"] == 4 * P) and np.all(o["visit_matrix"].reshape(n, -1).sum(1) == 4 * P)
    assert np.all((o["value_matrix"] >= 0) & (o["value_matrix"] <= o["visit_matrix"]))
    np.testing.assert_allclose(o["empirical"].sum(-1), 1.0, atol=1e-9)
    assert np.all((o["empirical_value"] >= 0) & (o["empirical_value"] <= 1)) and 0.3 < o["empirical_value"].mean() < 0.7
    assert len({o["value_matrix"][i].tobytes() for i in range(0, n, 37)}) > 100
    one = TreeSearch(eng, n, particles=1)
    one.set_roots(b, d, r, seeds)
    one.run(24, c=1.0, root_rolls=3, other_rolls=1)
    o1 = one.output()
    one.close()
    eng.close()
    for i in range(n):
        visit, value, nodes = oracle_search(orc, b[i], d[i], r[i], seeds[i], 24, 1.0, 3, 1)
        assert np.array_equal(o1["visit_matrix"][i], visit) and np.array_equal(o1["value_matrix"][i], value), i
        assert o1["nodes"][i] == nodes, i


def oracle_search_bandit(orc, net, discrete, b, d, r, seed, iters, kind, f1, f2, rr, orl):
    f = orc.lib.oracle_mcts_run_bandit
    f.restype = None
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint64, ctypes.c_uint32,
                  ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
                  ctypes.c_void_p]
    visit, value, nodes = np.zeros(81, np.uint32), np.zeros(81, np.float64), np.zeros(1, np.uint64)
    f(net.h if net is not None else None, int(discrete), b.ctypes.data, d.ctypes.data, int(r), int(seed), iters, kind, f1, f2, rr, orl,
      visit.ctypes.data, value.ctypes.data, nodes.ctypes.data)
    return visit.reshape(9, 9), value.reshape(9, 9), int(nodes[0])


@pytest.mark.parametrize("spec,kind,f1,f2", [("ucb1-2.0", 1, 2.0, 0.05), ("exp3-1.0-0.1", 3, 1.0, 0.1), ("exp3-0.3", 3, 0.3, 0.05)])
def test_other_bandits_equal_the_sequential_oracle(spec, kind, f1, f2):
    """UCB1 (ucb1.h) and Exp3 (exp3.h; the reference's default agent bandit is "exp3-1.0-0.1", search-test.cc:29) with
    MonteCarlo leaves, one descent in flight: every selection — Exp3's sampled ones included, drawn from the documented
    counter-based stream — matches the oracle, so the matrices and node counts are identical."""
    import oak_b200
    from oak_b200.search import TreeSearch
    orc = Oracle()
    n, iters = 6, 300
    b, d, r = midgame(orc, n, first=300)
    seeds = np.arange(1, n + 1, dtype=np.uint64) * np.uint64(0x2468ACE)
    eng = oak_b200.BattleEngine(64)
    ts = TreeSearch(eng, n, particles=1)
    ts.set_bandit(spec)
    ts.set_roots(b, d, r, seeds)
    ts.run(iters, root_rolls=3, other_rolls=1)
    out = ts.output()
    for i in range(n):
        visit, value, nodes = oracle_search_bandit(orc, None, False, b[i], d[i], r[i], seeds[i], iters, kind, f1, f2, 3, 1)
        assert np.array_equal(out["visit_matrix"][i], visit), i
        assert np.array_equal(out["value_matrix"][i], value), i
        assert out["nodes"][i] == nodes and out["iterations"][i] == iters
    ts.close()
    eng.close()


@pytest.mark.parametrize("spec,kind,f1,f2", [("pucb-1.5", 2, 1.5, 0.05), ("pexp3-1.0-0.1", 4, 1.0, 0.1)])
def test_contextual_bandits_use_the_network_priors(spec, kind, f1, f2):
    """PUCB / PExp3: priors = the network's policy logits over the legal choices at the root and at every new leaf
    (mcts.h:190-209,409-421; pucb.h:38-40, pexp3.h:44-48). One descent in flight against the oracle with the same network.
    Values and logits are floats computed in a different order on the GPU (~1e-6), so a near-tie can send the two searches
    apart: required as for UCB with network leaves — at least half the roots identical selection for selection, all of them
    statistically the same search — and the priors must matter: PUCB with this network differs from UCB with it."""
    import oak_b200
    from oak_b200.search import TreeSearch
    from oak_ref import NetOracle, make_net_blob
    orc = Oracle()
    n, iters = 8, 300
    b, d, r = midgame(orc, n, first=1200)
    seeds = np.arange(1, n + 1, dtype=np.uint64) * np.uint64(0x1357911)
    blob = make_net_blob(np.random.default_rng(11), scale=3.0)
    net = NetOracle(orc, blob)
    eng = oak_b200.BattleEngine(64)
    eng.load_network(blob)
    ts = TreeSearch(eng, n, particles=1)
    ts.set_leaf(1)
    ts.set_bandit(spec)
    ts.set_roots(b, d, r, seeds)
    ts.run(iters, root_rolls=3, other_rolls=1)
    out = ts.output()
    same = 0
    for i in range(n):
        visit, value, nodes = oracle_search_bandit(orc, net, False, b[i], d[i], r[i], seeds[i], iters, kind, f1, f2, 3, 1)
        assert out["iterations"][i] == iters and out["visit_matrix"][i].sum() == iters
        assert abs(out["empirical_value"][i] - value.sum() / iters) < 0.03
        if np.array_equal(out["visit_matrix"][i], visit) and out["nodes"][i] == nodes:
            same += 1
            assert np.abs(out["value_matrix"][i] - value).max() < 1e-4 * iters
    assert same >= n // 2, same
    if kind == 2:
        ts.set_bandit("ucb-1.5")
        ts.set_roots(b, d, r, seeds)
        ts.run(iters, root_rolls=3, other_rolls=1)
        assert not np.array_equal(ts.output()["visit_matrix"], out["visit_matrix"])
    ts.close()
    eng.close()


def test_iteration_budget_is_exact_and_contextual_needs_a_network():
    """`iterations` is run to the iteration, not rounded up to whole waves (the reference runs exactly its budget,
    mcts.h:237-241); pucb without network leaves is refused as the reference refuses it (search.cc:253-258)."""
    import oak_b200
    from oak_b200.engine import OakB200Error
    from oak_b200.search import TreeSearch
    orc = Oracle()
    b, d, r = midgame(orc, 5, first=40)
    eng = oak_b200.BattleEngine(5 * 64)
    ts = TreeSearch(eng, 5, particles=64)
    ts.set_roots(b, d, r, sample_seeds(1, 5))
    total = 0
    for it in (10, 1, 64, 100):
        ts.run(it)
        total += it
        o = ts.output()
        assert np.all(o["iterations"] == total) and np.all(o["visit_matrix"].reshape(5, -1).sum(1) == total)
    ts.set_bandit("pucb-1.0")
    with pytest.raises(OakB200Error):
        ts.run(8)
    with pytest.raises(ValueError):
        ts.set_bandit("thompson-1")
    ts.close()
    eng.close()
