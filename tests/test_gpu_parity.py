"""GPU parity: the CUDA path (through the C ABI, include/oakb200.h) against the oracle on the same
seeded inputs — bit-exact battle[384], durations[8], actions[16], results, legal choices, playout
scores/turns/steps. Run on the B200 box: python -m pytest tests -m gpu."""
import numpy as np
import pytest

from oak_ref import (Oracle, oracle_rollout_batch, oracle_sample_batch, pick_word, sample_seed, sample_seeds, sample_teams,
                     splitmix64)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orc():
    return Oracle()


@pytest.fixture(scope="module")
def eng():
    import oak_b200
    e = oak_b200.BattleEngine(1 << 20)
    yield e
    e.close()


def first_diff(a, b):
    bad = np.nonzero((a != b).reshape(a.shape[0], -1).any(axis=1))[0]
    return bad[:5].tolist()


def test_native_library_is_loaded(eng):
    # the product path is the CUDA library: it is loaded from the tree and a call through the C ABI launches a kernel
    import oak_b200
    lib = oak_b200.load_library()
    assert lib is not None and "liboakb200" in lib._name
    before = eng.launch_count
    eng.build_battles(sample_teams(0, 128), sample_seeds(0, 128))
    assert eng.launch_count > before


def test_build_battles_matches_reference_construction(orc, eng):
    n = 4096
    b, d, r = oracle_sample_batch(orc, 0, n)
    eng.build_battles(sample_teams(0, n), sample_seeds(0, n))
    gb, gd, gr = eng.download()
    assert np.array_equal(gb, b), first_diff(gb, b)
    assert np.array_equal(gd, d)
    assert np.array_equal(gr, r)


def test_lockstep_choices_and_update_bit_exact(orc, eng):
    """choices + update per step for a whole batch, every step compared (config C2 at small n)."""
    n, steps = 1024, 80
    b, d, r = oracle_sample_batch(orc, 0, n)
    eng.upload(b, d, r)
    seeds = sample_seeds(0, n)
    for step in range(steps):
        ch, cnt = eng.choices()
        c1 = np.zeros(n, np.uint8)
        c2 = np.zeros(n, np.uint8)
        act_ref = np.zeros((n, 16), np.uint8)
        for i in range(n):
            res = int(r[i])
            if res & 15:
                continue
            o1 = orc.choices(b[i], 0, (res >> 4) & 3)
            o2 = orc.choices(b[i], 1, (res >> 6) & 3)
            assert ch[i, 0, :cnt[i, 0]].tolist() == o1 and ch[i, 1, :cnt[i, 1]].tolist() == o2, (step, i)
            pick = pick_word(int(seeds[i]), step)
            c1[i] = o1[pick % len(o1)]
            c2[i] = o2[(pick >> 32) % len(o2)]
            rr, act = orc.update(b[i], int(c1[i]), int(c2[i]), d[i])
            r[i] = rr
            act_ref[i] = act
        live = (eng.download()[2] & 15) == 0
        gr, gact = eng.update(c1, c2, want_actions=True)
        gb, gd, gr2 = eng.download()
        assert np.array_equal(gr, r), (step, first_diff(gr, r))
        assert np.array_equal(gb, b), (step, first_diff(gb, b))
        assert np.array_equal(gd, d), (step, first_diff(gd, d))
        assert np.array_equal(gact[live], act_ref[live]), (step, first_diff(gact[live], act_ref[live]))


def test_step_random_device_matches_oracle_stream(orc, eng):
    import torch
    n, steps = 2048, 40
    b, d, r = oracle_sample_batch(orc, 5000, n)
    seeds = sample_seeds(5000, n)
    dev = torch.device("cuda", eng.device)
    tb, td, tr = (torch.from_numpy(x).to(dev) for x in (b, d, r))
    ts = torch.from_numpy(seeds.view(np.int64)).to(dev)
    chosen = torch.zeros((n, 2), dtype=torch.uint8, device=dev)
    for step in range(steps):
        eng.step_random_device(tb, td, tr, ts, step, chosen_out=chosen)
        eng.synchronize()
        for i in range(n):
            if int(r[i]) & 15:
                continue
            rr, _, c1, c2 = orc.step(b[i], d[i], int(r[i]), pick_word(int(seeds[i]), step))
            r[i] = rr
        assert np.array_equal(tr.cpu().numpy(), r), step
        assert np.array_equal(tb.cpu().numpy(), b), (step, first_diff(tb.cpu().numpy(), b))
        assert np.array_equal(td.cpu().numpy(), d), step


@pytest.mark.parametrize("track_chance", [True, False])
def test_rollouts_bit_exact_final_state(orc, eng, track_chance):
    n = 1 << 15
    b, d, r = oracle_sample_batch(orc, 0, n)
    seeds = sample_seeds(0, n)
    eng.upload(b, d, r)
    score, turns, steps = eng.rollout(seeds, write_state=True, track_chance=track_chance)
    gb, gd, gr = eng.download()
    os_, ot, ost = oracle_rollout_batch(orc, b, d, r, seeds)
    assert np.array_equal(score, os_) and np.array_equal(turns, ot) and np.array_equal(steps, ost)
    assert np.array_equal(gr, r)
    assert np.array_equal(gb, b), first_diff(gb, b)
    if track_chance:
        assert np.array_equal(gd, d)


def test_rollout_leaves_resident_state_untouched_without_write_flag(orc, eng):
    n = 4096
    b, d, r = oracle_sample_batch(orc, 100, n)
    eng.upload(b, d, r)
    eng.rollout(sample_seeds(100, n))
    gb, gd, gr = eng.download()
    assert np.array_equal(gb, b) and np.array_equal(gd, d) and np.array_equal(gr, r)


def random_team(rng):
    species = rng.integers(1, 152, 6).astype(np.uint8)
    moves = np.zeros(24, np.uint8)
    for k in range(6):
        moves[4 * k:4 * k + 4] = rng.choice(np.arange(1, 165), 4, replace=False)
    n = int(rng.integers(1, 7))
    species[n:] = 0
    moves[4 * n:] = 0
    return species, moves


def test_random_teams_all_moves_rollouts(orc, eng):
    """Fuzz-style coverage of every move effect (random species, random moves, ragged team sizes)."""
    rng = np.random.default_rng(7)
    n = 8192
    b = np.zeros((n, 384), np.uint8)
    d = np.zeros((n, 8), np.uint8)
    r = np.zeros(n, np.uint8)
    seeds = np.zeros(n, np.uint64)
    for i in range(n):
        s1, m1 = random_team(rng)
        s2, m2 = random_team(rng)
        seeds[i] = splitmix64(i)
        b[i] = orc.battle_init(s1, m1, s2, m2, int(seeds[i]) ^ 0x5555)
        r[i], _ = orc.update(b[i], 0, 0, d[i])
    eng.upload(b, d, r)
    score, turns, steps = eng.rollout(seeds, max_steps=3000, write_state=True, track_chance=True)
    gb, gd, gr = eng.download()
    os_, ot, ost = oracle_rollout_batch(orc, b, d, r, seeds, max_steps=3000)
    assert np.array_equal(steps, ost), first_diff(steps, ost)
    assert np.array_equal(score, os_) and np.array_equal(turns, ot)
    assert np.array_equal(gb, b), first_diff(gb, b)
    assert np.array_equal(gd, d) and np.array_equal(gr, r)


def test_calc_overrides_clamped_damage_rolls(orc, eng):
    """Oak's tree steps force the damage roll through calc overrides (mcts.h:569-604)."""
    n, steps = 512, 30
    b, d, r = oracle_sample_batch(orc, 777, n)
    seeds = sample_seeds(777, n)
    eng.upload(b, d, r)
    for step in range(steps):
        ov = np.zeros((n, 16), np.uint8)
        ov[:, 0] = 217 + 19 * (b[:, 382] % 3)
        ov[:, 8] = 217 + 19 * (b[:, 383] % 3)
        c1 = np.zeros(n, np.uint8)
        c2 = np.zeros(n, np.uint8)
        act_ref = np.zeros((n, 16), np.uint8)
        live = (r & 15) == 0
        for i in np.nonzero(live)[0]:
            o1 = orc.choices(b[i], 0, (int(r[i]) >> 4) & 3)
            o2 = orc.choices(b[i], 1, (int(r[i]) >> 6) & 3)
            pick = pick_word(int(seeds[i]), step)
            c1[i], c2[i] = o1[pick % len(o1)], o2[(pick >> 32) % len(o2)]
            r[i], act_ref[i] = orc.update(b[i], int(c1[i]), int(c2[i]), d[i], ov[i])
        gr, gact = eng.update(c1, c2, overrides=ov, want_actions=True)
        gb, gd, _ = eng.download()
        assert np.array_equal(gr, r) and np.array_equal(gb, b) and np.array_equal(gd, d), step
        assert np.array_equal(gact[live], act_ref[live]), step


def test_randomize_hidden_variables(orc, eng):
    import torch
    n = 4096
    b, d, r = oracle_sample_batch(orc, 0, n)
    seeds = sample_seeds(0, n)
    # advance a varying number of steps so that sleep/confusion/disable/binding durations are live
    eng.upload(b, d, r)
    dev = torch.device("cuda", eng.device)
    tb, td, tr = (torch.from_numpy(x).to(dev) for x in (b, d, r))
    ts = torch.from_numpy(seeds.view(np.int64)).to(dev)
    for step in range(25):
        eng.step_random_device(tb, td, tr, ts, step)
    eng.synchronize()
    b, d = tb.cpu().numpy().copy(), td.cpu().numpy().copy()
    assert (d != 0).any()
    # fresh per-iteration seed like mcts.h:255-257
    b[:, 376:384] = np.random.default_rng(1).integers(0, 256, (n, 8), dtype=np.uint8)
    tb = torch.from_numpy(b).to(dev)
    eng.randomize_hidden_device(tb, td)
    eng.synchronize()
    for i in range(n):
        orc.randomize_hidden(b[i], d[i])
    assert np.array_equal(tb.cpu().numpy(), b), first_diff(tb.cpu().numpy(), b)


def test_full_size_checksums_one_million_playouts(orc, eng):
    """Config C2 at full size: per-battle (score, turns, steps) equal and the libpkmn-style benchmark
    checksum (sum of turns, XOR of final RNG words; engine/docs/TESTING.md:236-249)."""
    n = 1 << 20
    eng.build_battles(sample_teams(0, n), sample_seeds(0, n))
    seeds = sample_seeds(0, n)
    score, turns, steps = eng.rollout(seeds, write_state=True, track_chance=False)
    gb, _, gr = eng.download()
    b, d, r = oracle_sample_batch(orc, 0, n)
    os_, ot, ost = oracle_rollout_batch(orc, b, d, r, seeds)
    assert int(turns.astype(np.uint64).sum()) == int(ot.astype(np.uint64).sum())
    assert np.array_equal(steps, ost) and np.array_equal(score, os_) and np.array_equal(turns, ot)
    x_gpu = np.bitwise_xor.reduce(gb[:, 376:384].copy().view(np.uint64).ravel())
    x_cpu = np.bitwise_xor.reduce(b[:, 376:384].copy().view(np.uint64).ravel())
    assert x_gpu == x_cpu
    assert np.array_equal(gb, b) and np.array_equal(gr, r)


def test_pipelined_host_rollout_equals_resident_rollout(orc):
    """oakb200_rollout_host (chunked upload overlapped with the playouts) returns exactly what upload + rollout returns,
    including a batch that does not divide into chunks."""
    import oak_b200
    n = 300_001
    b, d, r = oracle_sample_batch(orc, 0, n)
    seeds = sample_seeds(0, n)
    eng = oak_b200.BattleEngine(n)
    eng.upload(b, d, r)
    want = eng.rollout(seeds)
    got = eng.rollout_host(b, d, r, seeds)
    for a, w in zip(got, want):
        assert np.array_equal(a, w)
    got = eng.rollout_host(b[:1000], None, None, seeds[:1000])
    for a, w in zip(got, want):
        assert np.array_equal(a, w[:1000])
    # chunk sizes that are not a multiple of 128 records: the library rounds them, and the records that share a
    # sector with the previous chunk's tail still come out right
    import os
    for chunk in ("9999", "10001", "333"):
        os.environ["OAKB200_PIPE_CHUNK"] = chunk
        try:
            got = eng.rollout_host(b, d, r, seeds)
        finally:
            del os.environ["OAKB200_PIPE_CHUNK"]
        for a, w in zip(got, want):
            assert np.array_equal(a, w), chunk
    eng.close()


def test_pipelined_host_rollout_beyond_32_default_chunks(orc):
    """n > 2^21: the chunk becomes ceil(n / 32) (not a multiple of 128 before the fix). Inputs are the first 2^16
    sample battles tiled, so the expected outputs are those of one resident rollout, tiled."""
    import oak_b200
    base = 1 << 16
    n = (1 << 21) + 12_345
    b0, d0, r0 = oracle_sample_batch(orc, 0, base)
    s0 = sample_seeds(0, base)
    eng = oak_b200.BattleEngine(n)
    eng.upload(b0, d0, r0)
    want = eng.rollout(s0)
    idx = np.arange(n) % base
    got = eng.rollout_host(b0[idx], d0[idx], r0[idx], s0[idx])
    for a, w in zip(got, want):
        assert np.array_equal(a, w[idx])
    eng.close()


def test_records_aligned_to_16_bytes_only(orc):
    """The one-update kernels and the evaluation read a 32-byte-aligned record with 256-bit accesses and any other
    16-byte-aligned one with 128-bit accesses (include/oakb200.h asks for 16): both forms give the same bytes."""
    import torch
    import oak_b200
    n = 4096
    dev = torch.device("cuda:0")
    e = oak_b200.BattleEngine(n)
    b, d, r = oracle_sample_batch(orc, 0, n)
    seeds = torch.from_numpy(sample_seeds(0, n).view(np.int64)).to(dev)
    raw = torch.zeros(n * 384 + 64, dtype=torch.uint8, device=dev)
    outs = []
    for shift in (0, 16):  # torch allocations are 512-byte aligned: shift 16 leaves every record on 16 (mod 32)
        tb = raw[shift:shift + n * 384].view(n, 384)
        assert tb.data_ptr() % 32 == shift
        tb.copy_(torch.from_numpy(b).to(dev))
        td = torch.from_numpy(d).to(dev)
        tr = torch.from_numpy(r).to(dev)
        for s in range(24):
            e.step_random_device(tb, td, tr, seeds, s)
        torch.cuda.synchronize()
        outs.append((tb.cpu().numpy().copy(), td.cpu().numpy(), tr.cpu().numpy()))
    for x, y in zip(outs[0], outs[1]):
        assert np.array_equal(x, y)
    ob, od, orr = b.copy(), d.copy(), r.copy()  # and both equal the oracle's 24 playout steps
    sd = sample_seeds(0, n)
    for i in range(0, n, 64):
        for s in range(24):
            if (orr[i] & 15) == 0:
                orr[i], _, _, _ = orc.step(ob[i], od[i], int(orr[i]), pick_word(int(sd[i]), s))
        assert np.array_equal(outs[0][0][i], ob[i]) and outs[0][2][i] == orr[i], first_diff(outs[0][0][i:i + 1], ob[i:i + 1])
    e.close()
