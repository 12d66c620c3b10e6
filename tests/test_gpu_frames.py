"""GPU frame recorder / writer and policy sampling (through the C ABI) against the oracle, which is pinned to the
reference's own serialiser (tests/test_oracle_frames.py). Byte-exact."""
import ctypes

import numpy as np
import pytest

from oak_ref import MASK64, Oracle, frames_expected_game, parse_battle_data, random_frame_games, splitmix64

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def run_lockstep(eng, torch, G, drain_every, slots=None):
    """Drive the recorder the way a self-play loop does; -> list of (stream, games) per drain, and the drain order."""
    dev = torch.device("cuda", 0)
    games, T, L = G["battles"].shape[0], G["T"], G["L"]
    t_ = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    battles = t_(G["battles"])
    out, order = [], []
    alive = np.full(games, 0x50, np.uint8)
    eng.frames_begin_device(battles, t_(alive))
    for t in range(T):
        before = np.where(t < L, 0x50, G["final"]).astype(np.uint8)
        after = np.where(t + 1 < L, 0x50, G["final"]).astype(np.uint8)
        eng.frames_record_device(t_(before), t_(G["choice"][t, :, 0]), t_(G["choice"][t, :, 1]), t_(G["k"][t]),
                                 t_(G["ev"][t]), t_(G["nv"][t]), t_(G["emp"][t]), t_(G["nash"][t]), iterations=t_(G["iters"][t].view(np.int32)))
        eng.frames_finish_device(t_(after))
        if (t + 1) % drain_every == 0 or t + 1 == T:
            stream, n = eng.frames_drain()
            done = [g for g in range(games) if L[g] <= t + 1 and g not in order]
            assert n == len(done)
            order += done
            out.append(stream)
    return out, order


def test_recorder_stream_is_byte_identical(orc):
    import torch
    import oak_b200
    G = random_frame_games(np.random.default_rng(11), 300, 60)
    eng = oak_b200.BattleEngine(512)
    eng.frames_create(300, max_updates=64)
    streams, order = run_lockstep(eng, torch, G, drain_every=7)
    got = np.concatenate(streams)
    want = np.concatenate([frames_expected_game(orc.lib, "oracle_frames_write_game", G, g) for g in order])
    assert sorted(order) == list(range(300))
    assert np.array_equal(got, want)
    assert len(parse_battle_data(got)) == 300
    # the freed slots record a second wave of games
    G2 = random_frame_games(np.random.default_rng(12), 300, 20)
    streams, order = run_lockstep(eng, torch, G2, drain_every=1000)
    want = np.concatenate([frames_expected_game(orc.lib, "oracle_frames_write_game", G2, g) for g in range(300)])
    assert np.array_equal(np.concatenate(streams), want)
    assert eng.frames_pending() == (0, 0)
    eng.close()


def test_recorder_drops_overlong_games_and_reports_capacity(orc):
    import torch
    import oak_b200
    G = random_frame_games(np.random.default_rng(13), 64, 30)
    eng = oak_b200.BattleEngine(64)
    eng.frames_create(64, max_updates=16)
    streams, _ = None, None
    dev = torch.device("cuda", 0)
    t_ = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    eng.frames_begin_device(t_(G["battles"]), t_(np.full(64, 0x50, np.uint8)))
    for t in range(G["T"]):
        before = np.where(t < G["L"], 0x50, G["final"]).astype(np.uint8)
        after = np.where(t + 1 < G["L"], 0x50, G["final"]).astype(np.uint8)
        eng.frames_record_device(t_(before), t_(G["choice"][t, :, 0]), t_(G["choice"][t, :, 1]), t_(G["k"][t]), t_(G["ev"][t]),
                                 t_(G["nv"][t]), t_(G["emp"][t]), t_(G["nash"][t]), iterations=t_(G["iters"][t].view(np.int32)))
        eng.frames_finish_device(t_(after))
    keep = [g for g in range(64) if G["L"][g] <= 16]
    finished, dropped = eng.frames_pending()
    assert finished == len(keep) and dropped == 64 - len(keep)
    with pytest.raises(oak_b200.OakB200Error):
        eng.frames_drain(capacity=100)
    stream, n = eng.frames_drain()
    want = np.concatenate([frames_expected_game(orc.lib, "oracle_frames_write_game", G, g) for g in keep])
    assert n == len(keep) and np.array_equal(stream, want)
    eng.close()


def test_policy_sample_matches_oracle(orc):
    import torch
    import oak_b200
    f = orc.lib.oracle_policy_sample
    f.restype = ctypes.c_int
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_void_p]
    rng = np.random.default_rng(21)
    n = 4096
    counts = rng.integers(1, 10, (n, 2)).astype(np.uint8)
    choices = rng.integers(1, 256, (n, 2, 9)).astype(np.uint8)
    logits = (rng.standard_normal((n, 2, 9)) * 2).astype(np.float32)
    seeds = rng.integers(0, 1 << 63, n, dtype=np.uint64)
    dev = torch.device("cuda", 0)
    t_ = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    eng = oak_b200.BattleEngine(n)
    for temp, floor_p, step in ((1.0, 0.0, 0), (2.5, 0.0, 3), (1.0, 0.1, 17), (0.5, 0.05, 4000)):
        probs = torch.zeros((n, 2, 9), dtype=torch.float32, device=dev)
        c1 = torch.zeros(n, dtype=torch.uint8, device=dev)
        c2 = torch.zeros(n, dtype=torch.uint8, device=dev)
        eng.policy_sample_device(t_(choices), t_(counts), t_(logits), t_(seeds.view(np.int64)), step, probs, c1, c2, temp, floor_p)
        torch.cuda.synchronize()
        probs, c = probs.cpu().numpy(), np.stack([c1.cpu().numpy(), c2.cpu().numpy()], 1)
        mism = 0
        for i in range(n):
            for s in range(2):
                u = (splitmix64(int(seeds[i]) ^ ((0xD1B54A32D192ED03 * step) & MASK64) ^ s) >> 11) * 2.0 ** -53
                pr = np.zeros(9, np.float32)
                idx = f(logits[i, s].ctypes.data, int(counts[i, s]), temp, floor_p, u, pr.ctypes.data)
                np.testing.assert_allclose(probs[i, s], pr, rtol=2e-6, atol=1e-9)
                mism += int(c[i, s] != choices[i, s, idx])
        # expf / pow differ in the last ulp between libm and the device: a draw that lands on a boundary may flip
        assert mism <= 2, mism
    eng.close()
