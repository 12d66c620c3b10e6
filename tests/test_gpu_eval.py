"""GPU parity of the leaf-evaluation kernels (through the C ABI) against the oracle, which is itself
pinned to the reference's encoder and torch inference (tests/test_oracle_nn.py). Both kernels are held to
1e-4 relative (north_star allows 1e-3): mode 1 = fp32 CUDA cores, mode 2 = tcgen05 tensor cores with
split-bf16 operands (measured max |error| 4e-6 on logits of rms 0.25)."""
import os

import numpy as np
import pytest

from oak_ref import ROOT, NetOracle, Oracle, make_net_blob, oracle_sample_batch, pick_word, sample_seeds

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(ROOT, "tests", "golden", "nn_golden.npz"))
RTOL, ATOL = 1e-4, 2e-5


@pytest.fixture(scope="module")
def orc():
    return Oracle()


@pytest.fixture(scope="module")
def eng():
    import oak_b200
    e = oak_b200.BattleEngine(1 << 16)
    yield e
    e.close()


def compare(v, lg, cnt, ov, olg, ocnt):
    assert np.array_equal(cnt, ocnt)
    np.testing.assert_allclose(v, ov, rtol=RTOL, atol=ATOL)
    mask = np.arange(9)[None, None, :] < ocnt[:, :, None]
    np.testing.assert_allclose(lg[mask], olg[mask], rtol=RTOL, atol=ATOL)
    assert np.all(lg[~mask] == 0)


MODES = [pytest.param(1, id="fp32"), pytest.param(2, id="tcgen05")]


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("tag", ["relu", "clamp"])
def test_eval_matches_reference_golden(orc, eng, tag, mode):
    """Directly against the outputs of the reference's own torch inference (committed golden)."""
    b, d, r = (np.ascontiguousarray(G[k]) for k in ("battles", "durations", "results"))
    eng.upload(b, d, r)
    eng.load_network(G[f"net_{tag}"])
    eng.set_eval_mode(mode)
    v, lg, cnt = eng.evaluate()
    assert np.array_equal(cnt, G["counts"])
    np.testing.assert_allclose(v, G[f"value_{tag}"], rtol=RTOL, atol=ATOL)
    mask = np.arange(9)[None, None, :] < cnt[:, :, None]
    np.testing.assert_allclose(lg[mask], G[f"logits_{tag}"][mask], rtol=2e-4, atol=5e-5)


def midgame_states(orc, n, first=0):
    b, d, r = oracle_sample_batch(orc, first, n)
    seeds = sample_seeds(first, n)
    rng = np.random.default_rng(5)
    stops = rng.integers(0, 120, n)
    for i in range(n):
        for step in range(int(stops[i])):
            if int(r[i]) & 15:
                break
            r[i], _, _, _ = orc.step(b[i], d[i], int(r[i]), pick_word(int(seeds[i]), step))
    return b, d, r


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("activation", [1, 2])
def test_eval_default_dims_midgame_states(orc, eng, activation, mode):
    """Default network dimensions (198/427 -> 128 -> 59/83, 768 -> 64 -> 64, 32, 64 -> 315), random parameters
    in the reference file format, 8192 mid-game states incl. terminal / fainted-active / forced-switch ones."""
    n = 8192
    b, d, r = midgame_states(orc, n)
    blob = make_net_blob(np.random.default_rng(1234), activation=activation, scale=2.0)
    net = NetOracle(orc, blob)
    ov, olg, ocnt = net.eval_batch(b, d, r)
    eng.upload(b, d, r)
    eng.load_network(blob)
    eng.set_eval_mode(mode)
    v, lg, cnt = eng.evaluate()
    compare(v, lg, cnt, ov, olg, ocnt)


@pytest.mark.parametrize("mode", MODES)
def test_eval_ragged_batch_and_small_dims(orc, eng, mode):
    """A batch that is not a multiple of the 128-row tile, and a narrower network (zero-padded on the tensor path)."""
    n = 1000
    b, d, r = midgame_states(orc, n, first=50000)
    blob = make_net_blob(np.random.default_rng(7), phd=96, pod=40, ahd=112, aod=60, hd=48, vhd=24, pohd=32, activation=1)
    net = NetOracle(orc, blob)
    ov, olg, ocnt = net.eval_batch(b, d, r)
    eng.upload(b, d, r)
    eng.load_network(blob)
    eng.set_eval_mode(mode)
    v, lg, cnt = eng.evaluate()
    compare(v, lg, cnt, ov, olg, ocnt)


def test_eval_duplicate_active_moves(orc, eng):
    """Mimic can leave the same move twice in the active move slots; the reference's sparse encoder lists the
    column twice (battle.h:381-409), so the one-hot tile must carry 2.0 there."""
    n = 256
    b, d, r = midgame_states(orc, n, first=1000)
    b = b.copy()
    for i in range(n):
        for s in (0, 184):
            b[i, s + 144 + 24 + 2] = b[i, s + 144 + 24 + 0]  # active move slot 2 := slot 1 (id), keep its pp
            b[i, s + 144 + 24 + 3] = max(1, b[i, s + 144 + 24 + 3])
    blob = make_net_blob(np.random.default_rng(3), activation=1)
    ov, olg, ocnt = NetOracle(orc, blob).eval_batch(b, d, r)
    eng.upload(b, d, r)
    eng.load_network(blob)
    for mode in (1, 2):
        eng.set_eval_mode(mode)
        v, lg, cnt = eng.evaluate()
        compare(v, lg, cnt, ov, olg, ocnt)


def test_eval_mode_limits(eng):
    import oak_b200
    eng.load_network(make_net_blob(np.random.default_rng(0), pod=70))  # pokemon embedding wider than the tensor tile
    eng.set_eval_mode(2)
    eng.upload(np.zeros((1, 384), np.uint8))
    with pytest.raises(oak_b200.OakB200Error):
        eng.evaluate()
    eng.set_eval_mode(0)


def test_eval_rejects_bad_parameter_files(eng):
    import oak_b200
    eng.set_eval_mode(0)
    blob = make_net_blob(np.random.default_rng(0))
    with pytest.raises(oak_b200.OakB200Error):
        eng.load_network(np.concatenate([blob, np.zeros(3, np.uint8)]))
    with pytest.raises(oak_b200.OakB200Error):
        eng.load_network(blob[:1000])


def test_pipelined_host_evaluation_matches_upload_then_eval(orc, eng):
    """oakb200_eval_host (chunked upload / kernel / download on three streams) returns what oakb200_upload + oakb200_eval do,
    for a batch that is not a multiple of the chunk, with and without explicit durations / results."""
    import torch
    n = 50_000
    b, d, r = oracle_sample_batch(orc, 300_000, n)
    eng.load_network(make_net_blob(np.random.default_rng(21)))
    for mode in (2, 1):
        eng.set_eval_mode(mode)
        eng.upload(b, d, r)
        v0, l0, c0 = eng.evaluate()
        pb, pd, pr = (torch.from_numpy(x).pin_memory() for x in (b, d, r))
        v1, l1, c1 = eng.evaluate_host(pb, pd, pr)
        assert np.array_equal(v0, v1) and np.array_equal(l0, l1) and np.array_equal(c0, c1)
        v2, l2, c2 = eng.evaluate_host(b[:777], None, None)  # pageable memory, defaults
        assert np.array_equal(v2, v0[:777]) and np.array_equal(l2, l0[:777]) and np.array_equal(c2, c0[:777])
    eng.set_eval_mode(0)


def device_midgame_states(n, steps):
    """n sample battles advanced `steps` random-playout steps by the GPU step kernel (itself parity-tested against the oracle
    step for step in tests/test_gpu_parity.py): the bench's C3 input, cheap to make at full size."""
    import torch
    import oak_b200
    from oak_ref import sample_teams
    e = oak_b200.BattleEngine(n)
    seeds = sample_seeds(0, n)
    e.build_battles(sample_teams(0, n), seeds)
    b, d, r = (torch.from_numpy(x).cuda() for x in e.download())
    s = torch.from_numpy(seeds.view(np.int64)).cuda()
    for k in range(steps):
        e.step_random_device(b, d, r, s, k)
    e.synchronize()
    out = tuple(x.cpu().numpy() for x in (b, d, r))
    e.close()
    return out


def test_config_c3_full_batch_65536(orc, eng):
    """BASELINE configs[2] at its full size, the batch bench.py times: 65 536 positions = 512 tiles of 128, 3.46 rounds over
    148 persistent CTAs (the uneven last round and the head warps' one-tile-behind hand-over at the end of each CTA's list),
    through the resident path and through oakb200_eval_host (chunked upload / evaluate / download)."""
    n = 1 << 16
    b, d, r = device_midgame_states(n, 19)
    net = NetOracle(orc, make_net_blob(np.random.default_rng(1234)))
    ov, olg, ocnt = net.eval_batch(b, d, r)
    eng.load_network(make_net_blob(np.random.default_rng(1234)))
    eng.set_eval_mode(2)
    eng.upload(b, d, r)
    compare(*eng.evaluate(), ov, olg, ocnt)
    compare(*eng.evaluate_host(b, d, r), ov, olg, ocnt)
    eng.set_eval_mode(0)
