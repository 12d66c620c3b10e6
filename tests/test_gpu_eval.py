"""GPU parity of the leaf-evaluation kernel (through the C ABI) against the oracle, which is itself
pinned to the reference's encoder and torch inference (tests/test_oracle_nn.py). fp32, tolerance
1e-4 relative (north_star allows 1e-3)."""
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


@pytest.mark.parametrize("tag", ["relu", "clamp"])
def test_eval_matches_reference_golden(orc, eng, tag):
    """Directly against the outputs of the reference's own torch inference (committed golden)."""
    b, d, r = (np.ascontiguousarray(G[k]) for k in ("battles", "durations", "results"))
    eng.upload(b, d, r)
    eng.load_network(G[f"net_{tag}"])
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


@pytest.mark.parametrize("activation", [1, 2])
def test_eval_default_dims_midgame_states(orc, eng, activation):
    """Default network dimensions (198/427 -> 128 -> 59/83, 768 -> 64 -> 64, 32, 64 -> 315), random parameters
    in the reference file format, 8192 mid-game states incl. terminal / fainted-active / forced-switch ones."""
    n = 8192
    b, d, r = midgame_states(orc, n)
    blob = make_net_blob(np.random.default_rng(1234), activation=activation, scale=2.0)
    net = NetOracle(orc, blob)
    ov, olg, ocnt = net.eval_batch(b, d, r)
    eng.upload(b, d, r)
    eng.load_network(blob)
    v, lg, cnt = eng.evaluate()
    compare(v, lg, cnt, ov, olg, ocnt)


def test_eval_rejects_bad_parameter_files(eng):
    import oak_b200
    blob = make_net_blob(np.random.default_rng(0))
    with pytest.raises(oak_b200.OakB200Error):
        eng.load_network(np.concatenate([blob, np.zeros(3, np.uint8)]))
    with pytest.raises(oak_b200.OakB200Error):
        eng.load_network(blob[:1000])
