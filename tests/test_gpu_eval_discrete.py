"""Discrete ("int8 main net") leaf evaluation on the GPU, both kernels, against the oracle (whose integer net is pinned
bit-exactly to the reference's quantized layers, tests/test_oracle_nn_discrete.py).

What can and cannot be exact: the main net is integer arithmetic and is reproduced exactly. Its u8 inputs are
floor(127 * f) of the FLOAT embedding nets, and a float sum taken in another order (Eigen in the reference, a serial loop
in the oracle, lane-parallel / tensor-core accumulation here) can land on the other side of an integer boundary for a few
of the 768 inputs. So: an evaluation whose u8 embedding is reproduced must match the oracle bit for bit; the others may
differ by one input step, i.e. by |w8| / 8128 per flipped input."""
import numpy as np
import pytest

from oak_ref import (NetOracle, Oracle, blob_layers, dyadic_clamp_blob, make_net_blob, oracle_eval_discrete_batch, oracle_sample_batch, pick_word,
                     sample_seeds)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup():
    import oak_b200
    orc = Oracle()
    n = 4096
    b, d, r = oracle_sample_batch(orc, 0, n)
    seeds = sample_seeds(0, n)
    for i in range(n):  # mid-game positions of different depths
        for s_ in range(i % 40):
            if (r[i] & 15) == 0:
                r[i], _, _, _ = orc.step(b[i], d[i], int(r[i]), pick_word(int(seeds[i]), s_))
    blob = make_net_blob(np.random.default_rng(99), activation=2, scale=3.0)
    eng = oak_b200.BattleEngine(n)
    yield orc, eng, blob, b, d, r
    eng.close()


@pytest.mark.parametrize("mode", [1, 2])
def test_discrete_eval_matches_oracle(setup, mode):
    orc, eng, blob, b, d, r = setup
    net = NetOracle(orc, blob)
    ov, ol, oc = oracle_eval_discrete_batch(net, b, d, r)
    eng.load_network(blob, discrete=True)
    eng.set_eval_mode(mode)
    eng.upload(b, d, r)
    v, lg, cnt = eng.evaluate()
    assert np.array_equal(cnt, oc)
    mask = np.arange(9)[None, None, :] < oc[:, :, None]
    assert np.all(lg[~mask] == 0)
    same = (np.abs(v - ov) < 1e-6) & np.all(lg == ol, axis=(1, 2))  # value: expf vs std::exp differ in the last place
    # the integer net is exact: most positions are reproduced bit for bit ...
    assert same.mean() > 0.7, same.mean()
    # ... and the rest differ by a few input steps: one step moves an output by at most 127 * 127 / 64 / 8128 per layer hop
    wmax = max(np.abs(np.trunc(w * 64)).max() for w, _ in blob_layers(blob)[4:])
    step = wmax / 8128.0
    assert np.abs(lg - ol)[mask].max() <= 40 * step and np.abs(v - ov).max() <= 0.05
    # every output is a multiple of 1 / 8128 (integer accumulators), logits included
    assert np.allclose(lg[mask] * 8128.0, np.round(lg[mask] * 8128.0), atol=1e-2)


@pytest.mark.parametrize("mode", [1, 2])
def test_discrete_eval_is_bit_exact_when_the_embeddings_are(setup, mode):
    """With embedding nets that sum exactly in any order (dyadic weights, see dyadic_clamp_blob) the u8 inputs of the main
    net are the oracle's, and then EVERY value and logit must equal the oracle's bit for bit."""
    orc, eng, _, b, d, r = setup
    blob = dyadic_clamp_blob(np.random.default_rng(5))
    net = NetOracle(orc, blob)
    ov, ol, oc, emb = oracle_eval_discrete_batch(net, b, d, r, want_embeddings=True)
    assert emb.max() == 127 and (emb > 0).mean() > 0.2 and np.unique(emb).size > 100  # a live network, not all zeros
    eng.load_network(blob, discrete=True)
    eng.set_eval_mode(mode)
    eng.upload(b, d, r)
    v, lg, cnt = eng.evaluate()
    assert np.array_equal(cnt, oc)
    assert np.array_equal(lg, ol)  # integer accumulators / 8128: bit for bit
    assert np.abs(v - ov).max() < 1e-6  # sigmoid of an identical argument: CUDA expf vs glibc exp, last place


def test_discrete_outputs_agree_between_kernels(setup):
    orc, eng, blob, b, d, r = setup
    eng.load_network(blob, discrete=True)
    eng.upload(b, d, r)
    eng.set_eval_mode(1)
    v1, l1, c1 = eng.evaluate()
    eng.set_eval_mode(2)
    v2, l2, c2 = eng.evaluate()
    assert np.array_equal(c1, c2)
    assert ((np.abs(v1 - v2) < 1e-6) & np.all(l1 == l2, axis=(1, 2))).mean() > 0.7


def test_discrete_load_checks(setup):
    from oak_b200.engine import OakB200Error
    orc, eng, blob, b, d, r = setup
    relu = make_net_blob(np.random.default_rng(1), activation=1)
    with pytest.raises(OakB200Error):
        eng.load_network(relu, discrete=True)  # header must say clamp (search.cc:137-141)
    big = make_net_blob(np.random.default_rng(1), activation=2, scale=40.0)
    with pytest.raises(OakB200Error):
        eng.load_network(big, discrete=True)  # weights outside (-2, 2) (quantized/affine.h:88-95)
    eng.load_network(blob)  # the float network still loads afterwards
    eng.set_eval_mode(0)
