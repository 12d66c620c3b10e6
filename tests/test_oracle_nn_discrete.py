"""The oracle's int8 main net (nn_oracle.hpp: main_discrete) against the reference's own quantized layers: the committed
golden outputs (tools/gen_nn_discrete_golden.py, made with oracle/_ref) and, where _ref is built, fresh inputs. CPU only,
bit-exact (integer arithmetic, one float division at the end)."""
import os

import numpy as np
import pytest

from oak_ref import ROOT, NetOracle, Oracle, make_net_blob, oracle_main_discrete, ref_library, RefQuantizedMain


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def test_main_discrete_matches_reference_golden(orc):
    G = np.load(os.path.join(ROOT, "tests", "golden", "nn_discrete_golden.npz"))
    net = NetOracle(orc, G["blob"])
    for i in range(G["q"].shape[0]):
        v, l1, l2 = oracle_main_discrete(net, G["q"][i], G["idx1"][i], G["idx2"][i])
        assert v == G["value"][i] and np.array_equal(l1, G["l1"][i]) and np.array_equal(l2, G["l2"][i]), i
    assert np.unique(G["value"]).size > 200  # not a degenerate network


def test_main_discrete_matches_reference_live(orc):
    ref = ref_library()
    if ref is None or not hasattr(ref, "ref_quantized_main"):
        pytest.skip("oracle/_ref is only built where /root/reference exists")
    rng = np.random.default_rng(8)
    for scale in (1.0, 6.0):
        blob = make_net_blob(rng, activation=2, scale=scale)
        net = NetOracle(orc, blob)
        rq = RefQuantizedMain(ref, blob)
        for t in range(300):
            q = rng.integers(0, 128, 768).astype(np.uint8)
            m, n = int(rng.integers(1, 10)), int(rng.integers(1, 10))
            i1, i2 = rng.integers(0, 315, m).astype(np.uint16), rng.integers(0, 315, n).astype(np.uint16)
            a, b = rq(q, i1, i2), oracle_main_discrete(net, q, i1, i2)
            assert a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2]), (scale, t)
