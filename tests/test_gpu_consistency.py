"""The reference's `lab test-consistency` (src/oak/lab.py:21-78) with this library as `oak`: for every game of a data file,
oak.cpp_inference(oak.BattleFrames.from_bytes(buffer, n_frames), network, discrete, "0") against the reference's torch
BattleNetwork.inference on oak.EncodedBattleFrames.from_bytes of the same game. The torch half ran in the build container
on the reference's own src/oak/torch.py (tools/gen_consistency_golden.py -> tests/golden/consistency_golden.npz; checked
again there by tests/test_consistency_host.py); this is the GPU half and the assertion. lab.py's --eps is the caller's:
here 2e-4 on values in (0,1) and on logits of magnitude ~1 (fp32 on both sides, different summation orders)."""
import os
import sys

import numpy as np
import pytest

from oak_ref import ROOT, make_net_blob

pytestmark = pytest.mark.gpu
EPS = 2e-4


@pytest.mark.parametrize("tag", ["relu", "clamp"])
def test_lab_test_consistency(tag, tmp_path):
    import oak_b200.pyoak as oak
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from gen_consistency_golden import NET_SCALE, NET_SEED
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    C = np.load(os.path.join(ROOT, "tests", "golden", "consistency_golden.npz"))
    data_path, net_path = str(tmp_path / "games.battle.data"), str(tmp_path / "net.battle.net")
    G["stream"].tofile(data_path)
    make_net_blob(np.random.default_rng(NET_SEED), activation=1 if tag == "relu" else 2, scale=NET_SCALE).tofile(net_path)

    buffer_list = oak.read_battle_data(data_path)
    assert len(buffer_list) == 8
    worst_v = worst_l = 0.0
    for g, (buffer, n_frames) in enumerate(buffer_list):
        frames = oak.BattleFrames.from_bytes(buffer, n_frames)
        o2 = oak.cpp_inference(frames, net_path, False, "0")
        frames.close()
        value_diff = np.abs(C[f"value_{tag}_{g}"] - o2.value)
        logit_diff = np.abs(C[f"policy_logit_{tag}_{g}"] - o2.policy_logit)
        assert o2.value.shape == (n_frames, 1) and o2.policy_logit.shape == (n_frames, 2, 9)
        assert (value_diff < EPS).all() and (logit_diff < EPS).all(), (g, value_diff.max(), logit_diff.max())
        worst_v, worst_l = max(worst_v, float(value_diff.max())), max(worst_l, float(logit_diff.max()))
    print(f"max value diff {worst_v:.2e}, max logit diff {worst_l:.2e} over {len(buffer_list)} games")


def test_encoded_frames_host_mirror_feeds_numpy_wrappers():
    """EncodedBattleFrames.host(): the reference's field names as numpy arrays (what src/oak/torch.py:16-28 wraps)."""
    import oak_b200.pyoak as oak
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    f = oak.EncodedBattleFrames.from_bytes(G["stream"].tobytes())
    h = f.host()
    for name in ("k", "empirical_policies", "nash_policies", "empirical_value", "nash_value", "score", "hp", "pokemon", "active",
                 "choice_indices", "iterations"):
        a = getattr(h, name)
        assert isinstance(a, np.ndarray) and a.dtype == G[name].dtype and np.array_equal(a, G[name]), name
    assert h.size == f.size == 1107
    f.close()
