"""Training rows (EncodedFrames::from_bytes): the oracle restatement against the reference's own row builder — the
committed golden rows (made by tools/gen_training_rows_golden.py with oracle/_ref/ref_pyframes) and, where the reference
module is built, fresh games. CPU only."""
import os
import sys

import numpy as np
import pytest

from oak_ref import ROOT, ROW_FIELDS, Oracle, oracle_training_rows, play_frame_games, split_games


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def _ref_module():
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
    try:
        import ref_pyframes
        return ref_pyframes
    except ImportError:
        return None


def test_rows_match_reference_golden(orc):
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    R = oracle_training_rows(orc, G["stream"], threads=2)
    assert R["k"].shape[0] == 1107
    for name in ROW_FIELDS:
        assert R[name].dtype == G[name].dtype and R[name].shape == G[name].shape, name
        assert np.array_equal(R[name], G[name]), name  # bit-exact, floats included


def test_rows_cover_the_cases_that_matter(orc):
    """The golden games must actually contain fainted slots, status, volatile state, durations and forced moves."""
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    assert (G["hp"] == 0).any() and (G["hp"] == 1).any()
    assert G["pokemon"][..., 169:183].sum() > 0   # status
    assert G["active"][..., 26:42].sum() > 0      # volatile flags
    assert G["active"][..., 209:229].sum() > 0    # durations
    assert (G["choice"] == 1).any()               # the forced "move 0"
    assert (G["choice_indices"] == 315).any() and (G["choice_indices"] >= 164).any()
    assert set(np.unique(G["score"])) <= {0.0, 0.5, 1.0}


def test_rows_match_reference_module_on_fresh_games(orc):
    ref = _ref_module()
    if ref is None:
        pytest.skip("oracle/_ref/ref_pyframes is only built where /root/reference exists")
    stream = play_frame_games(orc, first=123_456, games=24, seed=5)
    R = oracle_training_rows(orc, stream, threads=4)
    at = 0
    for g in split_games(stream):
        n = int.from_bytes(bytes(g[4:6]), "little")
        want = ref.from_bytes(bytes(g), n)
        for name in ROW_FIELDS:
            assert np.array_equal(R[name][at:at + n], np.asarray(want[name])), name
        at += n
    assert at == R["k"].shape[0]


def test_bad_replay_is_reported(orc):
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    stream = G["stream"].copy()
    stream[390] ^= 3  # the first game's stored result no longer matches its replay
    with pytest.raises(AssertionError):
        oracle_training_rows(orc, stream, threads=1)


def test_uncompressed_frames_match_reference_golden(orc):
    """pyoak BattleFrames (py/battle/frames.h): battle / durations / result before every stored update."""
    from oak_ref import oracle_battle_frames
    S = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))["stream"]
    G = np.load(os.path.join(ROOT, "tests", "golden", "battle_frames_golden.npz"))
    R = oracle_battle_frames(orc, S)
    for name in ("battle", "durations", "result"):
        assert R[name].shape == G[name].shape and np.array_equal(R[name], G[name]), name
    assert not G["choices"].any()  # the reference leaves `choices` zero (frames.h:47-49)
    # the legal choices are those of the engine for that frame: counts agree with the stored policy lengths wherever the
    # stored k is the legal count (it always is for games this library records)
    k = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))["k"]
    legal = R["legal_choices"]
    cnt = np.maximum((legal != 0).sum(-1), 1)  # a lone pass (choice 0) counts as one
    assert np.array_equal(cnt, k[..., 0])
    assert (R["durations"] != 0).any() and len(np.unique(R["result"])) > 1
