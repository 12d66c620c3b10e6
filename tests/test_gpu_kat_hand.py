"""The reference's "RNG overrides" known answer (gen1/test.zig:10131-10327; tests/kat_hand.py) through the GPU step kernels:
every roll answered by the 16-byte calc overrides of oakb200_step / oakb200_step_log (R13: all Action fields, not only the
damage bytes Oak sets), expected log bytes and HP as the reference wrote them."""
import numpy as np
import pytest

import kat_hand as K
import oak_b200

pytestmark = pytest.mark.gpu


def test_rng_overrides_full_on_the_gpu():
    n = 33  # the same battle in every lane of a warp and one more
    eng = oak_b200.BattleEngine(n, device=0)

    def update_log(b, c1, c2, d, ov):
        eng.upload(np.tile(b, (n, 1)), np.tile(d, (n, 1)), None)
        res, act, logs, lens = eng.update_log(np.full(n, c1, np.uint8), np.full(n, c2, np.uint8),
                                              None if ov is None else np.tile(ov, (n, 1)))
        gb, gd, _ = eng.download()
        assert np.all(gb == gb[0]) and np.all(res == res[0]) and np.all(lens == lens[0]) and np.all(logs == logs[0])
        b[:] = gb[0]
        d[:] = gd[0]
        return int(res[0]), act[0], logs[0, :lens[0]].tobytes()

    hp = K.run_full(update_log)
    assert hp == (376, 547)
    # the plain (non-LOG) kernel with the same overrides reaches the same battle
    b = K.full_start()
    d = np.zeros(8, np.uint8)
    eng.upload(np.tile(b, (n, 1)), None, None)
    eng.update(np.zeros(n, np.uint8), np.zeros(n, np.uint8))
    for c1, c2, ov, _ in K.full_script():
        eng.update(np.full(n, c1, np.uint8), np.full(n, c2, np.uint8), overrides=np.tile(ov, (n, 1)))
    gb, _, _ = eng.download()
    assert (int(gb[5, 18]) | int(gb[5, 19]) << 8, int(gb[5, 202]) | int(gb[5, 203]) << 8) == hp
    eng.close()
