"""The playout pick's remainders by a small divisor (oak_b200/csrc/battle_kernels.cuh: rem_small / mod_small) use two constant
tables instead of division sequences. The tables are read from the source and the method is checked against `%` here, so an
edit of either cannot pass unnoticed on a box without a GPU (on the GPU the playout parity tests cover the same code)."""
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _table(name):
    src = open(os.path.join(ROOT, "oak_b200", "csrc", "battle_kernels.cuh")).read()
    body = re.search(r"__constant__ uint32_t %s\[10\] = \{(.*?)\};" % name, src, re.S).group(1)
    return [int(x.strip().rstrip("u"), 0) for x in body.replace("\n", " ").split(",") if x.strip()]


def test_tables_are_the_reciprocals_and_the_residues():
    recip, two32 = _table("SMALL_RECIP"), _table("SMALL_TWO32")
    assert len(recip) == 10 and len(two32) == 10
    assert recip[1] == 0xFFFFFFFF
    for d in range(2, 10):
        assert recip[d] == (1 << 32) // d
    for d in range(1, 10):
        assert two32[d] == (1 << 32) % d


def test_method_equals_the_remainder():
    recip, two32 = np.array(_table("SMALL_RECIP"), np.uint64), np.array(_table("SMALL_TWO32"), np.uint64)
    rng = np.random.default_rng(7)
    x = np.concatenate([np.arange(0, 1 << 16, dtype=np.uint64), (np.uint64(1 << 32) - np.arange(1, 1 << 12, dtype=np.uint64)),
                        rng.integers(0, 1 << 32, 1 << 18, dtype=np.uint64)])
    v = rng.integers(0, np.iinfo(np.uint64).max, 1 << 18, dtype=np.uint64)

    def rem_small(a, d):  # x - umulhi(x, M) * d, then one conditional subtraction
        r = a - ((a * recip[d]) >> np.uint64(32)) * np.uint64(d)
        assert (r < np.uint64(2 * d)).all()
        return np.where(r >= np.uint64(d), r - np.uint64(d), r)

    for d in range(1, 10):
        assert np.array_equal(rem_small(x, d), x % np.uint64(d))
        hi, lo = rem_small(v >> np.uint64(32), d), rem_small(v & np.uint64(0xFFFFFFFF), d)
        assert np.array_equal(rem_small(hi * two32[d] + lo, d), v % np.uint64(d))
