"""The reference's 13 statistical search known answers (cpp/src/search-test.cc:50-109) on the sequential search oracle
(oracle/mcts_oracle.hpp): its only reference-held anchors. GPU counterpart: tests/test_gpu_search.py."""
import ctypes

import numpy as np
import pytest

from oak_ref import Oracle
from test_gpu_search import SEARCH_KATS, kat_position


@pytest.mark.parametrize("case", range(len(SEARCH_KATS)))
def test_search_oracle_known_answers(case):
    orc = Oracle()
    p1, p2, expected, tol = SEARCH_KATS[case]
    b, d, r = kat_position(orc, p1, p2)
    f = orc.lib.oracle_mcts_run
    f.restype = None
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_float, ctypes.c_uint32,
                  ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    iters = 1 << 14
    visit, value = np.zeros(81, np.uint32), np.zeros(81, np.float64)
    f(b.ctypes.data, d.ctypes.data, int(r), 1234 + case, iters, 1.0, 3, 1, visit.ctypes.data, value.ctypes.data, None)
    assert visit.sum() == iters
    assert abs(value.sum() / iters - expected) <= tol + 1e-12, (p1, value.sum() / iters, expected)
