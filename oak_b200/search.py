"""Batched-leaf tree search over the oakb200 C ABI (oakb200_search_*): host trees and UCB bandits in the library's C++
host code, every battle operation on the GPU. Mirrors MCTS::Search::run -> MCTS::Output (reference
cpp/include/search/mcts.h:68-89,154-246) for many roots at once."""
import ctypes

import numpy as np

from .engine import OakB200Error, _np, _ptr


class TreeSearch:
    def __init__(self, engine, roots, particles=1):
        self.eng, self.lib = engine, engine.lib
        vp, sz, u32 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32
        lib = self.lib
        lib.oakb200_search_create.restype = ctypes.c_int
        lib.oakb200_search_create.argtypes = [vp, sz, u32, ctypes.POINTER(vp)]
        lib.oakb200_search_destroy.restype = None
        lib.oakb200_search_destroy.argtypes = [vp]
        lib.oakb200_search_set_roots.restype = ctypes.c_int
        lib.oakb200_search_set_roots.argtypes = [vp, sz, vp, vp, vp, vp]
        lib.oakb200_search_run.restype = ctypes.c_int
        lib.oakb200_search_run.argtypes = [vp, u32, ctypes.c_float, u32, u32, ctypes.c_int]
        lib.oakb200_search_output.restype = ctypes.c_int
        lib.oakb200_search_output.argtypes = [vp] * 9
        lib.oakb200_search_stats.restype = ctypes.c_int
        lib.oakb200_search_stats.argtypes = [vp, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
        h = vp()
        self._check(lib.oakb200_search_create(engine._h, int(roots), int(particles), ctypes.byref(h)))
        self._h, self.roots_max, self.particles, self.n = h, int(roots), int(particles), 0

    def _check(self, rc):
        if rc != 0:
            raise OakB200Error(f"oakb200 search error {rc}: {self.lib.oakb200_last_error(self.eng._h).decode()}")

    def close(self):
        if getattr(self, "_h", None):
            self.lib.oakb200_search_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_roots(self, battles, durations, results, seeds):
        battles = _np(battles, np.uint8)
        n = battles.shape[0]
        battles, durations = _np(battles, np.uint8, (n, 384)), _np(durations, np.uint8, (n, 8))
        results, seeds = _np(results, np.uint8, (n,)), _np(seeds, np.uint64, (n,))
        self._check(self.lib.oakb200_search_set_roots(self._h, n, _ptr(battles), _ptr(durations), _ptr(results), _ptr(seeds)))
        self.n = n

    BANDITS = {"ucb": 0, "ucb1": 1, "pucb": 2, "exp3": 3, "pexp3": 4}

    def set_bandit(self, spec):
        """The reference's bandit strings (cpp/src/search.cc:236-300): "ucb-<c>", "ucb1-<c>", "pucb-<c>", "exp3-<gamma>[-<alpha>]",
        "pexp3-<gamma>[-<alpha>]". pucb / pexp3 need set_leaf(1) (their priors are the network's policy)."""
        parts = spec.split("-")
        if len(parts) < 2 or parts[0] not in self.BANDITS:
            raise ValueError(f"Could not parse bandit string: {spec}")
        f1 = float(parts[1])
        f2 = float(parts[2]) if len(parts) >= 3 else 0.05
        fn = self.lib.oakb200_search_set_bandit
        fn.restype = ctypes.c_int
        fn.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_float]
        self._check(fn(self._h, self.BANDITS[parts[0]], f1, f2))

    def run(self, iterations, c=1.0, root_rolls=3, other_rolls=1, host_threads=0):
        self._check(self.lib.oakb200_search_run(self._h, int(iterations), float(c), int(root_rolls), int(other_rolls), int(host_threads)))

    def output(self):
        n = self.n
        out = dict(k=np.zeros((n, 2), np.uint8), choices=np.zeros((n, 2, 9), np.uint8), visit_matrix=np.zeros((n, 9, 9), np.uint32),
                   value_matrix=np.zeros((n, 9, 9), np.float64), iterations=np.zeros(n, np.uint64),
                   empirical_value=np.zeros(n, np.float64), empirical=np.zeros((n, 2, 9), np.float64), nodes=np.zeros(n, np.uint64))
        self._check(self.lib.oakb200_search_output(self._h, *[_ptr(out[k]) for k in ("k", "choices", "visit_matrix", "value_matrix",
                                                                                  "iterations", "empirical_value", "empirical", "nodes")]))
        return out

    def timing(self):
        """-> dict(total_ms, host_ms, device_wait_ms, waves) of the last run (oakb200_search_timing)."""
        out = (ctypes.c_double * 4)()
        f = self.lib.oakb200_search_timing
        f.restype = ctypes.c_int
        f.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        self._check(f(self._h, out))
        return dict(total_ms=out[0], host_ms=out[1], device_wait_ms=out[2], waves=int(out[3]))

    def set_leaf(self, mode):
        """0 = MonteCarlo playouts, 1 = the loaded network's value at every new leaf."""
        f = self.lib.oakb200_search_set_leaf
        f.restype = ctypes.c_int
        f.argtypes = [ctypes.c_void_p, ctypes.c_int]
        self._check(f(self._h, int(mode)))

    def nash(self):
        """process_output's Nash half (mcts.h:628-656): -> (nash[n,2,9] f32, nash_value[n] f32)."""
        f = self.lib.oakb200_search_nash
        f.restype = ctypes.c_int
        f.argtypes = [ctypes.c_void_p] * 3
        p, v = np.zeros((self.n, 2, 9), np.float32), np.zeros(self.n, np.float32)
        self._check(f(self._h, _ptr(p), _ptr(v)))
        return p, v

    def stats(self):
        w, lv = ctypes.c_uint64(), ctypes.c_uint64()
        self._check(self.lib.oakb200_search_stats(self._h, ctypes.byref(w), ctypes.byref(lv)))
        return w.value, lv.value


def solve_matrix(row_payoff, discretize_factor=256):
    """pyoak.solve_matrix (reference cpp/src/pyoak.cc:394-425): row player's payoffs in [0, 1] -> (p1, p2, value) of the
    one-sum matrix game after discretising to 1/discretize_factor. Host code of liboakb200 (no GPU needed)."""
    from .engine import load_library
    lib = load_library()
    lib.oakb200_solve_matrix.restype = ctypes.c_int
    lib.oakb200_solve_matrix.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 3
    a = np.asarray(row_payoff, np.float32)
    if a.ndim != 2:
        raise ValueError("Expecting 2d array")
    m, n = a.shape
    disc = np.ascontiguousarray((a * np.float32(discretize_factor)).astype(np.int32))
    p1, p2, v = np.zeros(m, np.float32), np.zeros(n, np.float32), np.zeros(1, np.float32)
    rc = lib.oakb200_solve_matrix(m, n, _ptr(disc), int(discretize_factor), _ptr(p1), _ptr(p2), _ptr(v))
    if rc != 0:
        raise OakB200Error(f"oakb200_solve_matrix: error {rc}")
    return p1, p2, float(v[0])
