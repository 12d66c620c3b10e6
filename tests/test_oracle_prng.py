"""PRNG known answers from the reference (common/rng.zig:73-77 PSRNG, :167-179 Gen56)."""
import ctypes

from kat_harness import oracle_lib


def test_psrng_known_answer():
    lib = oracle_lib()
    lib.oracle_psrng_range.restype = ctypes.c_uint32
    seed = ctypes.c_uint64(0x0001000200030004)
    assert lib.oracle_psrng_range(ctypes.byref(seed), 0, 256) == 121
    # chance(128, 256) is false because the draw is 226
    assert lib.oracle_psrng_range(ctypes.byref(seed), 0, 256) == 226


def test_gen56_lcg_known_answers():
    lib = oracle_lib()
    lib.oracle_psrng_next.restype = ctypes.c_uint32
    for start, n, want in [(0x0, 5, 0xC83FB970153A9227), (0x0, 10, 0x67795501267F125A),
                           (0x8000000000000000, 5, 0x483FB970153A9227), (0x8000000000000000, 10, 0xE7795501267F125A)]:
        seed = ctypes.c_uint64(start)
        for _ in range(n):
            lib.oracle_psrng_next(ctypes.byref(seed))
        assert seed.value == want
