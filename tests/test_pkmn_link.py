"""A translation unit compiled against the REFERENCE's own pkmn.h links against liboakb200.so and runs.
tests/pkmn_link/link_tu.c is built by tests/pkmn_link/build.py where /root/reference exists; the .so travels to the GPU box."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

from oak_ref import ROOT, Oracle, pick_word, sample_seed

sys.path.insert(0, os.path.join(ROOT, "tests", "pkmn_link"))
import build as tu_build  # noqa: E402

RECORD = 593


@pytest.fixture(scope="module")
def tu():
    import oak_b200
    oak_b200.load_library()
    so = tu_build.build()
    if so is None:
        pytest.skip("reference pkmn.h not available and no prebuilt unit")
    return ctypes.CDLL(so)


def test_unit_links_and_resolves_every_pkmn_symbol(tu):
    """gcc -Wl,--no-undefined already proved link-time resolution; here every address is non-null and lies in liboakb200.so."""
    tu.tu_symbols.restype = ctypes.c_size_t
    out = (ctypes.c_void_p * 64)()
    n = tu.tu_symbols(out, ctypes.c_size_t(64))
    assert n == 29 and all(out[i] for i in range(n))
    nm = subprocess.run(["nm", "-D", "--defined-only", os.path.join(ROOT, "oak_b200", "liboakb200.so")], capture_output=True, text=True).stdout
    for name in ("PKMN_GEN1_MAX_CHOICES", "PKMN_GEN1_CHOICES_SIZE", "PKMN_GEN1_MAX_LOGS", "PKMN_GEN1_LOGS_SIZE", "PKMN_OPTIONS",
                 "pkmn_error", "pkmn_choice_init", "pkmn_choice_type", "pkmn_choice_data", "pkmn_psrng_init", "pkmn_psrng_next",
                 "pkmn_gen1_battle_options_chance_probability", "pkmn_gen1_battle_options_calc_summaries", "pkmn_rational_reduce"):
        assert f" {name}\n" in nm, name


def test_constants_and_pure_helpers_through_the_reference_header(tu):
    out = (ctypes.c_size_t * 16)()
    tu.tu_constants(out)
    v = list(out)
    assert v[0:6] == [9, 16, 180, 256, 9, 180]
    assert v[6] == 0b1111  # showdown, log, chance, calc
    assert v[7] == 4  # sizeof(pkmn_result_kind): the accessors return full-width enums
    assert v[8:11] == [3, 1, 2] and v[11] == ((5 << 2) | 2) and v[12] == 0b01
    assert v[13:15] == [121, 226]  # common/rng.zig:73-77


@pytest.mark.gpu
def test_c1_trace_through_the_unit(tu):
    """Config C1 (sample battle 0, fixed seed) and two more battles played by C code written against pkmn.h: every update's
    choices, result, battle bytes, durations, actions and protocol-log bytes equal the oracle's."""
    orc = Oracle()
    tu.tu_playout.restype = ctypes.c_size_t
    tu.tu_playout.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint64, ctypes.c_size_t, ctypes.c_void_p,
                              ctypes.c_size_t, ctypes.c_size_t]
    for i in (0, 7, 19):
        b, d, r = orc.sample_battle(i)
        seed = sample_seed(i)
        trace = np.zeros(1200 * RECORD, np.uint8)
        steps = tu.tu_playout(b.ctypes.data, d.ctypes.data, int(r), seed, 1200, trace.ctypes.data, trace.size, 180)
        assert steps > 20
        result = int(r)
        for s in range(steps):
            rec = trace[s * RECORD:(s + 1) * RECORD]
            c1l = orc.choices(b, 0, (result >> 4) & 3)
            c2l = orc.choices(b, 1, (result >> 6) & 3)
            pick = pick_word(seed, s)
            c1, c2 = c1l[pick % len(c1l)], c2l[(pick >> 32) % len(c2l)]
            assert (int(rec[0]), int(rec[1])) == (c1, c2), (i, s)
            result, act, log = orc.update_log(b, c1, c2, d)
            assert int(rec[2]) == result, (i, s)
            assert np.array_equal(rec[5:389], b), (i, s)
            assert np.array_equal(rec[389:397], d) and np.array_equal(rec[397:413], act), (i, s)
            n = int(rec[3]) | int(rec[4]) << 8
            assert n == len(log) and rec[413:413 + n].tobytes() == log, (i, s, list(log), rec[413:413 + n].tolist())
        assert (result & 15) in (1, 2, 3)


@pytest.mark.gpu
def test_log_buffer_too_small_is_pkmn_error(tu):
    orc = Oracle()
    tu.tu_playout.restype = ctypes.c_size_t
    tu.tu_playout.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint8, ctypes.c_uint64, ctypes.c_size_t, ctypes.c_void_p,
                              ctypes.c_size_t, ctypes.c_size_t]
    b, d, r = orc.sample_battle(1)
    trace = np.zeros(4 * RECORD, np.uint8)
    steps = tu.tu_playout(b.ctypes.data, d.ctypes.data, int(r), 5, 4, trace.ctypes.data, trace.size, 3)
    assert steps == 1 and int(trace[2]) == 0x0C
