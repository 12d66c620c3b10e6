"""The C-ABI library loads anywhere and exports every symbol include/*.h declares; without a CUDA device it fails
loudly instead of computing anything (no CPU path behind the product API)."""
import ctypes
import os
import re

import pytest

import oak_b200
from oak_ref import ROOT


def declared_symbols():
    names = []
    for h in ("oakb200.h", "oakb200_pkmn.h"):
        text = open(os.path.join(ROOT, "include", h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"\b((?:oakb200|pkmn)_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_every_declared_symbol_is_exported():
    lib = oak_b200.load_library()
    names = declared_symbols()
    assert len(names) >= 40 and "oakb200_eval_device" in names and "pkmn_gen1_battle_update" in names
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # data symbols of pkmn.h (engine/src/lib/c.zig:4-14): Oak sizes its choice arrays with PKMN_GEN1_MAX_CHOICES
    for name, value in (("PKMN_GEN1_MAX_CHOICES", 9), ("PKMN_GEN1_CHOICES_SIZE", 16), ("PKMN_GEN1_MAX_LOGS", 180),
                        ("PKMN_GEN1_LOGS_SIZE", 256), ("PKMN_MAX_CHOICES", 9), ("PKMN_CHOICES_SIZE", 16), ("PKMN_MAX_LOGS", 180),
                        ("PKMN_LOGS_SIZE", 256)):
        assert ctypes.c_size_t.in_dll(lib, name).value == value, name
    assert bytes((ctypes.c_uint8 * 4).in_dll(lib, "PKMN_OPTIONS")) == b"\x01\x01\x01\x01"


def test_pure_helpers_need_no_device():
    lib = oak_b200.load_library()
    lib.pkmn_result_type.restype = lib.pkmn_result_p1.restype = lib.pkmn_result_p2.restype = ctypes.c_int  # enums, pkmn.h:108-112
    assert (lib.pkmn_result_type(0x51), lib.pkmn_result_p1(0x51), lib.pkmn_result_p2(0x91)) == (1, 1, 2)
    opts = (ctypes.c_uint8 * 128)()
    lib.pkmn_gen1_battle_options_chance_durations.restype = ctypes.c_void_p
    lib.pkmn_gen1_battle_options_chance_actions.restype = ctypes.c_void_p
    d = lib.pkmn_gen1_battle_options_chance_durations(opts)
    a = lib.pkmn_gen1_battle_options_chance_actions(opts)
    base = ctypes.addressof(opts)
    assert base <= a < base + 112 and base <= d < base + 120 and abs(a - d) >= 8


def test_no_device_means_an_error_not_a_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(oak_b200.OakB200Error):
        oak_b200.BattleEngine(16)
    lib = oak_b200.load_library()
    lib.pkmn_gen1_battle_update.restype = ctypes.c_uint8
    battle = (ctypes.c_uint8 * 384)()
    assert lib.pkmn_gen1_battle_update(battle, 0, 0, None) == 4  # PKMN_RESULT_ERROR
    out = (ctypes.c_uint8 * 9)()
    lib.pkmn_gen1_battle_choices.restype = ctypes.c_uint8
    assert lib.pkmn_gen1_battle_choices(battle, 0, 1, out, ctypes.c_size_t(9)) == 0


def test_rows_index_host_needs_no_gpu():
    """oakb200_rows_index_host is plain host code: header walk of a .battle.data stream, and its error paths."""
    import numpy as np
    from oak_ref import ROOT as R
    from oak_b200.engine import OakB200Error, rows_index_host
    G = np.load(os.path.join(R, "tests", "golden", "training_rows_golden.npz"))
    off, base, rows = rows_index_host(G["stream"])
    assert rows == 1107 and len(off) == 8 and off[0] == 0 and base[0] == 0
    assert int(off[1]) == int.from_bytes(G["stream"][0:4].tobytes(), "little")
    for bad in (G["stream"][:-1], G["stream"][:200], np.concatenate([G["stream"], np.zeros(3, np.uint8)])):
        try:
            rows_index_host(bad)
        except OakB200Error:
            continue
        raise AssertionError("a stream that does not split into whole games must be rejected")
    assert rows_index_host(np.zeros(0, np.uint8))[2] == 0
