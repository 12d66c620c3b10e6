"""oakb200_rows_build_static — the allocation-free variant of the training-row build (no counterpart in the reference)
— against oakb200_rows_build_device through the C ABI: the twelve row arrays and the
raw-frame arrays bit for bit (the tolerance of tests/test_gpu_training_rows.py: exact, floats included), on two calls in a
row into dirty outputs, the second on a different stream of games so that anything stale in the shared static scratch would
show. Collected last on purpose."""
import ctypes
import os

import numpy as np
import pytest

from oak_ref import ROOT

pytestmark = pytest.mark.gpu


def _alloc(torch, n, dev):
    from oak_b200.training_rows import _frame_shapes, _shapes
    shapes = dict(_shapes(torch))
    for k, v in _frame_shapes(torch).items():
        if k != "choices":
            shapes.setdefault(k, v)
    return {name: torch.full((n,) + shape, 7, dtype=dtype, device=dev) for name, (shape, dtype) in shapes.items()}


def _build_both(torch, eng, stream_np, dev):
    from oak_b200.engine import rows_index_host
    off, base, rows = rows_index_host(stream_np)
    games = len(off)
    d_stream = torch.from_numpy(stream_np).to(dev)
    d_off = torch.from_numpy(off.astype(np.uint64).view(np.int64)).to(dev)
    d_base = torch.from_numpy(base.astype(np.uint32).view(np.int32)).to(dev)
    want, got = _alloc(torch, rows, dev), _alloc(torch, rows, dev)
    bad = torch.full((1,), 5, dtype=torch.int32, device=dev)
    eng.rows_build_device(d_stream, stream_np.size, d_off, d_base, games, rows, want)
    assert eng.rows_status() == -1
    launches = eng.launch_count
    eng.rows_build_static(d_stream, stream_np.size, d_off, d_base, games, rows, got, first_bad_out=bad)
    eng.synchronize()
    assert eng.launch_count == launches + 2  # replay + encode: the library's own kernels ran
    assert int(bad.cpu().numpy().view(np.uint32)[0]) == 0xFFFFFFFF
    for name in want:
        assert torch.equal(want[name], got[name]), name
    return rows, (d_stream, d_off, d_base, got)


def test_static_scratch_build_equals_the_allocating_build_twice():
    import torch
    import oak_b200
    dev = torch.device("cuda", 0)
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    stream = np.ascontiguousarray(G["stream"], np.uint8)
    eng = oak_b200.BattleEngine(1, device=0)
    eng.set_stream(torch.cuda.current_stream(dev).cuda_stream)
    # first call: all 8 golden games (1 107 rows); also equal to the rows the reference's own builder produced
    rows, keep = _build_both(torch, eng, stream, dev)
    assert rows == 1107
    from oak_ref import ROW_FIELDS
    for name in ROW_FIELDS:
        a = keep[3][name].cpu().numpy()
        a = a.view(np.uint32) if name == "iterations" else a
        assert np.array_equal(a, G[name]), name
    # second call right after it: other games (the last five, so every row index holds a different state than before)
    from oak_b200.engine import rows_index_host
    off, _, _ = rows_index_host(stream)
    rows2, _ = _build_both(torch, eng, np.ascontiguousarray(stream[int(off[3]):]), dev)
    assert 0 < rows2 < rows
    # a corrupted stored result is reported through first_bad_game_out exactly as oakb200_rows_status reports it
    broken = stream.copy()
    broken[int(off[2]) + 390] ^= 3
    d_stream, d_off, d_base, got = keep
    d_broken = torch.from_numpy(broken).to(dev)
    bad = torch.zeros(1, dtype=torch.int32, device=dev)
    eng.rows_build_static(d_broken, broken.size, d_off, d_base, len(off), rows, got, first_bad_out=bad)
    eng.synchronize()
    assert int(bad.cpu()[0]) == 2
    eng.rows_build_device(d_broken, broken.size, d_off, d_base, len(off), rows, got)
    assert eng.rows_status() == 2
    eng.close()


def test_static_scratch_build_rejects_sizes_beyond_its_maximum():
    import torch
    import oak_b200
    eng = oak_b200.BattleEngine(1, device=0)
    launches = eng.launch_count
    one = torch.zeros(8, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(oak_b200.OakB200Error, match="OAKB200_ROWS_STATIC_MAX"):
        eng.rows_build_static(one, 8, one, one, 1, 16385, {})
    assert eng.launch_count == launches  # nothing was enqueued
    eng.close()
