#!/usr/bin/env python3
"""The torch half of the reference's `lab test-consistency` (src/oak/lab.py:21-78), run HERE with the reference's own
src/oak/torch.py imported under `oak = oak_b200.pyoak` (no stub module: the dims, tables and OutputBuffer are this repo's),
on the reference-generated rows of tests/golden/training_rows_golden.npz and a default-dimension network in the
reference's parameter file format drawn from a fixed numpy seed (tests/oak_ref.make_net_blob; the test redraws it and
checks its CRC against the fixture, so the 800 KB of weights are not committed).

Writes tests/golden/consistency_golden.npz = per game: value[n,1], policy_logit[n,2,9] (-inf -> 0 as lab.py:52 does).
The GPU test (tests/test_gpu_consistency.py) runs the other half — oak.cpp_inference on oak.BattleFrames.from_bytes of
the same games — and asserts lab.py's comparison; /root/reference is not on the GPU box, hence the fixture.
Run in the build container only.
"""
import importlib.util
import io
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
NET_SEED, NET_SCALE = 20261105, 2.0
REFERENCE_TORCH = "/root/reference/src/oak/torch.py"


def load_reference_torch():
    """-> (oak, oak.torch): the reference's torch.py executed with this repo's module standing in for `oak`."""
    import oak_b200.pyoak as oak
    sys.modules["oak"] = oak
    spec = importlib.util.spec_from_file_location("oak.torch", REFERENCE_TORCH)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["oak.torch"] = mod
    spec.loader.exec_module(mod)
    return oak, mod


def torch_half(oak, oak_torch, rows, net_bytes, discrete=False):
    """lab.py:36-52 for one game's rows (a namespace of numpy arrays named as oak.EncodedBattleFrames)."""
    network = oak_torch.BattleNetwork(activation=oak_torch.Activation.clamp if discrete else oak_torch.Activation.relu)
    network.read_parameters(io.BytesIO(net_bytes))
    frames = oak_torch.EncodedBattleFrames(rows)
    out = oak_torch.OutputBuffer(oak.OutputBuffer(rows.size))
    with torch.no_grad():
        network.inference(frames, out)
    out.policy_logit[torch.isneginf(out.policy_logit)] = 0.0
    return out.value.numpy().copy(), out.policy_logit.numpy().copy()


def game_rows(G, lo, hi):
    r = types.SimpleNamespace(size=hi - lo)
    for name in ("k", "empirical_policies", "nash_policies", "empirical_value", "nash_value", "score", "hp", "pokemon", "active",
                 "choice_indices"):
        setattr(r, name, np.ascontiguousarray(G[name][lo:hi]))
    return r


def game_bounds(stream):
    """[(byte offset, bytes, frames)] of the stored games (u32 bytes | u16 frames header, cpp/include/train/battle/frame.h)"""
    out, o = [], 0
    while o < len(stream):
        nb = int.from_bytes(stream[o:o + 4].tobytes(), "little")
        out.append((o, nb, int.from_bytes(stream[o + 4:o + 6].tobytes(), "little")))
        o += nb
    return out


def main():
    oak, oak_torch = load_reference_torch()
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    import zlib
    from oak_ref import make_net_blob
    save, lo = {}, 0
    for tag in ("relu", "clamp"):
        lo = 0
        net = make_net_blob(np.random.default_rng(NET_SEED), activation=1 if tag == "relu" else 2, scale=NET_SCALE).tobytes()
        save[f"net_crc_{tag}"] = np.array([zlib.crc32(net)], np.uint32)
        for g, (_, _, n) in enumerate(game_bounds(G["stream"])):
            v, pl = torch_half(oak, oak_torch, game_rows(G, lo, lo + n), net, discrete=tag == "clamp")
            save[f"value_{tag}_{g}"], save[f"policy_logit_{tag}_{g}"] = v, pl
            lo += n
    assert lo == G["k"].shape[0]
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "consistency_golden.npz"), **save)
    print("games", len(game_bounds(G["stream"])), "rows", lo)


if __name__ == "__main__":
    main()
