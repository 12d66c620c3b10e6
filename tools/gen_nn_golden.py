#!/usr/bin/env python3
"""Golden vectors for the leaf-evaluation half of the hot path, produced by THE REFERENCE'S OWN CODE.

  * encoder / policy index / pokemon key / hidden-variable draw / side construction:
    the reference's header-only C++ compiled in place into oracle/_ref/libref_oak.so (oracle/ref_shim.cpp)
  * network forward + parameter file format: the reference's PyTorch mirror /root/reference/src/oak/torch.py
    (BattleNetwork.write_parameters / inference), imported under a stub `oak` module carrying only the
    eleven integer attributes of cpp/src/pyoak.cc:586-596.

Writes tests/golden/nn_golden.npz (+ the generating states), committed. Run in the build container only.
"""
import ctypes
import importlib.util
import io
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oak_ref import Oracle, pick_word, sample_seed  # noqa: E402

ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_oak.so"))
ref.ref_encode_pokemon_sparse.restype = ctypes.c_uint32
ref.ref_encode_active_pokemon_sparse.restype = ctypes.c_uint32
ref.ref_pokemon_key.restype = ctypes.c_uint8
ref.ref_policy_index.restype = ctypes.c_uint16


def P(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def ref_pokemon(pk, sleep):
    v = np.zeros(512, np.float32)
    ix = np.zeros(512, np.uint16)
    n = ref.ref_encode_pokemon_sparse(P(pk), ctypes.c_uint8(sleep), P(v), P(ix))
    return v[:n].copy(), ix[:n].copy()


def ref_active_pokemon(pk, act, dur):
    v = np.zeros(512, np.float32)
    ix = np.zeros(512, np.uint16)
    n = ref.ref_encode_active_pokemon_sparse(P(pk), P(act), ctypes.c_uint32(dur), P(v), P(ix))
    return v[:n].copy(), ix[:n].copy()


def load_reference_torch():
    oak = types.ModuleType("oak")
    for k, v in dict(pokemon_in_dim=198, active_in_dim=427, pokemon_hidden_dim=128, pokemon_out_dim=59, active_hidden_dim=128,
                     active_out_dim=83, side_out_dim=384, hidden_dim=64, value_hidden_dim=32, policy_hidden_dim=64,
                     policy_out_dim=315, build_policy_hidden_dim=128, build_value_hidden_dim=128).items():
        setattr(oak, k, v)
    for name in ("EncodedBattleFrames", "OutputBuffer", "BuildTrajectories", "BattleFrames"):
        setattr(oak, name, object)
    sys.modules["oak"] = oak
    spec = importlib.util.spec_from_file_location("oak_reference_torch", "/root/reference/src/oak/torch.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def collect_states(orc, count):
    """Mid-playout states with live statuses / volatiles / durations."""
    rng = np.random.default_rng(11)
    out = []
    i = 0
    while len(out) < count:
        b, d, r = orc.sample_battle(i)
        seed = sample_seed(i)
        stop = int(rng.integers(0, 90))
        step = 0
        while (r & 15) == 0 and step < stop:
            r, _, _, _ = orc.step(b, d, r, pick_word(seed, step))
            step += 1
        if (r & 15) == 0:
            out.append((b.copy(), d.copy(), r))
        i += 1
    return out


def main():
    orc = Oracle()
    rt = load_reference_torch()
    states = collect_states(orc, 96)
    M = len(states)

    nets = {}
    torch.manual_seed(1234)
    # small dims keep the committed fixture small; the format and code path are dimension-agnostic
    for tag, act in (("relu", rt.Activation.relu), ("clamp", rt.Activation.clamp)):
        net = rt.BattleNetwork(phd=24, ahd=40, pod=11, aod=13, hd=32, vhd=16, pohd=24, activation=act)
        with torch.no_grad():
            for p_ in net.parameters():
                p_.mul_(3.0)  # spread activations so clamp/relu both bite
        buf = io.BytesIO()
        net.write_parameters(buf)
        nets[tag] = (net, np.frombuffer(buf.getvalue(), np.uint8).copy())

    battles = np.stack([s[0] for s in states])
    durations = np.stack([s[1] for s in states])
    results = np.array([s[2] for s in states], np.uint8)

    pokemon = np.zeros((M, 2, 6, 198), np.float32)
    active = np.zeros((M, 2, 1, 427), np.float32)   # Active(229) ‖ Pokemon(198), as ActivePokemon::write orders it
    hp = np.zeros((M, 2, 6, 1), np.float32)
    choice_idx = np.full((M, 2, 9), 315, np.int64)
    counts = np.zeros((M, 2), np.uint8)
    keys = np.zeros((M, 2, 6), np.uint8)
    enc_n = np.zeros((M, 2, 6), np.int32)
    for i, (b, d, r) in enumerate(states):
        dd = d.view(np.uint32)
        for s in range(2):
            side = b[184 * s:184 * (s + 1)]
            order = side[176:182]
            dur = int(dd[s])
            for slot in range(1, 7):
                pid = int(order[slot - 1])
                if pid == 0:
                    continue
                pk = np.ascontiguousarray(side[24 * (pid - 1):24 * pid])
                cur, mx = int(pk[18]) | int(pk[19]) << 8, int(pk[0]) | int(pk[1]) << 8
                sleep = (dur >> (3 * (slot - 1))) & 7
                keys[i, s, slot - 1] = ref.ref_pokemon_key(P(pk), ctypes.c_uint8(sleep))
                if cur == 0:
                    continue
                hp[i, s, slot - 1, 0] = np.float32(cur) / np.float32(mx)
                if slot == 1:
                    v, ix = ref_active_pokemon(pk, np.ascontiguousarray(side[144:176]), dur)
                    active[i, s, 0, ix] = v
                else:
                    v, ix = ref_pokemon(pk, sleep)
                    pokemon[i, s, slot - 1, ix] = v
                enc_n[i, s, slot - 1] = len(v)
            req = (r >> (4 + 2 * s)) & 3
            ch = orc.choices(b, s, req)
            counts[i, s] = len(ch)
            for k, c in enumerate(ch):
                choice_idx[i, s, k] = ref.ref_policy_index(P(np.ascontiguousarray(side)), ctypes.c_uint8(c))

    out = dict(battles=battles, durations=durations, results=results, counts=counts, choice_idx=choice_idx, keys=keys, enc_n=enc_n,
               ref_pokemon_dense=pokemon[:16], ref_active_dense=active[:16], ref_hp=hp)
    # inputs in the reference's EncodedBattleFrames layout (py/battle/encoded-frames.h:29-37,52-107):
    # active[.,.,0,:229] = Active encoding, pokemon[.,.,0,:] = the active's stored Pokemon encoding
    in_active = np.ascontiguousarray(active[:, :, :, :229])
    in_pokemon = pokemon.copy()
    in_pokemon[:, :, 0, :] = active[:, :, 0, 229:]
    for tag, (net, raw) in nets.items():
        inp = types.SimpleNamespace(size=M, pokemon=torch.from_numpy(in_pokemon), active=torch.from_numpy(in_active),
                                    hp=torch.from_numpy(hp), choice_indices=torch.from_numpy(choice_idx))
        sd = net.side_out_dim
        outp = types.SimpleNamespace(size=M, pokemon=torch.zeros(M, 2, 5, net.pokemon_out_dim),
                                     active_pokemon=torch.zeros(M, 2, 1, net.active_out_dim), sides=torch.zeros(M, 2, 1, sd),
                                     value=torch.zeros(M, 1), logit=torch.zeros(M, 2, 316), policy_logit=torch.zeros(M, 2, 9))
        with torch.no_grad():
            net.inference(inp, outp)  # the reference's own BattleNetwork.inference (src/oak/torch.py:399-437)
        out[f"net_{tag}"] = raw
        out[f"value_{tag}"] = outp.value.numpy().reshape(M)
        out[f"logits_{tag}"] = outp.policy_logit.numpy()
        out[f"embedding_{tag}"] = outp.sides.numpy().reshape(M, -1)
    path = os.path.join(ROOT, "tests", "golden", "nn_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", M, "states")


if __name__ == "__main__":
    main()
