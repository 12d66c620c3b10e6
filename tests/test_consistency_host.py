"""The reference's own src/oak/torch.py imports and runs with `oak = oak_b200.pyoak` (dims, tables, OutputBuffer from this
repo), and the torch half of `lab test-consistency` (src/oak/lab.py:21-78) reproduces the committed fixture the GPU half
(tests/test_gpu_consistency.py) is compared with. Needs /root/reference (build container only): skipped elsewhere."""
import os
import sys
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_constant_tables_have_the_reference_shapes():
    import oak_b200.pyoak as oak
    assert (oak.pokemon_in_dim, oak.active_in_dim, oak.policy_out_dim) == (198, 427, 315)
    assert oak.side_out_dim == (1 + oak.active_out_dim) + 5 * (1 + oak.pokemon_out_dim) == 384
    assert (oak.build_policy_hidden_dim, oak.build_value_hidden_dim, oak.build_max_actions) == (128, 128, 339)
    assert len(oak.species_names) == 152 and oak.species_names[25] == "Pikachu" and oak.species_names[151] == "Mew"
    assert len(oak.move_names) == 166 and oak.move_names[0] == "None" and oak.move_names[165] == "Struggle"
    # species_move_list: (species, 0) then (species, move) for each learnable move; the table is its inverse (tensorizer.h:31-99)
    assert len(oak.species_move_list) == 3749 and len(oak.species_move_table) == 152 and len(oak.species_move_table[0]) == 166
    for index, (species, move) in enumerate(oak.species_move_list):
        assert oak.species_move_table[species][move] == index
    assert len(oak.pokemon_dim_labels) == 198 and len(oak.active_dim_labels) == 229 and len(oak.policy_dim_labels) == 316  # pyoak.cc:634 appends ""


def test_seeded_network_is_the_fixtures():
    from gen_consistency_golden import NET_SCALE, NET_SEED
    from oak_ref import make_net_blob
    C = np.load(os.path.join(ROOT, "tests", "golden", "consistency_golden.npz"))
    for tag, act in (("relu", 1), ("clamp", 2)):
        net = make_net_blob(np.random.default_rng(NET_SEED), activation=act, scale=NET_SCALE).tobytes()
        assert zlib.crc32(net) == int(C[f"net_crc_{tag}"][0])


@pytest.mark.skipif(not os.path.exists("/root/reference/src/oak/torch.py"), reason="the reference tree is not on this machine")
def test_reference_torch_module_runs_on_this_oak_module():
    import torch
    import gen_consistency_golden as gen
    from oak_ref import make_net_blob
    saved = {k: sys.modules.get(k) for k in ("oak", "oak.torch")}
    try:
        oak, oak_torch = gen.load_reference_torch()
        assert oak_torch.BattleNetwork.active_in_dim == 427
        build = oak_torch.BuildNetwork()  # sized by len(oak.species_move_list) and the build_* dims
        assert sum(p.numel() for p in build.parameters()) > 3749 * 128
        G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
        C = np.load(os.path.join(ROOT, "tests", "golden", "consistency_golden.npz"))
        net = make_net_blob(np.random.default_rng(gen.NET_SEED), activation=1, scale=gen.NET_SCALE).tobytes()
        lo = 0
        for g, (_, _, n) in enumerate(gen.game_bounds(G["stream"])[:3]):
            v, pl = gen.torch_half(oak, oak_torch, gen.game_rows(G, lo, lo + n), net)
            assert np.allclose(v, C[f"value_relu_{g}"], rtol=0, atol=1e-6) and np.allclose(pl, C[f"policy_logit_relu_{g}"], rtol=0, atol=1e-5)
            assert np.abs(pl).max() > 0.05 and v.std() > 1e-3  # the comparison is not vacuous
            lo += n
        assert torch.get_num_threads() >= 1
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def test_oracle_inference_agrees_with_the_reference_torch_network_on_reference_rows():
    """The CPU half of lab test-consistency on the oracle: oracle/nn_oracle.hpp's inference path on the raw frames the
    reference's Frames::from_bytes produced (battle_frames_golden.npz) against the reference's torch network on the rows its
    EncodedFrames::from_bytes produced, all 1107 rows of 8 games. Pins the encoder, the parameter format, the network and
    the policy index — including the forced 'move 0' choice, where the reference indexes moves[-1] (policy.h:34-36)."""
    from gen_consistency_golden import NET_SCALE, NET_SEED, game_bounds
    from oak_ref import NetOracle, Oracle, make_net_blob
    orc = Oracle()
    G = np.load(os.path.join(ROOT, "tests", "golden", "training_rows_golden.npz"))
    F = np.load(os.path.join(ROOT, "tests", "golden", "battle_frames_golden.npz"))
    C = np.load(os.path.join(ROOT, "tests", "golden", "consistency_golden.npz"))
    forced = 0
    for tag, act in (("relu", 1), ("clamp", 2)):
        net = NetOracle(orc, make_net_blob(np.random.default_rng(NET_SEED), activation=act, scale=NET_SCALE))
        lo = 0
        for g, (_, _, n) in enumerate(game_bounds(G["stream"])):
            b, d = np.ascontiguousarray(F["battle"][lo:lo + n]), np.ascontiguousarray(F["durations"][lo:lo + n])
            r = np.ascontiguousarray(F["result"][lo:lo + n, 0])
            v, lg, cnt = net.eval_batch(b, d, r)
            assert np.array_equal(cnt, G["k"][lo:lo + n, :, 0])
            legal = np.arange(9)[None, None, :] < cnt[:, :, None]
            lg = np.where(legal, lg, 0.0)
            assert np.abs(v - C[f"value_{tag}_{g}"][:, 0]).max() < 2e-6, (tag, g)
            assert np.abs(lg - C[f"policy_logit_{tag}_{g}"]).max() < 2e-5, (tag, g)
            # rows whose only choice is the forced move 0 and whose moves[-1] read is not zero: the case the index fix is for
            ci = G["choice_indices"][lo:lo + n]
            forced += int(((cnt == 1) & (ci[:, :, 0] != 0) & (ci[:, :, 0] < 164)).sum())
            lo += n
    assert forced > 0
