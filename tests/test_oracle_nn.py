"""Pin the oracle's leaf evaluation (encoder, policy index, pokemon key, parameter file format, forward)
against golden vectors produced by the reference's own code (tools/gen_nn_golden.py: the reference's
header-only C++ encoder compiled in place + its PyTorch mirror src/oak/torch.py)."""
import ctypes
import os

import numpy as np
import pytest

from oak_ref import ROOT, NetOracle, Oracle, make_net_blob

G = np.load(os.path.join(ROOT, "tests", "golden", "nn_golden.npz"))


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def dense(n, v, ix, size):
    out = np.zeros(size, np.float32)
    out[ix[:n]] = v[:n]
    return out


def test_encoder_key_and_policy_index_match_reference_golden(orc):
    lib = orc.lib
    lib.oracle_enc_pokemon_sparse.restype = ctypes.c_uint32
    lib.oracle_enc_active_pokemon_sparse.restype = ctypes.c_uint32
    lib.oracle_pokemon_key.restype = ctypes.c_uint8
    lib.oracle_policy_index.restype = ctypes.c_uint16
    b_all, d_all, r_all = G["battles"], G["durations"], G["results"]
    v = np.zeros(512, np.float32)
    ix = np.zeros(512, np.uint16)
    checked = 0
    for i in range(b_all.shape[0]):
        dd = d_all[i].view(np.uint32)
        for s in range(2):
            side = np.ascontiguousarray(b_all[i, 184 * s:184 * (s + 1)])
            for slot in range(1, 7):
                pid = int(side[176 + slot - 1])
                if pid == 0:
                    continue
                pk = np.ascontiguousarray(side[24 * (pid - 1):24 * pid])
                sleep = (int(dd[s]) >> (3 * (slot - 1))) & 7
                assert lib.oracle_pokemon_key(pk.ctypes.data, ctypes.c_uint8(sleep)) == G["keys"][i, s, slot - 1]
                if int(pk[18]) | int(pk[19]) << 8 == 0:
                    continue
                if slot == 1:
                    n = lib.oracle_enc_active_pokemon_sparse(pk.ctypes.data, np.ascontiguousarray(side[144:176]).ctypes.data,
                                                             ctypes.c_uint32(int(dd[s])), v.ctypes.data, ix.ctypes.data)
                    if i < 16:
                        assert np.array_equal(dense(n, v, ix, 427), G["ref_active_dense"][i, s, 0])
                else:
                    n = lib.oracle_enc_pokemon_sparse(pk.ctypes.data, ctypes.c_uint8(sleep), v.ctypes.data, ix.ctypes.data)
                    if i < 16:
                        assert np.array_equal(dense(n, v, ix, 198), G["ref_pokemon_dense"][i, s, slot - 1])
                assert n == G["enc_n"][i, s, slot - 1]
                checked += 1
            ch = orc.choices(b_all[i], s, (int(r_all[i]) >> (4 + 2 * s)) & 3)
            assert len(ch) == G["counts"][i, s]
            for k, c in enumerate(ch):
                assert lib.oracle_policy_index(side.ctypes.data, ctypes.c_uint8(c)) == G["choice_idx"][i, s, k]
    assert checked > 800


@pytest.mark.parametrize("tag", ["relu", "clamp"])
def test_forward_matches_reference_torch_inference(orc, tag):
    net = NetOracle(orc, G[f"net_{tag}"])
    v, lg, cnt = net.eval_batch(np.ascontiguousarray(G["battles"]), np.ascontiguousarray(G["durations"]),
                                np.ascontiguousarray(G["results"]), threads=2)
    assert np.array_equal(cnt, G["counts"])
    np.testing.assert_allclose(v, G[f"value_{tag}"], rtol=1e-5, atol=1e-6)
    for i in range(v.shape[0]):
        for s in range(2):
            k = cnt[i, s]
            np.testing.assert_allclose(lg[i, s, :k], G[f"logits_{tag}"][i, s, :k], rtol=1e-4, atol=1e-5)
    # and the 2 x side_dim battle embedding fed to the main net
    b, d, r = G["battles"][5], G["durations"][5], int(G["results"][5])
    c1 = orc.choices(b, 0, (r >> 4) & 3)
    c2 = orc.choices(b, 1, (r >> 6) & 3)
    _, _, _, emb = net.eval_one(np.ascontiguousarray(b), np.ascontiguousarray(d), c1, c2)
    np.testing.assert_allclose(emb, G[f"embedding_{tag}"][5], rtol=1e-5, atol=1e-6)


def test_parameter_blob_must_hit_eof_exactly(orc):
    blob = make_net_blob(np.random.default_rng(0), phd=8, pod=4, ahd=8, aod=4, hd=8, vhd=4, pohd=8)
    NetOracle(orc, blob)
    with pytest.raises(ValueError):
        NetOracle(orc, np.concatenate([blob, np.zeros(1, np.uint8)]))  # network.h:61-64
    with pytest.raises(ValueError):
        NetOracle(orc, blob[:-4])


REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_oak.so")


@pytest.mark.skipif(not os.path.exists(REF_SO), reason="oracle/_ref not built (needs /root/reference)")
def test_oak_side_restatements_against_live_reference_code(orc):
    """randomize_hidden_variables (durations.h:25-97) and init_side (init.h:144-154) of the oracle vs the
    reference's own compiled code on many states."""
    from oak_ref import oracle_sample_batch, pick_word, sample_seed, sample_teams
    ref = ctypes.CDLL(REF_SO)
    rng = np.random.default_rng(3)
    n = 400
    b, d, r = oracle_sample_batch(orc, 0, n)
    teams = sample_teams(0, n)
    for i in range(n):
        for s in range(2):
            side = np.zeros(184, np.uint8)
            ref.ref_init_side(np.ascontiguousarray(teams[i, s]).ctypes.data, side.ctypes.data)
            fresh = orc.battle_init(teams[i, 0, :, 0], teams[i, 0, :, 1:].reshape(-1), teams[i, 1, :, 0], teams[i, 1, :, 1:].reshape(-1), 1)
            assert np.array_equal(side, fresh[184 * s:184 * (s + 1)])
        seed = sample_seed(i)
        for step in range(int(rng.integers(5, 60))):
            if int(r[i]) & 15:
                break
            r[i], _, _, _ = orc.step(b[i], d[i], int(r[i]), pick_word(seed, step))
        b[i, 376:384] = rng.integers(0, 256, 8, dtype=np.uint8)
        want = b[i].copy()
        ref.ref_randomize_hidden_variables(want.ctypes.data, d[i].ctypes.data)
        got = b[i].copy()
        orc.randomize_hidden(got, d[i])
        assert np.array_equal(got, want), i


def test_memoised_path_equals_the_cache_free_function():
    """The reference's PokemonCache / ActivePokemonCache path (oracle CachedBattleNet, the bench's CPU baseline) against the
    cache-free forward: benched-Pokémon embeddings come from the 240-entry tables and must agree everywhere; the active memo's
    key omits the duration one-hots (cache.h:186-195), so positions are grouped per battle and the comparison allows the few
    collisions that property causes."""
    from oak_ref import NetOracle, Oracle, make_net_blob, oracle_sample_batch, pick_word, sample_seeds
    orc = Oracle()
    net = NetOracle(orc, make_net_blob(np.random.default_rng(5)))
    n = 96
    b, d, r = oracle_sample_batch(orc, 0, n)
    seeds = sample_seeds(0, n)
    B, D, R, G = [], [], [], []
    for i in range(n):  # 30 positions along each battle's playout: one group (= one fill_cache) per battle
        for s in range(30):
            if (r[i] & 15) != 0:
                break
            B.append(b[i].copy()); D.append(d[i].copy()); R.append(r[i]); G.append(i)
            r[i], _, _, _ = orc.step(b[i], d[i], int(r[i]), pick_word(int(seeds[i]), s))
    B, D, R = np.array(B), np.array(D), np.array(R, np.uint8)
    v0, l0, c0 = net.eval_batch(B, D, R, threads=2)
    v1, l1, c1 = net.eval_batch_cached(B, D, R, np.array(G, np.uint32), threads=2)
    assert np.array_equal(c0, c1)
    same = np.isclose(v0, v1, rtol=1e-5, atol=1e-6) & np.all(np.isclose(l0, l1, rtol=1e-5, atol=1e-6), axis=(1, 2))
    assert same.mean() > 0.97, same.mean()
