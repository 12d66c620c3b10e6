// ORACLE — TEST INFRASTRUCTURE ONLY (see gen1_oracle.hpp).
//
// CPU restatement of Oak's leaf evaluation: battle encoding -> embedding nets -> main MLP.
//   encoder        cpp/include/encode/battle/battle.h:17-551 (sparse writers), key.h:22-30,65-71, policy.h:29-58
//   affine / nets  cpp/include/nn/affine.h:34-103, nn/ffn.h:39-75, nn/battle/main-net.h:36-107
//   network        cpp/include/nn/battle/network.h:52-175 (read_parameters, value_policy_inference,
//                  write_battle_embedding), file header cpp/src/search.cc:127-147
// Evaluation here is cache-free: the reference memoises embeddings (nn/battle/cache.h), whose values
// equal a direct evaluation for the Pokemon cache (keyed by everything that varies) and equal the
// FIRST evaluation for the active cache (its key omits the confusion/disable/attacking/binding
// duration bits, cache.h:186-221); the cache-free value is the function the memo approximates.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

#include "gen1_oracle.hpp"

namespace oracle {

static const int POKEMON_IN = 198, ACTIVE_IN = 229, ACTIVE_POKEMON_IN = 427, POLICY_DIM = 315;

struct Sparse {
  float v[ACTIVE_POKEMON_IN];
  uint16_t i[ACTIVE_POKEMON_IN];
  uint32_t n = 0;
  void put(uint16_t index, float value) {
    v[n] = value;
    i[n] = index;
    n++;
  }
};

// battle.h:17-45 (max_hp_value = compute_stat(Chansey base hp 250, hp) = 703)
static inline void enc_stats(const Stats& s, Sparse& o, uint16_t& off) {
  o.put(off + 0, s.hp / 703.0f);
  o.put(off + 1, s.atk / 999.0f);
  o.put(off + 2, s.def / 999.0f);
  o.put(off + 3, s.spe / 999.0f);
  o.put(off + 4, s.spc / 999.0f);
  off += 5;
}
// battle.h:74-83
static inline void enc_moves(const MoveSlot* ms, Sparse& o, uint16_t& off) {
  for (int k = 0; k < 4; k++)
    if (ms[k].id != MOVE_Struggle && ms[k].id != MOVE_None && ms[k].pp) o.put(off + ms[k].id - 1, 1.0f);
  off += 164;
}
// battle.h:106-123
static inline int status_index(uint8_t status, uint8_t sleeps) {
  if ((status & 7) == 0) return __builtin_ctz(status) - 3;
  if (!(status & 0x80)) return 3 + sleeps;
  return 14 - (status & 7);
}
static inline void enc_status(uint8_t status, uint8_t sleeps, Sparse& o, uint16_t& off) {  // :155-162
  if (status) o.put(off + status_index(status, sleeps), 1.0f);
  off += 14;
}
static inline void enc_types(uint8_t types, Sparse& o, uint16_t& off) {  // :181-195
  o.put(off + (types & 15), 1.0f);
  if ((types >> 4) != (types & 15)) o.put(off + (types >> 4), 1.0f);
  off += 15;
}
static inline void enc_pokemon(const Pokemon& p, uint8_t sleep, Sparse& o, uint16_t& off) {  // :208-214
  enc_stats(p.stats, o, off);
  enc_moves(p.moves, o, off);
  enc_status(p.status, sleep, o, off);
  enc_types(p.types, o, off);
}
static inline void enc_active(const Active& a, uint32_t duration, Sparse& o, uint16_t& off) {  // :480-489
  enc_stats(a.stats, o, off);
  enc_types(a.types, o, off);
  for (int k = 0; k < 6; k++) {  // Boosts :262-283
    const uint8_t* b = BOOSTS[bget(a, k) + 6];
    o.put(off + k, ((float)b[0] / b[1]) * (k < 4 ? 1 / 4.0f : 1 / 3.0f));
  }
  off += 6;
  {  // Volatiles :319-349 (MultiHit and Flinch are not encoded)
    static const int flags[16] = {V_Bide, V_Thrashing, V_Charging, V_Binding, V_Invulnerable, V_Confusion, V_Mist, V_FocusEnergy,
                                  V_Substitute, V_Recharging, V_Rage, V_LeechSeed, V_Toxic, V_LightScreen, V_Reflect, V_Transform};
    float vals[19];
    for (int k = 0; k < 16; k++) vals[k] = vget(a, flags[k]) ? 1.0f : 0.0f;
    vals[16] = vfget(a, VF_state) / 65535.0f;
    vals[17] = vfget(a, VF_substitute) / 177.0f;  // chansey_sub = 706 / 4 + 1
    vals[18] = vfget(a, VF_toxic) / 16.0f;
    for (int k = 0; k < 19; k++)
      if (vals[k] != 0.0f) o.put(off + k, vals[k]);
    off += 19;
  }
  enc_moves(a.moves, o, off);
  {  // Duration :381-409
    const uint32_t c = dget(duration, D_confusion), d = dget(duration, D_disable), at = dget(duration, D_attacking),
                   b = dget(duration, D_binding);
    if (c) o.put(off + c - 1, 1.0f);
    off += 5;
    if (d) o.put(off + d - 1, 1.0f);
    off += 8;
    if (at) o.put(off + at - 1, 1.0f);
    off += 3;
    if (b) o.put(off + b - 1, 1.0f);
    off += 4;
  }
}
static inline void enc_active_pokemon(const Pokemon& p, const Active& a, uint32_t duration, Sparse& o) {  // :544-551
  uint16_t off = 0;
  enc_active(a, duration, o, off);
  enc_pokemon(p, (uint8_t)sleeps_get(duration, 0), o, off);
}

static inline uint8_t pokemon_key(const Pokemon& p, uint8_t sleep) {  // key.h:22-30,65-71
  uint8_t key = 0;
  for (int k = 0; k < 4; k++) key |= (p.moves[k].pp != 0) << k;
  if (p.status) key |= (uint8_t)((status_index(p.status, sleep) + 1) << 4);
  return key;
}
static inline uint16_t policy_index(const Side& side, uint8_t choice) {  // policy.h:29-58
  const int type = choice & 3, data = choice >> 2;
  if (type == 1) {
    // data == 0 (the forced "move 0" of recharge / locked-in turns) indexes moves[-1] in the reference: the two bytes before
    // Pokemon::moves, i.e. the low byte of stats.spc. That is what the compiled reference reads (its torch mirror gathers
    // the same index from the training rows, src/oak/lab.py test-consistency), so it is what we read.
    const Pokemon& st = Engine<false, false, false, true, false>::stored(const_cast<Side&>(side));
    const uint16_t id = data >= 1 ? st.moves[data - 1].id : (uint16_t)(st.stats.spc & 0xFF);
    return id == 0 ? 0 : id - 1;
  }
  if (type == 2) return (uint16_t)(164 + Engine<false, false, false, true, false>::get(const_cast<Side&>(side), data).species - 1);
  return 0;
}

// ---- parameters (affine.h:34-69; order network.h:52-70, main-net.h:36-55; header search.cc:127-147)
struct AffineP {
  uint32_t in = 0, out = 0;
  std::vector<float> bias, w;  // w[out][in] row-major
  const uint8_t* read(const uint8_t* p, const uint8_t* end) {
    if (!p || p + 8 > end) return nullptr;
    memcpy(&in, p, 4);
    memcpy(&out, p + 4, 4);
    p += 8;
    const size_t nb = (size_t)out * 4, nw = (size_t)out * in * 4;
    if (p + nb + nw > end) return nullptr;
    bias.resize(out);
    w.resize((size_t)out * in);
    memcpy(bias.data(), p, nb);
    memcpy(w.data(), p + nb, nw);
    return p + nb + nw;
  }
  void act(float* y, int activation) const {
    for (uint32_t o = 0; o < out; o++) {
      if (activation >= 1) y[o] = y[o] > 0 ? y[o] : 0;
      if (activation == 2) y[o] = y[o] < 1 ? y[o] : 1;
    }
  }
  void dense(const float* x, float* y, int activation) const {  // affine.h:71-84
    for (uint32_t o = 0; o < out; o++) {
      float acc = 0;
      for (uint32_t k = 0; k < in; k++) acc += w[(size_t)o * in + k] * x[k];
      y[o] = acc + bias[o];
    }
    act(y, activation);
  }
  void sparse(const Sparse& s, float* y, int activation) const {  // affine.h:86-102
    for (uint32_t o = 0; o < out; o++) y[o] = bias[o];
    for (uint32_t k = 0; k < s.n; k++)
      for (uint32_t o = 0; o < out; o++) y[o] += w[(size_t)o * in + s.i[k]] * s.v[k];
    act(y, activation);
  }
};

struct BattleNet {
  int activation = 1;  // 1 relu, 2 clamp (header byte 0 + 1)
  AffineP pk0, pk1, ac0, ac1, fc0, fc1, v2, v3, p1a, p1b, p2a, p2b;
  bool load(const uint8_t* data, size_t len) {
    if (len < 8) return false;
    activation = data[0] + 1;
    const uint8_t *p = data + 8, *end = data + len;
    AffineP* all[12] = {&pk0, &pk1, &ac0, &ac1, &fc0, &fc1, &v2, &v3, &p1a, &p1b, &p2a, &p2b};
    for (AffineP* a : all) p = a->read(p, end);
    if (p != end) return false;  // must hit EOF exactly (network.h:61-64)
    return pk0.in == POKEMON_IN && ac0.in == ACTIVE_POKEMON_IN && fc0.in == 2 * ((1 + ac1.out) + 5 * (1 + pk1.out)) &&
           (activation == 1 || activation == 2);
  }

  // network.h:131-175 write_battle_embedding, cache-free
  void embed(const Battle& b, const uint8_t* durations, std::vector<float>& out) const {
    const uint32_t pod = pk1.out, aod = ac1.out, side_dim = (1 + aod) + 5 * (1 + pod);
    out.assign(2 * side_dim, 0.0f);
    uint32_t d[2];
    memcpy(d, durations, 8);
    std::vector<float> h(pk0.out > ac0.out ? pk0.out : ac0.out);
    for (int s = 0; s < 2; s++) {
      const Side& side = b.sides[s];
      float* e = out.data() + s * side_dim;
      const Pokemon& st = side.pokemon[side.order[0] - 1];
      if (st.hp != 0) {
        e[0] = (float)st.hp / st.stats.hp;
        Sparse sp;
        enc_active_pokemon(st, side.active, d[s], sp);
        ac0.sparse(sp, h.data(), activation);
        ac1.dense(h.data(), e + 1, activation);
      }
      for (int slot = 2; slot <= 6; slot++) {
        float* se = e + (1 + aod) + (slot - 2) * (1 + pod);
        const uint8_t id = side.order[slot - 1];
        if (id == 0) continue;
        const Pokemon& pk = side.pokemon[id - 1];
        if (pk.hp == 0) continue;
        se[0] = (float)pk.hp / pk.stats.hp;
        Sparse sp;
        uint16_t off = 0;
        enc_pokemon(pk, (uint8_t)sleeps_get(d[s], slot - 1), sp, off);
        pk0.sparse(sp, h.data(), activation);
        pk1.dense(h.data(), se + 1, activation);
      }
    }
  }

  // network.h:102-123 value_policy_inference + main-net.h:66-107
  float eval(const Battle& b, const uint8_t* durations, int m, int n, const uint8_t* c1, const uint8_t* c2, float* l1, float* l2,
             std::vector<float>* embedding_out = nullptr) const {
    std::vector<float> e;
    embed(b, durations, e);
    if (embedding_out) *embedding_out = e;
    return main(e, b, m, n, c1, c2, l1, l2);
  }
  // main-net.h:66-107 on a given battle embedding
  float main(const std::vector<float>& e, const Battle& b, int m, int n, const uint8_t* c1, const uint8_t* c2, float* l1, float* l2) const {
    std::vector<float> b0(fc0.out), b1(fc1.out), vb(v2.out), pb(p1a.out > p2a.out ? p1a.out : p2a.out);
    fc0.dense(e.data(), b0.data(), activation);
    fc1.dense(b0.data(), b1.data(), activation);
    v2.dense(b1.data(), vb.data(), activation);
    float value;
    v3.dense(vb.data(), &value, 0);
    for (int s = 0; s < 2; s++) {
      const AffineP &a = s == 0 ? p1a : p2a, &o = s == 0 ? p1b : p2b;
      a.dense(b1.data(), pb.data(), activation);
      const int cnt = s == 0 ? m : n;
      const uint8_t* ch = s == 0 ? c1 : c2;
      float* lo = s == 0 ? l1 : l2;
      for (int k = 0; k < cnt; k++) {
        const uint16_t idx = policy_index(b.sides[s], ch[k]);
        float acc = 0;
        for (uint32_t j = 0; j < o.in; j++) acc += o.w[(size_t)idx * o.in + j] * pb[j];
        lo[k] = acc + o.bias[idx];
      }
    }
    return 1.0f / (1.0f + std::exp(-value));
  }

  // ---- "discrete" mode (cpp/src/search.cc:105-123): the float embedding nets feed an int8 main net.
  //   embedding -> u8        static_cast<uint8_t>(127 * f), hp as percent * 127   nn/battle/cache.h:100-108,228-236, network.h:147,165
  //   parameters             int8 = float * 64, int32 bias = float * 64 * 127     quantized/affine.h:74-96
  //   layer                  int32 = bias + sum u8 * int8                          quantized/affine.h:99-186
  //   activation             clamp(x >> 6, 0, 127)                                 quantized/clipped_relu.h:53-98
  //   outputs                value_fc3 / (127 * 64) -> sigmoid; policy rows / (127 * 64)   quantized/main-net.h:104-147
  // The network must carry clamped activations (search.cc:133-141).
  static void quantize_embedding(const std::vector<float>& e, std::vector<uint8_t>& q) {
    q.resize(e.size());
    for (size_t i = 0; i < e.size(); i++) q[i] = (uint8_t)(127 * e[i]);
  }
  static int32_t qbias(float b) { return (int32_t)(b * 64 * 127); }
  static int8_t qweight(float w) { return (int8_t)(w * 64); }
  static void qdense(const AffineP& a, const uint8_t* x, int32_t* y) {
    for (uint32_t o = 0; o < a.out; o++) {
      int32_t acc = qbias(a.bias[o]);
      for (uint32_t k = 0; k < a.in; k++) acc += (int32_t)x[k] * qweight(a.w[(size_t)o * a.in + k]);
      y[o] = acc;
    }
  }
  static void crelu(const int32_t* y, uint8_t* x, uint32_t n) {
    for (uint32_t i = 0; i < n; i++) {
      const int32_t v = y[i] >> 6;
      x[i] = (uint8_t)(v < 0 ? 0 : v > 127 ? 127 : v);
    }
  }
  // integer main net on a given u8 battle embedding; returns value_fc3 / 8128 (before the sigmoid)
  float main_discrete(const uint8_t* q, int m, int n, const uint16_t* idx1, const uint16_t* idx2, float* l1, float* l2) const {
    std::vector<int32_t> y(320);
    std::vector<uint8_t> a0(fc0.out), a1(fc1.out), av(v2.out), ap(p1a.out > p2a.out ? p1a.out : p2a.out);
    qdense(fc0, q, y.data());
    crelu(y.data(), a0.data(), fc0.out);
    qdense(fc1, a0.data(), y.data());
    crelu(y.data(), a1.data(), fc1.out);
    qdense(v2, a1.data(), y.data());
    crelu(y.data(), av.data(), v2.out);
    int32_t v;
    qdense(v3, av.data(), &v);
    const float conversion = 127 * (1 << 6);
    for (int s = 0; s < 2; s++) {
      const AffineP &a = s == 0 ? p1a : p2a, &o = s == 0 ? p1b : p2b;
      qdense(a, a1.data(), y.data());
      crelu(y.data(), ap.data(), a.out);
      const int cnt = s == 0 ? m : n;
      const uint16_t* idx = s == 0 ? idx1 : idx2;
      float* lo = s == 0 ? l1 : l2;
      for (int k = 0; k < cnt; k++) {
        int32_t acc = qbias(o.bias[idx[k]]);
        for (uint32_t j = 0; j < o.in; j++) acc += (int32_t)ap[j] * qweight(o.w[(size_t)idx[k] * o.in + j]);
        lo[k] = acc / conversion;
      }
    }
    return v / conversion;
  }
  float eval_discrete(const Battle& b, const uint8_t* durations, int m, int n, const uint8_t* c1, const uint8_t* c2, float* l1, float* l2,
                      std::vector<uint8_t>* embedding_out = nullptr) const {
    std::vector<float> e;
    embed(b, durations, e);
    std::vector<uint8_t> q;
    quantize_embedding(e, q);
    if (embedding_out) *embedding_out = q;
    uint16_t i1[9], i2[9];
    for (int k = 0; k < m; k++) i1[k] = policy_index(b.sides[0], c1[k]);
    for (int k = 0; k < n; k++) i2[k] = policy_index(b.sides[1], c2[k]);
    const float value = main_discrete(q.data(), m, n, i1, i2, l1, l2);
    return 1.0f / (1.0f + std::exp(-value));
  }
};

// ---- the reference's inference path WITH its embedding memo (nn/battle/cache.h:80-131,186-221; network.h:48-50,131-175):
// per party Pokémon a table of all 16 has-PP x 15 status keys (240 embeddings) filled once per search from the root battle
// (`fill_cache`), and a map from (active block, key of the stored Pokémon) to the active embedding filled on demand. The
// main net is then the only dense work per evaluation. Used as the CPU baseline of the leaf-evaluation bench leg; note the
// reference's active key omits the confusion / disable / attacking / binding duration one-hots, so a memo hit can return the
// embedding of another duration — tests compare it with the cache-free function where the keys do not collide.
struct CachedBattleNet {
  const BattleNet& net;
  std::vector<float> pokemon[2][6];  // [240][pod]
  struct ActiveKey {
    uint8_t bytes[33];
    bool operator<(const ActiveKey& o) const { return memcmp(bytes, o.bytes, 33) < 0; }
  };
  std::map<ActiveKey, std::vector<float>> active[2][6];

  explicit CachedBattleNet(const BattleNet& n) : net(n) {}

  void fill(const Battle& root) {  // BattleCache::fill -> PokemonCache::fill (cache.h:80-124)
    static const uint8_t STATUS[8] = {0, 1u << ST_PSN, 1u << ST_BRN, 1u << ST_FRZ, 1u << ST_PAR, 0x81, 0x82, 0x83};
    const uint32_t pod = net.pk1.out;
    std::vector<float> h(net.pk0.out);
    for (int s = 0; s < 2; s++)
      for (int i = 0; i < 6; i++) {
        active[s][i].clear();
        pokemon[s][i].assign((size_t)240 * pod, 0.0f);
        Pokemon p = root.sides[s].pokemon[i];
        if (p.species == 0) continue;
        for (int m = 0; m < 16; m++) {
          for (int k = 0; k < 4; k++) p.moves[k].pp = (uint8_t)(m & (1 << k));
          auto entry = [&](uint8_t sleep) {
            Sparse sp;
            uint16_t off = 0;
            enc_pokemon(p, sleep, sp, off);
            net.pk0.sparse(sp, h.data(), net.activation);
            net.pk1.dense(h.data(), pokemon[s][i].data() + (size_t)pokemon_key(p, sleep) * pod, net.activation);
          };
          for (uint8_t st : STATUS) {
            p.status = st;
            entry(0);
          }
          p.status = 1;  // Status::Sleep1
          for (uint8_t sleep = 1; sleep <= 7; sleep++) entry(sleep);
        }
      }
  }

  void embed(const Battle& b, const uint8_t* durations, std::vector<float>& out) {  // network.h:131-175
    const uint32_t pod = net.pk1.out, aod = net.ac1.out, side_dim = (1 + aod) + 5 * (1 + pod);
    out.assign(2 * side_dim, 0.0f);
    uint32_t d[2];
    memcpy(d, durations, 8);
    for (int s = 0; s < 2; s++) {
      const Side& side = b.sides[s];
      float* e = out.data() + s * side_dim;
      const int id0 = side.order[0] - 1;
      const Pokemon& st = side.pokemon[id0];
      if (st.hp != 0) {
        e[0] = (float)st.hp / st.stats.hp;
        ActiveKey key;
        memcpy(key.bytes, &side.active, 32);
        key.bytes[32] = pokemon_key(st, (uint8_t)sleeps_get(d[s], 0));
        auto it = active[s][id0].find(key);
        if (it == active[s][id0].end()) {
          Sparse sp;
          enc_active_pokemon(st, side.active, d[s], sp);
          std::vector<float> h(net.ac0.out), emb(aod);
          net.ac0.sparse(sp, h.data(), net.activation);
          net.ac1.dense(h.data(), emb.data(), net.activation);
          it = active[s][id0].emplace(key, std::move(emb)).first;
        }
        memcpy(e + 1, it->second.data(), aod * 4);
      }
      for (int slot = 2; slot <= 6; slot++) {
        float* se = e + (1 + aod) + (slot - 2) * (1 + pod);
        const uint8_t id = side.order[slot - 1];
        if (id == 0) continue;
        const Pokemon& pk = side.pokemon[id - 1];
        if (pk.hp == 0) continue;
        se[0] = (float)pk.hp / pk.stats.hp;
        memcpy(se + 1, pokemon[s][id - 1].data() + (size_t)pokemon_key(pk, (uint8_t)sleeps_get(d[s], slot - 1)) * pod, pod * 4);
      }
    }
  }

  float eval(const Battle& b, const uint8_t* durations, int m, int n, const uint8_t* c1, const uint8_t* c2, float* l1, float* l2) {
    std::vector<float> e;
    embed(b, durations, e);
    return net.main(e, b, m, n, c1, c2, l1, l2);
  }
};

}  // namespace oracle
