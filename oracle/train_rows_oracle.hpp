// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference's training-row builder: replay one stored game
// from its first battle with the stored choices and, before each update, write the dense rows the learner trains on.
// Pinned against the reference's own header compiled in place (oracle/ref_pyframes.cpp: EncodedFrames::from_bytes) by
// tests/test_oracle_training_rows.py and the committed tests/golden/training_rows_golden.npz.
//   EncodedFrames::write / uncompress_from_bytes   cpp/include/py/battle/encoded-frames.h:53-133
//   Target::write, Update::write_to_tensor         cpp/include/py/battle/target.h:56-70, compressed-frame.h:141-163
//   CompressedFrames::read                         cpp/include/train/battle/compressed-frame.h:201-224
//   dense encoders                                 cpp/include/encode/battle/battle.h:22-29,64-71,146-151,171-178,
//                                                  201-206,251-264,297-326,375-400,465-478
//   Policy::get_index                              cpp/include/encode/battle/policy.h:29-58
//   PKMN::score                                    cpp/include/libpkmn/pkmn.h:174-190
#pragma once

#include <cstdint>
#include <cstring>

#include "gen1_oracle.hpp"

namespace rows_oracle {

using namespace oracle;

static const int POKEMON_DIM = 198, ACTIVE_DIM = 229, POLICY_DIM = 315;

struct Rows {  // row-major arrays of n frames, shapes as pyoak's EncodedBattleFrames (pyoak.cc:655-673)
  float* pokemon;            // [n][2][6][198]
  float* active;             // [n][2][1][229]
  float* hp;                 // [n][2][6][1]
  int64_t* choice_indices;   // [n][2][9]
  uint8_t* k;                // [n][2][1]
  uint8_t* choice;           // [n][2][1]
  uint32_t* iterations;      // [n][1]
  float* empirical_policies; // [n][2][9]
  float* nash_policies;      // [n][2][9]
  float* empirical_value;    // [n][1]
  float* nash_value;         // [n][1]
  float* score;              // [n][1]
};

inline int status_index(uint8_t status, uint8_t sleep) {  // battle.h:106-123
  if ((status & 7) == 0) {
    int bit = 0;
    while (!((status >> bit) & 1)) bit++;
    return bit - 3;  // Psn/Tox 0 (Tox has bit 3 and bit 7), Brn 1, Frz 2, Par 3
  }
  if (!(status & 0x80)) return 3 + sleep;  // natural sleep: turns slept 1..7 -> 4..10
  return 14 - (status & 7);                // Rest: 3,2,1 turns left -> 11,12,13
}

inline void stats_dense(const uint16_t* s, float* t) {  // battle.h:22-29; max hp = Chansey's 703
  t[0] = s[0] / 703.0f;
  for (int i = 1; i < 5; i++) t[i] = s[i] / 999.0f;
}
inline void moves_dense(const uint8_t* slots, float* t) {  // battle.h:64-71 — a later slot with the same id wins
  for (int i = 0; i < 4; i++) {
    const uint8_t id = slots[2 * i], pp = slots[2 * i + 1];
    if (id != 0 && id != 165) t[id - 1] = pp != 0;
  }
}
inline void pokemon_dense(const Pokemon& p, uint8_t sleep, float* t) {  // battle.h:201-206
  uint16_t s[5] = {p.stats.hp, p.stats.atk, p.stats.def, p.stats.spe, p.stats.spc};
  stats_dense(s, t);
  moves_dense(reinterpret_cast<const uint8_t*>(p.moves), t + 5);
  if (p.status) t[169 + status_index(p.status, sleep)] = 1;
  t[183 + (p.types & 15)] = 1;
  t[183 + (p.types >> 4)] = 1;
}
inline void active_dense(const Active& a, uint32_t dur, float* t) {  // battle.h:465-478
  uint16_t s[5] = {a.stats.hp, a.stats.atk, a.stats.def, a.stats.spe, a.stats.spc};
  stats_dense(s, t);
  t[5 + (a.types & 15)] = 1;
  t[5 + (a.types >> 4)] = 1;
  const uint8_t* ab = reinterpret_cast<const uint8_t*>(&a);
  const uint32_t boosts = ab[12] | (ab[13] << 8) | (ab[14] << 16);
  for (int i = 0; i < 6; i++) {  // battle.h:251-264, data.zig boosts table
    int st = (int)((boosts >> (4 * i)) & 15);
    if (st > 7) st -= 16;
    static const int num[13] = {25, 28, 33, 40, 50, 66, 1, 15, 2, 25, 3, 35, 4};
    static const int den[13] = {100, 100, 100, 100, 100, 100, 1, 10, 1, 10, 1, 10, 1};
    t[20 + i] = ((float)num[st + 6] / den[st + 6]) * (i < 4 ? (float)(1 / 4.0) : (float)(1 / 3.0));
  }
  uint64_t v;
  memcpy(&v, ab + 16, 8);
  // battle.h:297-326: Bide Thrashing (MultiHit Flinch skipped) Charging Binding Invulnerable Confusion Mist FocusEnergy
  // Substitute Recharging Rage LeechSeed Toxic LightScreen Reflect Transform
  float* tv = t + 26;
  for (int j = 0; j < 16; j++) tv[j] = (float)((v >> (j < 2 ? j : j + 2)) & 1);
  tv[16] = (float)((v >> 24) & 0xFFFF) / 65535.0f;  // state
  tv[17] = (float)((v >> 40) & 0xFF) / 177.0f;      // substitute hp / (706 / 4 + 1)
  tv[18] = (float)(((v >> 59) & 31) / 16.0);        // toxic counter
  moves_dense(ab + 24, t + 45);
  // battle.h:472-475 clears t[move id] for a disabled move AFTER t has moved past the move block: the store lands at
  // or beyond the Duration block of this row, always on a zero that a later write fills in, so it changes nothing in a
  // freshly cleared buffer (EncodedFrames(sz) clears) — nothing to restate.
  float* td = t + 209;  // battle.h:375-400: confusion 5, disable 8, attacking 3, binding 4
  const uint32_t conf = (dur >> 18) & 7, dis = (dur >> 21) & 15, att = (dur >> 25) & 7, bind = (dur >> 28) & 7;
  if (conf) td[conf - 1] = 1;
  if (dis) td[5 + dis - 1] = 1;
  if (att) td[13 + att - 1] = 1;
  if (bind) td[16 + bind - 1] = 1;
}

inline uint16_t row_policy_index(const Side& side, uint8_t choice) {  // policy.h:29-58
  const int type = choice & 3, data = choice >> 2;
  const Pokemon& stored = side.pokemon[side.order[0] - 1];
  if (type == 1) {
    // data == 0 (the forced "move 0" of recharge / locked moves) indexes moves[-1] in the reference: the slot before
    // Pokemon::moves, i.e. the low byte of stats.spc. That is what the compiled reference reads, so it is what we read.
    const uint8_t id = data >= 1 ? stored.moves[data - 1].id : (uint8_t)(stored.stats.spc & 0xFF);
    return id == 0 ? 0 : (uint16_t)(id - 1);
  }
  if (type == 2) return (uint16_t)(164 + side.pokemon[side.order[data - 1] - 1].species - 1);
  return 0;
}

// one row (EncodedFrames::write, encoded-frames.h:53-110); the four dense arrays must be zero on entry
inline void write_row(const Battle& b, const uint32_t* durations, uint8_t result, uint32_t k1, uint32_t k2, float* pokemon, float* active,
                      float* hp, int64_t* choice_indices) {
  Battle copy = b;
  Options o{};
  OakEngine e(copy, o);
  for (int s = 0; s < 2; s++) {
    const Side& side = b.sides[s];
    const uint32_t dur = durations[s];
    const Pokemon& stored = side.pokemon[side.order[0] - 1];
    if (stored.hp != 0) {
      hp[s * 6] = (float)stored.hp / stored.stats.hp;
      active_dense(side.active, dur, active + s * ACTIVE_DIM);
      pokemon_dense(stored, dur & 7, pokemon + (s * 6) * POKEMON_DIM);
    }
    for (int slot = 2; slot <= 6; slot++) {
      const uint8_t id = side.order[slot - 1];
      if (id == 0) continue;
      const Pokemon& p = side.pokemon[id - 1];
      if (p.hp == 0) continue;
      hp[s * 6 + slot - 1] = (float)p.hp / p.stats.hp;
      pokemon_dense(p, (dur >> (3 * (slot - 1))) & 7, pokemon + (s * 6 + slot - 1) * POKEMON_DIM);
    }
    uint8_t ch[9];
    const uint8_t n = e.choices(s, (result >> (4 + 2 * s)) & 3, ch);
    const uint32_t k = s ? k2 : k1;
    for (int i = 0; i < 9; i++) choice_indices[s * 9 + i] = POLICY_DIM;
    for (uint32_t i = 0; i < k && i < n; i++) choice_indices[s * 9 + i] = row_policy_index(side, ch[i]);
  }
}

// One stored game (compressed-frame.h:201-224 layout) -> rows [first, first + n_updates). Returns n_updates, or -1 if
// the replay does not end on the stored result (encoded-frames.h:132 asserts that).
inline int build_rows(const uint8_t* game, Rows& r, size_t first) {
  uint32_t n_bytes;
  uint16_t n_updates;
  memcpy(&n_bytes, game, 4);
  memcpy(&n_updates, game + 4, 2);
  Battle b;
  memcpy(&b, game + 6, 384);
  const uint8_t stored_result = game[390];
  const float score = (stored_result & 15) == RES_WIN ? 1.0f : (stored_result & 15) == RES_LOSE ? 0.0f : 0.5f;
  Options o{};
  uint8_t result = RESULT_DEFAULT;  // PKMN::result() (pkmn.h:228-233): None, both players to move
  const uint8_t* p = game + 391;
  for (uint32_t u = 0; u < n_updates; u++) {
    const size_t f = first + u;
    const uint32_t m = (p[0] & 15) + 1, n = (p[0] >> 4) + 1;
    const uint8_t c1 = p[1], c2 = p[2];
    memset(r.pokemon + f * 12 * POKEMON_DIM, 0, sizeof(float) * 12 * POKEMON_DIM);
    memset(r.active + f * 2 * ACTIVE_DIM, 0, sizeof(float) * 2 * ACTIVE_DIM);
    memset(r.hp + f * 12, 0, sizeof(float) * 12);
    memset(r.empirical_policies + f * 18, 0, sizeof(float) * 18);
    memset(r.nash_policies + f * 18, 0, sizeof(float) * 18);
    r.k[f * 2] = (uint8_t)m;
    r.k[f * 2 + 1] = (uint8_t)n;
    r.choice[f * 2] = c1;
    r.choice[f * 2 + 1] = c2;
    memcpy(&r.iterations[f], p + 3, 4);
    uint16_t ev, nv;
    memcpy(&ev, p + 7, 2);
    memcpy(&nv, p + 9, 2);
    r.empirical_value[f] = ev / 65535.0f;
    r.nash_value[f] = nv / 65535.0f;
    r.score[f] = score;
    const uint8_t* q = p + 11;
    auto probs = [&](float* dst, uint32_t cnt) {
      for (uint32_t i = 0; i < cnt; i++) {
        uint16_t x;
        memcpy(&x, q + 2 * i, 2);
        dst[i] = x / 65535.0f;
      }
      q += 2 * cnt;
    };
    probs(r.empirical_policies + f * 18, m);
    probs(r.nash_policies + f * 18, m);
    probs(r.empirical_policies + f * 18 + 9, n);
    probs(r.nash_policies + f * 18 + 9, n);
    p = q;
    write_row(b, o.durations, result, m, n, r.pokemon + f * 12 * POKEMON_DIM, r.active + f * 2 * ACTIVE_DIM, r.hp + f * 12,
              r.choice_indices + f * 18);
    memset(o.actions, 0, 16);  // PKMN::update: options_set(null) then update (pkmn.h:134-138)
    memset(o.overrides, 0, 16);
    OakEngine e(b, o);
    result = e.update(c1, c2);
  }
  if ((size_t)(p - game) != n_bytes) return -2;
  return result == stored_result ? (int)n_updates : -1;
}

// One stored game -> the uncompressed frames of pyoak's BattleFrames (Frames::write / uncompress_from_bytes,
// cpp/include/py/battle/frames.h:39-69): the battle, durations and result each update was applied to, and the legal
// choices of both players (which the reference computes at frames.h:47 and leaves unstored). Rows [first, first + n).
inline int build_frames(const uint8_t* game, size_t first, uint8_t* battle, uint8_t* durations, uint8_t* result, uint8_t* legal) {
  uint16_t n_updates;
  memcpy(&n_updates, game + 4, 2);
  Battle b;
  memcpy(&b, game + 6, 384);
  Options o{};
  uint8_t res = RESULT_DEFAULT;
  const uint8_t* p = game + 391;
  for (uint32_t u = 0; u < n_updates; u++) {
    const size_t f = first + u;
    memcpy(battle + f * 384, &b, 384);
    memcpy(durations + f * 8, o.durations, 8);
    result[f] = res;
    Battle copy = b;
    Options oc{};
    OakEngine ec(copy, oc);
    for (int s = 0; s < 2; s++) {
      uint8_t ch[9] = {0};
      const uint8_t k = ec.choices(s, (res >> (4 + 2 * s)) & 3, ch);
      for (int i = 0; i < 9; i++) legal[f * 18 + s * 9 + i] = i < k ? ch[i] : 0;
    }
    const uint32_t m = (p[0] & 15) + 1, n = (p[0] >> 4) + 1;
    memset(o.actions, 0, 16);
    memset(o.overrides, 0, 16);
    OakEngine e(b, o);
    res = e.update(p[1], p[2]);
    p += 11 + 4 * (m + n);
  }
  return res == game[390] ? (int)n_updates : -1;
}

}  // namespace rows_oracle
