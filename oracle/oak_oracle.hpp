// ORACLE — TEST INFRASTRUCTURE ONLY (see gen1_oracle.hpp).
//
// CPU restatement of the Oak-side pieces of the hot path that sit directly on libpkmn:
//   battle construction   cpp/include/libpkmn/pkmn.h:50-57, init.h:35-40,90-154, data/moves.h:1794-1796
//   random-playout loop   cpp/include/search/mcts.h:448-496 (same loop in cpp/src/dur.cc:11-34)
//   hidden-variable draw  cpp/include/search/durations.h:25-97
// plus the synthetic workload of SURVEY.md §8(d) (teams, seeds, counter-based choice stream).
#pragma once

#include "gen1_oracle.hpp"
#include "oak_teams.h"

namespace oracle {

// init.h:35-40  (DV 15, max stat experience => evs/4 = 63)
static inline uint16_t oak_compute_stat(uint8_t base, bool hp, uint8_t level) {
  const uint32_t core = 2 * ((uint32_t)base + 15) + 63;
  const uint32_t factor = hp ? (uint32_t)level + 10 : 5;
  return (uint16_t)(core * level / 100 + factor);
}
// data/moves.h:1794-1796
static inline uint8_t oak_max_pp(uint8_t move) {
  const int v = GEN1_MOVE_PP[move] / 5 * 8;
  return (uint8_t)(v < 61 ? v : 61);
}

// init.h:90-154 + pkmn.h:50-57: level 100, full HP/PP, no status, order[i] = i+1, rng = seed verbatim
static inline void oak_side_init(const uint8_t* species, const uint8_t* moves, Side& side) {
  memset(&side, 0, sizeof(side));
  for (int i = 0; i < 6; i++) {
    Pokemon& pk = side.pokemon[i];
    pk.species = species[i];
    if (pk.species == 0) {
      if (i == 0) side.order[i] = 1;
      continue;
    }
    pk.level = 100;
    const uint8_t* base = &GEN1_SPECIES_STATS[5 * pk.species];
    pk.stats.hp = oak_compute_stat(base[0], true, pk.level);
    pk.stats.atk = oak_compute_stat(base[1], false, pk.level);
    pk.stats.def = oak_compute_stat(base[2], false, pk.level);
    pk.stats.spe = oak_compute_stat(base[3], false, pk.level);
    pk.stats.spc = oak_compute_stat(base[4], false, pk.level);
    for (int m = 0; m < 4; m++) {
      pk.moves[m].id = moves[4 * i + m];
      pk.moves[m].pp = pk.moves[m].id ? oak_max_pp(pk.moves[m].id) : 0;
    }
    pk.hp = pk.stats.hp;
    pk.types = GEN1_SPECIES_TYPES[pk.species];
    if (i == 0 || pk.hp) side.order[i] = (uint8_t)(i + 1);
  }
}

static inline void oak_battle_init(const uint8_t* p1_species, const uint8_t* p1_moves, const uint8_t* p2_species,
                                   const uint8_t* p2_moves, uint64_t seed, Battle& b) {
  memset(&b, 0, sizeof(b));
  oak_side_init(p1_species, p1_moves, b.sides[0]);
  oak_side_init(p2_species, p2_moves, b.sides[1]);
  b.rng = seed;
}

static inline uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static inline uint64_t oak_sample_seed(uint64_t i) { return (i + 1) * 0x9E3779B97F4A7C15ull; }
// choice-stream word for step `step` of a playout seeded with `seed`
static inline uint64_t oak_pick(uint64_t seed, uint64_t step) { return splitmix64(seed ^ (0xD1B54A32D192ED03ull * step)); }

// Synthetic battle i: teams[i & 15] vs teams[(i >> 4) & 15], rng = seed_i, then update(0, 0)
// (cpp/src/generate.cc:238-240); durations start at zero.
static inline void oak_sample_battle(uint64_t i, Battle& b, uint8_t* durations, uint8_t* result) {
  uint8_t sp[2][6], mv[2][24];
  for (int s = 0; s < 2; s++) {
    const unsigned char(*team)[5] = OAK_OU_SAMPLE_TEAMS[s == 0 ? (i & 15) : ((i >> 4) & 15)];
    for (int k = 0; k < 6; k++) {
      sp[s][k] = team[k][0];
      for (int m = 0; m < 4; m++) mv[s][4 * k + m] = team[k][1 + m];
    }
  }
  oak_battle_init(sp[0], mv[0], sp[1], mv[1], oak_sample_seed(i), b);
  Options o{};
  OakEngine e(b, o);
  *result = e.update(0, 0);
  memcpy(durations, o.durations, 8);
}

// One playout step: the body of the loop at mcts.h:467-479 with the pick word supplied by the
// caller: c1 = choices1[pick % m], c2 = choices2[(pick >> 32) % n], options_set(null), update.
static inline uint8_t oak_step(Battle& b, uint8_t* durations, uint8_t result, uint64_t pick, uint8_t* actions_out,
                               uint8_t* c1_out, uint8_t* c2_out) {
  Options o{};
  memcpy(o.durations, durations, 8);
  OakEngine e(b, o);
  uint8_t ch1[9], ch2[9];
  const uint8_t m = e.choices(P1, (result >> 4) & 3, ch1);
  const uint8_t n = e.choices(P2, (result >> 6) & 3, ch2);
  const uint8_t c1 = ch1[pick % m];
  const uint8_t c2 = ch2[(pick >> 32) % n];
  const uint8_t r = e.update(c1, c2);
  memcpy(durations, o.durations, 8);
  if (actions_out) memcpy(actions_out, o.actions, 16);
  if (c1_out) *c1_out = c1;
  if (c2_out) *c2_out = c2;
  return r;
}

// Random playout to a terminal result (mcts.h:448-496); returns the number of updates made.
static inline uint32_t oak_rollout(Battle& b, uint8_t* durations, uint8_t* result, uint64_t seed, uint32_t max_steps) {
  uint32_t step = 0;
  while ((*result & 15) == 0 && step < max_steps) {
    *result = oak_step(b, durations, *result, oak_pick(seed, step), nullptr, nullptr, nullptr);
    step++;
  }
  return step;
}

// durations.h:25-97. Every draw reads the same battle.rng value without advancing it.
static inline void oak_randomize_hidden_variables(Battle& b, const uint8_t* durations) {
  static const uint8_t multi[4][40] = {
      {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 3, 3, 3, 3, 3, 3, 4, 4, 4, 4},
      {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 3, 3, 3, 3, 3, 3, 3, 3},
      {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2},
      {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1}};
  uint32_t d[2];
  memcpy(d, durations, 8);
  const uint64_t rng = b.rng;
  for (int s = 0; s < 2; s++) {
    Side& side = b.sides[s];
    Active& a = side.active;
    const uint32_t confusion = dget(d[s], D_confusion);
    if (confusion) {
      const uint8_t max = (uint8_t)(6 - (confusion + (confusion == 1)));
      vfset(a, VF_confusion, (uint8_t)((rng % max) + 1 + (confusion == 1)));
    }
    const uint32_t disable = dget(d[s], D_disable);
    if (disable) {
      const uint8_t max = (uint8_t)(9 - disable);
      vfset(a, VF_disable_duration, (uint8_t)((rng % max) + 1));
    }
    const uint32_t attacking = dget(d[s], D_attacking);
    if (attacking) {
      if (vget(a, V_Bide) || vget(a, V_Thrashing)) {
        vfset(a, VF_attacks, attacking == 3 ? 1 : 4 - (attacking + (uint32_t)(rng % 2)));
      }
    }
    const uint32_t binding = dget(d[s], D_binding);
    if (binding) vfset(a, VF_attacks, multi[binding - 1][rng % 40]);
    for (int i = 0; i < 6; i++) {
      const uint32_t sleep = sleeps_get(d[s], i);
      if (sleep) {
        Pokemon& pk = side.pokemon[side.order[i] - 1];
        if ((pk.status & 7) != 0 && !(pk.status & 0x80)) {
          const uint8_t max = (uint8_t)(8 - sleep);
          pk.status = (uint8_t)((pk.status & 0xF8) | (uint8_t)((rng % max) + 1));
        }
      }
    }
  }
}

}  // namespace oracle
