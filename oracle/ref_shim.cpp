// ORACLE — TEST INFRASTRUCTURE ONLY.
// Thin C entry points over the REFERENCE'S OWN header-only C++ (compiled in place from
// /root/reference/cpp/include, never copied): the battle encoder, policy index, pokemon key,
// hidden-variable randomizer and battle construction. Built into oracle/_ref/libref_oak.so by
// oracle/Makefile when /root/reference is present; used to pin the oracle's restatements of those
// pieces and to generate tests/golden fixtures.
#include <cstdint>
#include <cstring>

#include <encode/battle/battle.h>
#include <encode/battle/key.h>
#include <encode/battle/policy.h>
#include <libpkmn/data.h>
#include <libpkmn/init.h>
#include <search/durations.h>

extern "C" {

// Encode::Battle::Pokemon::write dense (battle.h:200-206); out[198] must be zeroed by the caller
void ref_encode_pokemon_dense(const uint8_t* pokemon24, uint8_t sleep, float* out) {
  PKMN::Pokemon p;
  memcpy(&p, pokemon24, 24);
  Encode::Battle::Pokemon::write(p, sleep, out);
}
// Encode::Battle::Active::write dense (battle.h:464-478); out[229]
void ref_encode_active_dense(const uint8_t* active32, uint32_t duration, float* out) {
  PKMN::ActivePokemon a;
  memcpy(&a, active32, 32);
  PKMN::Duration d;
  memcpy(&d, &duration, 4);
  Encode::Battle::Active::write(a, d, out);
}
// sparse ActivePokemon::write (battle.h:544-551): returns n; values/indices sized >= 427
uint32_t ref_encode_active_pokemon_sparse(const uint8_t* pokemon24, const uint8_t* active32, uint32_t duration, float* values,
                                          uint16_t* indices) {
  PKMN::Pokemon p;
  memcpy(&p, pokemon24, 24);
  PKMN::ActivePokemon a;
  memcpy(&a, active32, 32);
  PKMN::Duration d;
  memcpy(&d, &duration, 4);
  float* v = values;
  uint16_t* ix = indices;
  Encode::Battle::ActivePokemon::write(p, a, d, v, ix);
  return (uint32_t)(v - values);
}
// sparse Pokemon::write (battle.h:208-214)
uint32_t ref_encode_pokemon_sparse(const uint8_t* pokemon24, uint8_t sleep, float* values, uint16_t* indices) {
  PKMN::Pokemon p;
  memcpy(&p, pokemon24, 24);
  float* v = values;
  uint16_t* ix = indices;
  Encode::Battle::Pokemon::write(p, sleep, v, ix);
  return (uint32_t)(v - values);
}
uint8_t ref_pokemon_key(const uint8_t* pokemon24, uint8_t sleep) {
  PKMN::Pokemon p;
  memcpy(&p, pokemon24, 24);
  return Encode::Battle::pokemon_key(p, sleep);
}
uint16_t ref_policy_index(const uint8_t* side184, uint8_t choice) {
  PKMN::Side s;
  memcpy(&s, side184, 184);
  return Encode::Battle::Policy::get_index(s, choice);
}
void ref_randomize_hidden_variables(uint8_t* battle384, const uint8_t* durations8) {
  pkmn_gen1_battle b;
  pkmn_gen1_chance_durations d;
  memcpy(&b, battle384, 384);
  memcpy(&d, durations8, 8);
  MCTS::randomize_hidden_variables(b, d);
  memcpy(battle384, &b, 384);
}
// PKMN::Init::init_side (init.h:144-154) for 6 x {species, move x4} at level 100
void ref_init_side(const uint8_t* sets30, uint8_t* side184) {
  struct S {
    PKMN::Data::Species species;
    std::array<PKMN::Data::Move, 4> moves;
  };
  std::array<S, 6> sets{};
  for (int i = 0; i < 6; i++) {
    sets[i].species = static_cast<PKMN::Data::Species>(sets30[5 * i]);
    for (int m = 0; m < 4; m++) sets[i].moves[m] = static_cast<PKMN::Data::Move>(sets30[5 * i + 1 + m]);
  }
  const auto side = PKMN::Init::init_side(sets);
  memcpy(side184, &side, 184);
}
}

// ---- training frames: the reference's own serialiser (train/battle/compressed-frame.h:36-199) ----
#include <train/battle/compressed-frame.h>

extern "C" {
// One game -> bytes. k/choice are [n][2]; probabilities [n][2][9] and values are the doubles a search would output.
uint32_t ref_frames_write_game(const uint8_t* battle384, uint8_t result, uint32_t n_updates, const uint8_t* k, const uint8_t* choice,
                               const uint32_t* iterations, const double* emp_value, const double* nash_value, const double* emp,
                               const double* nash, char* out) {
  using Frames = Train::Battle::CompressedFrames;
  pkmn_gen1_battle b;
  memcpy(&b, battle384, 384);
  Frames frames{b};
  frames.result = result;
  for (uint32_t u = 0; u < n_updates; u++) {
    Frames::Update up{};
    up.iterations = iterations[u];
    up.empirical_value = Train::Battle::compress_probs<double, uint16_t>(emp_value[u]);
    up.nash_value = Train::Battle::compress_probs<double, uint16_t>(nash_value[u]);
    Frames::Update::Side* sides[2] = {&up.p1, &up.p2};
    for (int s = 0; s < 2; s++) {
      sides[s]->k = k[2 * u + s];
      sides[s]->choice = choice[2 * u + s];
      for (int i = 0; i < sides[s]->k; i++) {
        sides[s]->empirical[i] = Train::Battle::compress_probs<double, uint16_t>(emp[(u * 2 + s) * 9 + i]);
        sides[s]->nash[i] = Train::Battle::compress_probs<double, uint16_t>(nash[(u * 2 + s) * 9 + i]);
      }
    }
    frames.updates.push_back(up);
  }
  frames.write(out);
  return frames.n_bytes();
}
// read back through the reference's reader: returns n_updates, fills result and the last update's fields (sanity)
uint32_t ref_frames_read_game(const char* bytes, uint8_t* battle384, uint8_t* result) {
  Train::Battle::CompressedFrames frames{};
  frames.read(bytes);
  memcpy(battle384, frames.battle.bytes, 384);
  *result = frames.result;
  return (uint32_t)frames.updates.size();
}
}
