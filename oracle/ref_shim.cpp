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

// ---- "discrete" main net: the reference's own int8 layers (nn/battle/quantized/{affine,clipped_relu}.h, AVX2) ----
// quantized/main-net.h itself includes nn/affine.h (Eigen, an empty submodule here), so its propagate() glue — the fixed
// order of layer calls, main-net.h:104-147 — is restated below around the reference's layer classes; parameters go in
// through each layer's own try_copy_parameters (the float -> int8 / int32 conversion under test).
#include <stdexcept>
#include <string>

#include <nn/battle/quantized/affine.h>
#include <nn/battle/quantized/clipped_relu.h>

namespace {
struct FakeVec {
  const float* p;
  const float* data() const { return p; }
};
struct FakeAffine {  // the members AffineTransform::try_copy_parameters reads from NN::Affine
  uint32_t in_dim, out_dim;
  FakeVec weights, biases;
};
}  // namespace

extern "C" {
// Default shape 768 -> 64 -> 64, value 32, policy 64 -> 320 (315 rows zero-padded, main-net.h:56-63).
// w[i] / b[i] for i = fc0, fc1, value_fc2, value_fc3, p1_fc2, p1_fc3 (320 x 64), p2_fc2, p2_fc3 (320 x 64), row-major [out][in].
// Returns value_fc3 / (127 * 64) (before the sigmoid); p1[m], p2[n] = the chosen policy rows / (127 * 64). -1e9 on a
// parameter outside (-2, 2).
float ref_quantized_main(const float* const* w, const float* const* b, const uint8_t* input768, int m, int n, const uint16_t* idx1,
                         const uint16_t* idx2, float* p1, float* p2) {
  using namespace NN::Battle::Quantized;
  static AffineTransform<768, 64> fc0;
  static ClippedReLU<64> ac0;
  static AffineTransform<64, 64> fc1;
  static ClippedReLU<64> ac1;
  static AffineTransform<64, 32> v2;
  static ClippedReLU<32> vac2;
  static AffineTransform<32, 1> v3;
  static AffineTransform<64, 64> p1a, p2a;
  static ClippedReLU<64> p1ac, p2ac;
  static AffineTransform<64, 320> p1b, p2b;
  // try_copy_parameters builds two std::strings per weight for its range message: convert once per parameter set
  static const float* loaded = nullptr;
  if (loaded != w[0]) try {
    loaded = nullptr;
    fc0.try_copy_parameters(FakeAffine{768, 64, {w[0]}, {b[0]}});
    fc1.try_copy_parameters(FakeAffine{64, 64, {w[1]}, {b[1]}});
    v2.try_copy_parameters(FakeAffine{64, 32, {w[2]}, {b[2]}});
    v3.try_copy_parameters(FakeAffine{32, 1, {w[3]}, {b[3]}});
    p1a.try_copy_parameters(FakeAffine{64, 64, {w[4]}, {b[4]}});
    p1b.try_copy_parameters(FakeAffine{64, 320, {w[5]}, {b[5]}});
    p2a.try_copy_parameters(FakeAffine{64, 64, {w[6]}, {b[6]}});
    p2b.try_copy_parameters(FakeAffine{64, 320, {w[7]}, {b[7]}});
    loaded = w[0];
  } catch (const std::exception&) {
    return -1e9f;
  }
  alignas(64) uint8_t in[768];
  memcpy(in, input768, 768);
  alignas(64) int32_t o0[64], o1[64], ov2[32], ov3[32], op1[64], op2[64];
  alignas(64) uint8_t a0[64] = {}, a1[64] = {}, av2[32] = {}, ap1[64] = {}, ap2[64] = {};
  fc0.propagate(in, o0);
  ac0.propagate(o0, a0);
  fc1.propagate(a0, o1);
  ac1.propagate(o1, a1);
  v2.propagate(a1, ov2);
  vac2.propagate(ov2, av2);
  v3.propagate(av2, ov3);
  p1a.propagate(a1, op1);
  p1ac.propagate(op1, ap1);
  p2a.propagate(a1, op2);
  p2ac.propagate(op2, ap2);
  constexpr float conversion = 127 * (1 << 6);
  for (int i = 0; i < m; i++) p1[i] = p1b.propagate_single(ap1, idx1[i]) / conversion;
  for (int i = 0; i < n; i++) p2[i] = p2b.propagate_single(ap2, idx2[i]) / conversion;
  return ov3[0] / conversion;
}
}
