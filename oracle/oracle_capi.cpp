// ORACLE — TEST INFRASTRUCTURE ONLY (see gen1_oracle.hpp). C entry points for ctypes.
#include "gen1_oracle.hpp"
#include "oak_oracle.hpp"

using namespace oracle;

extern "C" {

// Oak-flag update (pkmn_gen1_battle_update equivalent; reference bindings/c.zig:174-184).
// The caller clears o->actions / o->overrides itself, as pkmn_gen1_battle_options_set does
// (bindings/c.zig:116-148, chance.zig:367-373).
uint8_t oracle_gen1_update(uint8_t* battle, uint8_t c1, uint8_t c2, Options* o) {
  OakEngine e(*reinterpret_cast<Battle*>(battle), *o);
  return e.update(c1, c2);
}

// pkmn_gen1_battle_choices equivalent (bindings/c.zig:186-198). Flag-independent.
uint8_t oracle_gen1_choices(const uint8_t* battle, uint8_t player, uint8_t request, uint8_t* out) {
  Options o{};
  OakEngine e(*const_cast<Battle*>(reinterpret_cast<const Battle*>(battle)), o);
  return e.choices(player, request, out);
}

// libpkmn-default-flag update driven by a scripted roll list (gen1/test.zig harness :10415-10554).
uint8_t oracle_gen1_update_kat(uint8_t* battle, uint8_t c1, uint8_t c2, Options* o, const uint32_t* rolls,
                               size_t len, size_t* index, int* overrun) {
  Rolls32 fx{rolls, len, *index, false};
  KatEngine e(*reinterpret_cast<Battle*>(battle), *o, &fx);
  const uint8_t r = e.update(c1, c2);
  *index = fx.index;
  *overrun = fx.overrun;
  return r;
}

// libpkmn-default-flag update with the live PRNG (used to diff flag sets on roll-free scenarios).
uint8_t oracle_gen1_update_default(uint8_t* battle, uint8_t c1, uint8_t c2, Options* o) {
  DefaultEngine e(*reinterpret_cast<Battle*>(battle), *o);
  return e.update(c1, c2);
}

// PSRNG known-answer helpers (rng.zig:34-59)
uint32_t oracle_psrng_next(uint64_t* seed) {
  *seed = 0x5D588B656C078965ull * *seed + 0x269EC3ull;
  return (uint32_t)(*seed >> 32);
}
uint32_t oracle_psrng_range(uint64_t* seed, uint32_t from, uint32_t to) {
  return (uint32_t)(((uint64_t)oracle_psrng_next(seed) * (to - from)) >> 32) + from;
}

// Oak-level helpers (oak_oracle.hpp)
void oracle_oak_battle_init(const uint8_t* p1_species, const uint8_t* p1_moves, const uint8_t* p2_species,
                            const uint8_t* p2_moves, uint64_t seed, uint8_t* out) {
  oak_battle_init(p1_species, p1_moves, p2_species, p2_moves, seed, *reinterpret_cast<Battle*>(out));
}
void oracle_oak_sample_battle(uint64_t i, uint8_t* battle_out, uint8_t* durations_out, uint8_t* result_out) {
  oak_sample_battle(i, *reinterpret_cast<Battle*>(battle_out), durations_out, result_out);
}
// One synthetic-stream step: choices x2 + pick + options_set(null) + update.
uint8_t oracle_oak_step(uint8_t* battle, uint8_t* durations, uint8_t result, uint64_t pick, uint8_t* actions_out,
                        uint8_t* c1_out, uint8_t* c2_out) {
  return oak_step(*reinterpret_cast<Battle*>(battle), durations, result, pick, actions_out, c1_out, c2_out);
}
// Full random playout of synthetic battle i from state (battle, durations, result).
void oracle_oak_rollout(uint8_t* battle, uint8_t* durations, uint8_t* result, uint64_t seed, uint32_t max_steps,
                        uint32_t* steps_out) {
  *steps_out = oak_rollout(*reinterpret_cast<Battle*>(battle), durations, result, seed, max_steps);
}
void oracle_oak_randomize_hidden(uint8_t* battle, const uint8_t* durations) {
  oak_randomize_hidden_variables(*reinterpret_cast<Battle*>(battle), durations);
}
}
