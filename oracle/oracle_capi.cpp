// ORACLE — TEST INFRASTRUCTURE ONLY (see gen1_oracle.hpp). C entry points for ctypes.
#include <atomic>
#include <thread>
#include <vector>

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

// Batch of random playouts on `threads` host threads (one battle stream per std::thread, as the
// reference's generate.cc:527-530 does). battles/durations/results are [n][384]/[n][8]/[n] AoS and are
// advanced in place to their terminal state. Used as the checker at full size and as the timed CPU
// baseline (bench.py cpu_baseline / --impl reference).
void oracle_oak_rollout_batch(uint8_t* battles, uint8_t* durations, uint8_t* results, const uint64_t* seeds, size_t n,
                              uint32_t max_steps, int threads, uint8_t* score2, uint16_t* turns, uint32_t* steps) {
  std::atomic<size_t> next{0};
  auto work = [&]() {
    for (;;) {
      const size_t i0 = next.fetch_add(256);
      if (i0 >= n) return;
      const size_t i1 = i0 + 256 < n ? i0 + 256 : n;
      for (size_t i = i0; i < i1; i++) {
        Battle& b = *reinterpret_cast<Battle*>(battles + i * 384);
        const uint32_t st = oak_rollout(b, durations + i * 8, results + i, seeds[i], max_steps);
        const uint8_t t = results[i] & 15;
        if (score2) score2[i] = t == RES_WIN ? 2 : (t == RES_TIE ? 1 : (t == RES_LOSE ? 0 : 255));
        if (turns) turns[i] = b.turn;
        if (steps) steps[i] = st;
      }
    }
  };
  if (threads <= 1) {
    work();
    return;
  }
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++) pool.emplace_back(work);
  for (auto& t : pool) t.join();
}

// Synthetic battles first..first+n built on `threads` host threads.
void oracle_oak_sample_batch(uint64_t first, size_t n, int threads, uint8_t* battles, uint8_t* durations, uint8_t* results) {
  std::atomic<size_t> next{0};
  auto work = [&]() {
    for (;;) {
      const size_t i0 = next.fetch_add(1024);
      if (i0 >= n) return;
      const size_t i1 = i0 + 1024 < n ? i0 + 1024 : n;
      for (size_t i = i0; i < i1; i++)
        oak_sample_battle(first + i, *reinterpret_cast<Battle*>(battles + i * 384), durations + i * 8, results + i);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 0; t < (threads < 1 ? 1 : threads); t++) pool.emplace_back(work);
  for (auto& t : pool) t.join();
}
}
