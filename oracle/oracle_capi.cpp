// ORACLE — TEST INFRASTRUCTURE ONLY (see gen1_oracle.hpp). C entry points for ctypes.
#include <atomic>
#include <thread>
#include <vector>

#include "gen1_oracle.hpp"
#include "oak_oracle.hpp"
#include "nn_oracle.hpp"

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

// The same two with the binary protocol log (common/protocol.zig:313-863) appended to buf[*pos..cap); a log that does
// not fit returns 0b1100 (bindings/c.zig:6,181) with the battle left wherever the update was when the write failed.
uint8_t oracle_gen1_update_log(uint8_t* battle, uint8_t c1, uint8_t c2, Options* o, uint8_t* buf, size_t cap, size_t* pos) {
  LogBuf lg{buf, cap, *pos};
  OakLogEngine e(*reinterpret_cast<Battle*>(battle), *o, nullptr, &lg);
  uint8_t r;
  try {
    r = e.update(c1, c2);
  } catch (const LogOverflow&) {
    r = RESULT_LOG_OVERFLOW;
  }
  *pos = lg.pos;
  return r;
}
uint8_t oracle_gen1_update_kat_log(uint8_t* battle, uint8_t c1, uint8_t c2, Options* o, const uint32_t* rolls, size_t len,
                                   size_t* index, int* overrun, uint8_t* buf, size_t cap, size_t* pos) {
  Rolls32 fx{rolls, len, *index, false};
  LogBuf lg{buf, cap, *pos};
  KatLogEngine e(*reinterpret_cast<Battle*>(battle), *o, &fx, &lg);
  uint8_t r;
  try {
    r = e.update(c1, c2);
  } catch (const LogOverflow&) {
    r = RESULT_LOG_OVERFLOW;
  }
  *index = fx.index;
  *overrun = fx.overrun;
  *pos = lg.pos;
  return r;
}

// libpkmn-default-flag update with the live PRNG and the protocol log
uint8_t oracle_gen1_update_default_log(uint8_t* battle, uint8_t c1, uint8_t c2, Options* o, uint8_t* buf, size_t cap, size_t* pos) {
  LogBuf lg{buf, cap, *pos};
  Engine<true, true, true, false, false, true> e(*reinterpret_cast<Battle*>(battle), *o, nullptr, &lg);
  uint8_t r;
  try {
    r = e.update(c1, c2);
  } catch (const LogOverflow&) {
    r = RESULT_LOG_OVERFLOW;
  }
  *pos = lg.pos;
  return r;
}

// libpkmn-default-flag update with the live PRNG (used to diff flag sets on roll-free scenarios).
uint8_t oracle_gen1_update_default(uint8_t* battle, uint8_t c1, uint8_t c2, Options* o) {
  DefaultEngine e(*reinterpret_cast<Battle*>(battle), *o);
  return e.update(c1, c2);
}

// One update under a chosen flag set with the live PRNG, reporting the flag-sensitive sites it reached.
// which = 0: Oak's flags {miss=false, advance=false, ebc=false, key=true}; 1: libpkmn's defaults {miss, advance, ebc, key=false}.
// counters[0] = advance calls (counted whether or not the build honours them), [1] = hit rolls that landed on the accuracy.
uint8_t oracle_gen1_update_flags(uint8_t* battle, uint8_t c1, uint8_t c2, Options* o, int which, uint32_t* counters) {
  uint8_t r;
  if (which == 0) {
    OakEngine e(*reinterpret_cast<Battle*>(battle), *o);
    r = e.update(c1, c2);
    counters[0] = e.n_advance;
    counters[1] = e.n_hit_edge;
  } else {
    DefaultEngine e(*reinterpret_cast<Battle*>(battle), *o);
    r = e.update(c1, c2);
    counters[0] = e.n_advance;
    counters[1] = e.n_hit_edge;
  }
  return r;
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

// ---- leaf evaluation (nn_oracle.hpp)
uint32_t oracle_enc_pokemon_sparse(const uint8_t* pokemon24, uint8_t sleep, float* values, uint16_t* indices) {
  Pokemon p;
  memcpy(&p, pokemon24, 24);
  Sparse sp;
  uint16_t off = 0;
  enc_pokemon(p, sleep, sp, off);
  memcpy(values, sp.v, sp.n * 4);
  memcpy(indices, sp.i, sp.n * 2);
  return sp.n;
}
uint32_t oracle_enc_active_pokemon_sparse(const uint8_t* pokemon24, const uint8_t* active32, uint32_t duration, float* values,
                                          uint16_t* indices) {
  Pokemon p;
  Active a;
  memcpy(&p, pokemon24, 24);
  memcpy(&a, active32, 32);
  Sparse sp;
  enc_active_pokemon(p, a, duration, sp);
  memcpy(values, sp.v, sp.n * 4);
  memcpy(indices, sp.i, sp.n * 2);
  return sp.n;
}
uint8_t oracle_pokemon_key(const uint8_t* pokemon24, uint8_t sleep) {
  Pokemon p;
  memcpy(&p, pokemon24, 24);
  return pokemon_key(p, sleep);
}
uint16_t oracle_policy_index(const uint8_t* side184, uint8_t choice) {
  Side s;
  memcpy(&s, side184, 184);
  return policy_index(s, choice);
}
void* oracle_net_create(const uint8_t* data, size_t len) {
  BattleNet* n = new BattleNet();
  if (!n->load(data, len)) {
    delete n;
    return nullptr;
  }
  return n;
}
void oracle_net_destroy(void* h) { delete static_cast<BattleNet*>(h); }
void oracle_net_dims(void* h, uint32_t* out7) {
  const BattleNet* n = static_cast<BattleNet*>(h);
  const uint32_t v[7] = {n->pk0.out, n->pk1.out, n->ac0.out, n->ac1.out, n->fc0.out, n->v2.out, n->p1a.out};
  memcpy(out7, v, sizeof(v));
}
// value_policy_inference for explicit choice lists; embedding_out nullable (2 * side_dim floats)
float oracle_net_eval(void* h, const uint8_t* battle, const uint8_t* durations, int m, int n, const uint8_t* c1, const uint8_t* c2,
                      float* l1, float* l2, float* embedding_out) {
  const BattleNet* net = static_cast<BattleNet*>(h);
  std::vector<float> emb;
  const float v = net->eval(*reinterpret_cast<const Battle*>(battle), durations, m, n, c1, c2, l1, l2, embedding_out ? &emb : nullptr);
  if (embedding_out) memcpy(embedding_out, emb.data(), emb.size() * 4);
  return v;
}
// Batched leaf evaluation as the GPU kernel defines it: legal choices are derived from each battle's
// result byte (choices x2), then value + logits of those choices. logits[n][2][9], counts[n][2].
void oracle_net_eval_batch(void* h, const uint8_t* battles, const uint8_t* durations, const uint8_t* results, size_t n, int threads,
                           float* values, float* logits, uint8_t* counts) {
  const BattleNet* net = static_cast<BattleNet*>(h);
  std::atomic<size_t> next{0};
  auto work = [&]() {
    for (;;) {
      const size_t i0 = next.fetch_add(64);
      if (i0 >= n) return;
      const size_t i1 = i0 + 64 < n ? i0 + 64 : n;
      for (size_t i = i0; i < i1; i++) {
        Battle b;
        memcpy(&b, battles + i * 384, 384);
        Options o{};
        OakEngine e(b, o);
        uint8_t c1[9], c2[9];
        const int m = e.choices(P1, (results[i] >> 4) & 3, c1);
        const int k = e.choices(P2, (results[i] >> 6) & 3, c2);
        float* lo = logits + i * 18;
        for (int j = 0; j < 18; j++) lo[j] = 0.0f;
        values[i] = net->eval(b, durations + i * 8, m, k, c1, c2, lo, lo + 9);
        counts[2 * i] = (uint8_t)m;
        counts[2 * i + 1] = (uint8_t)k;
      }
    }
  };
  std::vector<std::thread> pool;
  for (int t = 0; t < (threads < 1 ? 1 : threads); t++) pool.emplace_back(work);
  for (auto& t : pool) t.join();
}
}

extern "C" {
// The same batch through the reference's memoised path (CachedBattleNet). group[i] identifies the battle (team pair) position i
// comes from; positions must be sorted by group. Each thread takes whole groups: fill_cache from the group's first position
// (what a search does once at its root, network.h:48-50), then every position of the group through the memo, `repeat` times.
void oracle_net_eval_cached_batch(void* h, const uint8_t* battles, const uint8_t* durations, const uint8_t* results,
                                  const uint32_t* group, size_t n, int threads, int repeat, float* values, float* logits,
                                  uint8_t* counts) {
  const BattleNet* net = static_cast<BattleNet*>(h);
  std::vector<size_t> starts;
  for (size_t i = 0; i < n; i++)
    if (i == 0 || group[i] != group[i - 1]) starts.push_back(i);
  starts.push_back(n);
  std::atomic<size_t> next{0};
  auto work = [&]() {
    CachedBattleNet cached(*net);
    for (;;) {
      const size_t g = next.fetch_add(1);
      if (g + 1 >= starts.size()) return;
      Battle root;
      memcpy(&root, battles + starts[g] * 384, 384);
      cached.fill(root);
      // repeat > 1: the group's positions are evaluated `repeat` times under the one fill, i.e. repeat x as many evaluations
      // per root as the batch holds (a search makes thousands per root); the active memo is warm after the first pass, as it
      // is for a search revisiting its nodes
      for (int rep = 0; rep < (repeat < 1 ? 1 : repeat); rep++)
      for (size_t i = starts[g]; i < starts[g + 1]; i++) {
        Battle b;
        memcpy(&b, battles + i * 384, 384);
        Options o{};
        OakEngine e(b, o);
        uint8_t c1[9], c2[9];
        const int m = e.choices(P1, (results[i] >> 4) & 3, c1);
        const int k = e.choices(P2, (results[i] >> 6) & 3, c2);
        float* lo = logits + i * 18;
        for (int j = 0; j < 18; j++) lo[j] = 0.0f;
        values[i] = cached.eval(b, durations + i * 8, m, k, c1, c2, lo, lo + 9);
        counts[2 * i] = (uint8_t)m;
        counts[2 * i + 1] = (uint8_t)k;
      }
    }
  };
  std::vector<std::thread> pool;
  for (int t = 0; t < (threads < 1 ? 1 : threads); t++) pool.emplace_back(work);
  for (auto& t : pool) t.join();
}
}

// ---- training frames + policy sampling (frames_oracle.hpp) ----
#include "frames_oracle.hpp"
extern "C" {
uint32_t oracle_frames_write_game(const uint8_t* battle384, uint8_t result, uint32_t n_updates, const uint8_t* k, const uint8_t* choice,
                                  const uint32_t* iterations, const double* emp_value, const double* nash_value, const double* emp,
                                  const double* nash, uint8_t* out) {
  return frames_oracle::write_game(battle384, result, n_updates, k, choice, iterations, emp_value, nash_value, emp, nash, out);
}
int oracle_policy_sample(const float* logits, int k, double temp, double floor_p, double u, float* probs9) {
  return frames_oracle::policy_sample(logits, k, temp, floor_p, u, probs9);
}
}

// ---- tree search (mcts_oracle.hpp) ----
#include "mcts_oracle.hpp"
extern "C" {
void oracle_mcts_run(const uint8_t* battle384, const uint8_t* durations8, uint8_t result, uint64_t seed, uint32_t iterations, float c,
                     uint32_t root_rolls, uint32_t other_rolls, uint32_t* visit81, double* value81, uint64_t* nodes) {
  mcts_oracle::Search s;
  s.c = c;
  s.root_rolls = root_rolls;
  s.other_rolls = other_rolls;
  s.seed = seed;
  s.run(battle384, durations8, result, iterations);
  memcpy(visit81, s.visit, sizeof(s.visit));
  memcpy(value81, s.value, sizeof(s.value));
  if (nodes) *nodes = s.nodes;
}
}

// ---- training rows (train_rows_oracle.hpp) ----
#include "train_rows_oracle.hpp"
extern "C" {
// Plays sample battle `index` (oak_oracle.hpp) to the end or `max_updates` with playout picks from `seed` and writes it
// as one stored game with made-up search outputs (a counter-based stream), through frames_oracle::write_game.
// Returns bytes written (0 if the game did not finish within max_updates: such games are not stored).
uint32_t oracle_play_frame_game(uint64_t index, uint64_t seed, uint32_t max_updates, uint8_t* out) {
  oracle::Battle b;
  uint8_t dur[8], result;
  oracle::oak_sample_battle(index, b, dur, &result);
  uint8_t first[384];
  memcpy(first, &b, 384);
  std::vector<uint8_t> k, choice;
  std::vector<uint32_t> iters;
  std::vector<double> ev, nv, emp, nash;
  uint64_t s = oracle::splitmix64(seed ^ 0xA5A5A5A5ull);
  auto unit = [&]() {
    s = oracle::splitmix64(s);
    return (double)(s >> 11) * (1.0 / 9007199254740992.0);
  };
  uint32_t step = 0;
  while ((result & 15) == 0 && step < max_updates) {
    oracle::Options o{};
    memcpy(o.durations, dur, 8);
    oracle::OakEngine e(b, o);
    uint8_t ch1[9], ch2[9];
    const uint8_t m = e.choices(oracle::P1, (result >> 4) & 3, ch1), n = e.choices(oracle::P2, (result >> 6) & 3, ch2);
    const uint64_t pick = oracle::oak_pick(seed, step);
    const uint8_t c1 = ch1[pick % m], c2 = ch2[(pick >> 32) % n];
    k.push_back(m);
    k.push_back(n);
    choice.push_back(c1);
    choice.push_back(c2);
    iters.push_back((uint32_t)(oracle::splitmix64(s + step) >> 40));
    ev.push_back(unit());
    nv.push_back(unit());
    for (int side = 0; side < 2; side++) {
      double pe[9] = {}, pn[9] = {}, se = 0, sn = 0;
      const int cnt = side ? n : m;
      for (int i = 0; i < cnt; i++) se += pe[i] = unit() + 1e-3, sn += pn[i] = unit() + 1e-3;
      for (int i = 0; i < 9; i++) emp.push_back(pe[i] / se), nash.push_back(pn[i] / sn);
    }
    result = e.update(c1, c2);
    memcpy(dur, o.durations, 8);
    step++;
  }
  if ((result & 15) == 0) return 0;
  return frames_oracle::write_game(first, result, step, k.data(), choice.data(), iters.data(), ev.data(), nv.data(), emp.data(),
                                   nash.data(), out);
}

// Every game of a .battle.data stream -> rows (game g's rows start at the running sum of n_updates). `threads` host
// threads over games. Returns the number of rows, or -(g + 1) for the first game whose replay fails.
int64_t oracle_training_rows(const uint8_t* stream, uint64_t n_bytes, int threads, float* pokemon, float* active, float* hp,
                             int64_t* choice_indices, uint8_t* k, uint8_t* choice, uint32_t* iterations, float* emp, float* nash, float* ev,
                             float* nv, float* score) {
  std::vector<uint64_t> off, first;
  uint64_t at = 0, rows = 0;
  while (at + 391 <= n_bytes) {
    uint32_t nb;
    uint16_t nu;
    memcpy(&nb, stream + at, 4);
    memcpy(&nu, stream + at + 4, 2);
    if (nb < 391 || at + nb > n_bytes) break;
    off.push_back(at);
    first.push_back(rows);
    rows += nu;
    at += nb;
  }
  if (at != n_bytes) return INT64_MIN;
  rows_oracle::Rows r{pokemon, active, hp, choice_indices, k, choice, iterations, emp, nash, ev, nv, score};
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  if (threads <= 0) threads = 1;
  std::atomic<int64_t> bad{0};
  std::atomic<size_t> next{0};
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++)
    pool.emplace_back([&]() {
      for (;;) {
        const size_t g = next.fetch_add(1);
        if (g >= off.size()) return;
        if (rows_oracle::build_rows(stream + off[g], r, first[g]) < 0) {
          int64_t expect = 0;
          bad.compare_exchange_strong(expect, -(int64_t)(g + 1));
        }
      }
    });
  for (auto& t : pool) t.join();
  return bad.load() ? bad.load() : (int64_t)rows;
}
}

// Uncompressed frames of every game of a stream (rows_oracle::build_frames). Returns rows or -(g + 1).
extern "C" int64_t oracle_battle_frames(const uint8_t* stream, uint64_t n_bytes, uint8_t* battle, uint8_t* durations, uint8_t* result,
                                        uint8_t* legal) {
  uint64_t at = 0, rows = 0, g = 0;
  while (at + 391 <= n_bytes) {
    uint32_t nb;
    memcpy(&nb, stream + at, 4);
    if (nb < 391 || at + nb > n_bytes) break;
    const int got = rows_oracle::build_frames(stream + at, rows, battle, durations, result, legal);
    if (got < 0) return -(int64_t)(g + 1);
    rows += (uint64_t)got;
    at += nb;
    g++;
  }
  return at == n_bytes ? (int64_t)rows : INT64_MIN;
}

// ---- "discrete" evaluation (nn_oracle.hpp: eval_discrete) ----
extern "C" {
// the integer main net alone on a given u8 battle embedding (pins against oracle/_ref: ref_quantized_main)
float oracle_net_main_discrete(void* h, const uint8_t* q, int m, int n, const uint16_t* idx1, const uint16_t* idx2, float* l1, float* l2) {
  return static_cast<BattleNet*>(h)->main_discrete(q, m, n, idx1, idx2, l1, l2);
}
// batched, as oracle_net_eval_batch; embeddings_out nullable ([n][fc0.in] u8)
void oracle_net_eval_discrete_batch(void* h, const uint8_t* battles, const uint8_t* durations, const uint8_t* results, size_t n, int threads,
                                    float* values, float* logits, uint8_t* counts, uint8_t* embeddings_out) {
  const BattleNet* net = static_cast<BattleNet*>(h);
  std::atomic<size_t> next{0};
  auto work = [&]() {
    for (;;) {
      const size_t i0 = next.fetch_add(64);
      if (i0 >= n) return;
      const size_t i1 = i0 + 64 < n ? i0 + 64 : n;
      for (size_t i = i0; i < i1; i++) {
        Battle b;
        memcpy(&b, battles + i * 384, 384);
        Options o{};
        OakEngine e(b, o);
        uint8_t c1[9], c2[9];
        const int m = e.choices(P1, (results[i] >> 4) & 3, c1);
        const int k = e.choices(P2, (results[i] >> 6) & 3, c2);
        float* lo = logits + i * 18;
        for (int j = 0; j < 18; j++) lo[j] = 0.0f;
        std::vector<uint8_t> q;
        values[i] = net->eval_discrete(b, durations + i * 8, m, k, c1, c2, lo, lo + 9, embeddings_out ? &q : nullptr);
        if (embeddings_out) memcpy(embeddings_out + i * net->fc0.in, q.data(), q.size());
        counts[2 * i] = (uint8_t)m;
        counts[2 * i + 1] = (uint8_t)k;
      }
    }
  };
  std::vector<std::thread> pool;
  for (int t = 0; t < (threads < 1 ? 1 : threads); t++) pool.emplace_back(work);
  for (auto& t : pool) t.join();
}
}

// tree search with network-value leaves (mcts_oracle.hpp: Search::net)
extern "C" void oracle_mcts_run_net(void* net, int discrete, const uint8_t* battle384, const uint8_t* durations8, uint8_t result, uint64_t seed,
                                    uint32_t iterations, float c, uint32_t root_rolls, uint32_t other_rolls, uint32_t* visit81, double* value81,
                                    uint64_t* nodes) {
  mcts_oracle::Search s;
  s.c = c;
  s.root_rolls = root_rolls;
  s.other_rolls = other_rolls;
  s.seed = seed;
  s.net = static_cast<BattleNet*>(net);
  s.net_discrete = discrete != 0;
  s.run(battle384, durations8, result, iterations);
  memcpy(visit81, s.visit, sizeof(s.visit));
  memcpy(value81, s.value, sizeof(s.value));
  if (nodes) *nodes = s.nodes;
}

// the same search with any of the reference's bandits (kind 0 ucb, 1 ucb1, 2 pucb, 3 exp3, 4 pexp3; f1 = c or gamma, f2 = alpha);
// net may be null (MonteCarlo leaves; not with the contextual kinds)
extern "C" void oracle_mcts_run_bandit(void* net, int discrete, const uint8_t* battle384, const uint8_t* durations8, uint8_t result,
                                       uint64_t seed, uint32_t iterations, int kind, float f1, float f2, uint32_t root_rolls,
                                       uint32_t other_rolls, uint32_t* visit81, double* value81, uint64_t* nodes) {
  mcts_oracle::Search s;
  s.params.kind = kind;
  s.params.c = s.params.gamma = f1;
  s.params.alpha = f2;
  s.params_set = true;
  s.root_rolls = root_rolls;
  s.other_rolls = other_rolls;
  s.seed = seed;
  s.net = static_cast<BattleNet*>(net);
  s.net_discrete = discrete != 0;
  s.run(battle384, durations8, result, iterations);
  memcpy(visit81, s.visit, sizeof(s.visit));
  memcpy(value81, s.value, sizeof(s.value));
  if (nodes) *nodes = s.nodes;
}
