// TEST-ONLY: compiles the CUDA engine source (oak_b200/csrc/gen1_engine.cuh) with the host compiler
// so that it can be differential-tested against the oracle in the GPU-less build container. The
// record is scattered with STRIDE = 3 words (a 3-thread "column tile" of which we use column 1), so
// the column addressing (word handles + folded byte offsets) is exercised exactly as on the GPU.
// Nothing in the product links this.
#include <cstring>

#include "../../oak_b200/csrc/gen1_engine.cuh"

using namespace oakb200;

static const int ST = 3;
struct Tile {
  uint32_t w[96 * ST];
  explicit Tile(const uint8_t* battle) {
    memset(w, 0xAB, sizeof(w));
    for (int i = 0; i < 96; i++) memcpy(&w[i * ST + 1], battle + 4 * i, 4);
  }
  void store(uint8_t* battle) const {
    for (int i = 0; i < 96; i++) memcpy(battle + 4 * i, &w[i * ST + 1], 4);
  }
  uint8_t* column() { return reinterpret_cast<uint8_t*>(&w[1]); }
};

template <bool CALC, bool CHANCE>
static uint8_t run_update(uint8_t* battle, uint8_t c1, uint8_t c2, uint8_t* durations, uint8_t* actions, const uint8_t* overrides) {
  Engine<ST, CALC, CHANCE> e;
  Tile tile(battle);
  uint32_t d[2];
  memcpy(d, durations, 8);
  e.begin(tile.column(), d[0], d[1]);
  if (CALC && overrides) memcpy(e.ovr, overrides, 16);
  const uint32_t r = e.update(c1, c2);
  e.end();
  tile.store(battle);
  d[0] = e.dur[0];
  d[1] = e.dur[1];
  memcpy(durations, d, 8);
  if (actions) memcpy(actions, e.act, 16);
  return (uint8_t)r;
}

// SPLIT layout of the engine (gen1_state.cuh): 48-word column scattered with STRIDE = 3, constant fields read from a frozen
// copy of the record as the kernels read them from global memory; the dynamic fields are merged back afterwards.
template <bool CALC, bool CHANCE>
static uint8_t run_update_split(uint8_t* battle, uint8_t c1, uint8_t c2, uint8_t* durations, uint8_t* actions, const uint8_t* overrides) {
  Engine<ST, CALC, CHANCE, false, true> e;
  alignas(8) uint32_t frozen[96];
  memcpy(frozen, battle, 384);
  uint32_t w[48 * ST];
  memset(w, 0xAB, sizeof(w));
  split_fill<ST, PlainLoad>(w + 1, frozen);
  uint32_t d[2];
  memcpy(d, durations, 8);
  e.begin(reinterpret_cast<uint8_t*>(w + 1), d[0], d[1]);
  e.m.set_const(reinterpret_cast<const uint8_t*>(frozen));
  if (CALC && overrides) memcpy(e.ovr, overrides, 16);
  const uint32_t r = e.update(c1, c2);
  e.end();
  alignas(8) uint32_t out[96];
  memcpy(out, frozen, 384);
  split_merge<ST, PlainLoad>(w + 1, out);
  memcpy(battle, out, 384);
  d[0] = e.dur[0];
  d[1] = e.dur[1];
  memcpy(durations, d, 8);
  if (actions) memcpy(actions, e.act, 16);
  return (uint8_t)r;
}
extern "C" {
uint8_t shim_update(uint8_t* battle, uint8_t c1, uint8_t c2, uint8_t* durations, uint8_t* actions, const uint8_t* overrides) {
  return overrides ? run_update<true, true>(battle, c1, c2, durations, actions, overrides)
                   : run_update<false, true>(battle, c1, c2, durations, actions, nullptr);
}
uint8_t shim_update_nochance(uint8_t* battle, uint8_t c1, uint8_t c2) {
  uint8_t d[8] = {0};
  return run_update<false, false>(battle, c1, c2, d, nullptr, nullptr);
}
uint8_t shim_update_split(uint8_t* battle, uint8_t c1, uint8_t c2, uint8_t* durations, uint8_t* actions, const uint8_t* overrides) {
  return overrides ? run_update_split<true, true>(battle, c1, c2, durations, actions, overrides)
                   : run_update_split<false, true>(battle, c1, c2, durations, actions, nullptr);
}
uint8_t shim_choices_split(const uint8_t* battle, int player, int request, uint8_t* out) {
  Engine<ST, false, true, false, true> e;
  alignas(8) uint32_t frozen[96];
  memcpy(frozen, battle, 384);
  uint32_t w[48 * ST];
  split_fill<ST, PlainLoad>(w + 1, frozen);
  e.begin(reinterpret_cast<uint8_t*>(w + 1), 0, 0);
  e.m.set_const(reinterpret_cast<const uint8_t*>(frozen));
  int n;
  const uint64_t list = e.choices(player, request, &n);
  for (int i = 0; i < n; i++) out[i] = (uint8_t)Engine<ST, false, true, false, true>::choice_at(list, i);
  return (uint8_t)n;
}
// LOG build of the engine: the update's protocol log goes to log[0..cap); *len = bytes written; an overflow gives 0b1100
uint8_t shim_update_log(uint8_t* battle, uint8_t c1, uint8_t c2, uint8_t* durations, uint8_t* actions, const uint8_t* overrides,
                        uint8_t* log, uint32_t cap, uint32_t* len) {
  Engine<ST, true, true, true> e;
  Tile tile(battle);
  uint32_t d[2];
  memcpy(d, durations, 8);
  e.begin(tile.column(), d[0], d[1]);
  e.sink.buf = log;
  e.sink.cap = cap;
  e.sink.pos = 0;
  e.sink.overflow = 0;
  if (overrides) memcpy(e.ovr, overrides, 16);
  uint32_t r = e.update(c1, c2);
  if (e.sink.overflow) r = RESULT_LOG_OVERFLOW;
  e.end();
  tile.store(battle);
  d[0] = e.dur[0];
  d[1] = e.dur[1];
  memcpy(durations, d, 8);
  if (actions) memcpy(actions, e.act, 16);
  *len = e.sink.pos;
  return (uint8_t)r;
}
uint8_t shim_choices(const uint8_t* battle, int player, int request, uint8_t* out) {
  Engine<ST, false, true> e;
  Tile tile(battle);
  e.begin(tile.column(), 0, 0);
  int n;
  const uint64_t list = e.choices(player, request, &n);
  for (int i = 0; i < n; i++) {
    const uint32_t c = Engine<ST, false, true>::choice_at(list, i);
    out[i] = (uint8_t)(((c >> 2) << 2) | (c & 3));
  }
  return (uint8_t)n;
}
}
