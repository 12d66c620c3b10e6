// libpkmn-compatible scalar symbols (include/oakb200_pkmn.h) on top of the batched C ABI: a process-wide context
// with a batch of one. Same kernels as the batched path; no CPU engine is involved.
#include "../../include/oakb200_pkmn.h"

#include <cstdlib>
#include <cstring>
#include <mutex>

#include "../../include/oakb200.h"

namespace {

// layout of our 128-byte pkmn_gen1_battle_options
struct Opts {
  uint8_t actions[16];
  uint8_t durations[8];
  uint8_t overrides[16];
  uint8_t has_calc;
};
static_assert(sizeof(Opts) <= 128, "fits pkmn_gen1_battle_options");

std::mutex g_mu;
oakb200_ctx* g_ctx = nullptr;
bool g_failed = false;

oakb200_ctx* scalar_ctx() {  // call with g_mu held
  if (g_ctx || g_failed) return g_ctx;
  const char* dev = getenv("OAKB200_SCALAR_DEVICE");
  if (oakb200_create(dev ? atoi(dev) : 0, 1, &g_ctx) != OAKB200_OK) {
    if (g_ctx) oakb200_destroy(g_ctx);
    g_ctx = nullptr;
    g_failed = true;
  }
  return g_ctx;
}

}  // namespace

extern "C" {

uint8_t pkmn_result_type(pkmn_result result) { return result & 15u; }
uint8_t pkmn_result_p1(pkmn_result result) { return (result >> 4) & 3u; }
uint8_t pkmn_result_p2(pkmn_result result) { return result >> 6; }

void pkmn_gen1_battle_options_set(pkmn_gen1_battle_options* options, const pkmn_gen1_log_options*, const pkmn_gen1_chance_options* chance,
                                  const pkmn_gen1_calc_options* calc) {
  Opts* o = reinterpret_cast<Opts*>(options->bytes);
  if (chance) {
    memcpy(o->actions, chance->actions.bytes, 16);
    memcpy(o->durations, chance->durations.bytes, 8);
  } else {
    memset(o->actions, 0, 16);
  }
  if (calc) {
    memcpy(o->overrides, calc->overrides.bytes, 16);
    o->has_calc = 1;
  } else {
    memset(o->overrides, 0, 16);
    o->has_calc = 0;
  }
}

pkmn_gen1_chance_actions* pkmn_gen1_battle_options_chance_actions(const pkmn_gen1_battle_options* options) {
  return reinterpret_cast<pkmn_gen1_chance_actions*>(const_cast<uint8_t*>(options->bytes) + offsetof(Opts, actions));
}
pkmn_gen1_chance_durations* pkmn_gen1_battle_options_chance_durations(const pkmn_gen1_battle_options* options) {
  return reinterpret_cast<pkmn_gen1_chance_durations*>(const_cast<uint8_t*>(options->bytes) + offsetof(Opts, durations));
}

pkmn_result pkmn_gen1_battle_update(pkmn_gen1_battle* battle, pkmn_choice c1, pkmn_choice c2, pkmn_gen1_battle_options* options) {
  std::lock_guard<std::mutex> lock(g_mu);
  oakb200_ctx* ctx = scalar_ctx();
  if (!ctx) return 4;  // PKMN_RESULT_ERROR
  Opts* o = options ? reinterpret_cast<Opts*>(options->bytes) : nullptr;
  const uint8_t zero[8] = {0};
  uint8_t result = 0, durations[8], actions[16];
  if (oakb200_upload(ctx, 1, battle->bytes, o ? o->durations : zero, nullptr) != OAKB200_OK) return 4;
  if (oakb200_step(ctx, &c1, &c2, (o && o->has_calc) ? o->overrides : nullptr, &result, actions) != OAKB200_OK) return 4;
  if (oakb200_download(ctx, 1, battle->bytes, durations, nullptr) != OAKB200_OK) return 4;
  if (o) {
    memcpy(o->durations, durations, 8);
    memcpy(o->actions, actions, 16);
  }
  return result;
}

uint8_t pkmn_gen1_battle_choices(const pkmn_gen1_battle* battle, uint8_t player, uint8_t request, pkmn_choice out[], size_t len) {
  if (player > 1 || request > 2 || !out || len == 0) return 0;
  std::lock_guard<std::mutex> lock(g_mu);
  oakb200_ctx* ctx = scalar_ctx();
  if (!ctx) return 0;
  // the batched call derives both players' request kinds from the result byte: put `request` in this player's field
  const uint8_t result = (uint8_t)(request << (player ? 6 : 4));
  uint8_t choices[18], counts[2];
  if (oakb200_upload(ctx, 1, battle->bytes, nullptr, &result) != OAKB200_OK) return 0;
  if (oakb200_choices(ctx, choices, counts) != OAKB200_OK) return 0;
  uint8_t n = counts[player];
  if (n > len) n = (uint8_t)len;
  memcpy(out, choices + 9 * player, n);
  return n;
}

}  // extern "C"
