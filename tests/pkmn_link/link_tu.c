/* TEST-ONLY translation unit compiled against the REFERENCE's own header (engine/src/include/pkmn.h, given with -I) and
 * linked with -loakb200: shows that code written for libpkmn binds to the B200 library with no source change.
 *   (1) tu_symbols(): takes the address of every name libpkmn exports (engine/src/lib/c.zig:4-46), data symbols included
 *       (PKMN_GEN1_MAX_CHOICES is what cpp/include/search/mcts.h:163,166,338,... sizes its choice arrays with);
 *   (2) tu_playout(): the reference's playout loop (cpp/include/search/mcts.h:448-496; same shape in cpp/src/dur.cc:11-34,
 *       cpp/src/log.cc:464-498) through the scalar entry points with chance tracking and the protocol log on, recording a
 *       per-update trace for the test to compare with the oracle. The pick stream is SURVEY §8d's counter-based one.
 * Built by tests/pkmn_link/build.py where /root/reference exists; the .so travels to the GPU box. */
#include <string.h>

#include "pkmn.h"

static uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

size_t tu_symbols(const void **out, size_t cap) {
  const void *syms[] = {
      (const void *)&PKMN_OPTIONS, (const void *)&PKMN_MAX_CHOICES, (const void *)&PKMN_CHOICES_SIZE, (const void *)&PKMN_MAX_LOGS,
      (const void *)&PKMN_LOGS_SIZE, (const void *)&PKMN_GEN1_MAX_CHOICES, (const void *)&PKMN_GEN1_CHOICES_SIZE,
      (const void *)&PKMN_GEN1_MAX_LOGS, (const void *)&PKMN_GEN1_LOGS_SIZE, (const void *)pkmn_choice_init,
      (const void *)pkmn_choice_type, (const void *)pkmn_choice_data, (const void *)pkmn_result_type, (const void *)pkmn_result_p1,
      (const void *)pkmn_result_p2, (const void *)pkmn_error, (const void *)pkmn_psrng_init, (const void *)pkmn_psrng_next,
      (const void *)pkmn_rational_init, (const void *)pkmn_rational_reduce, (const void *)pkmn_rational_numerator,
      (const void *)pkmn_rational_denominator, (const void *)pkmn_gen1_battle_options_set,
      (const void *)pkmn_gen1_battle_options_chance_probability, (const void *)pkmn_gen1_battle_options_chance_actions,
      (const void *)pkmn_gen1_battle_options_chance_durations, (const void *)pkmn_gen1_battle_options_calc_summaries,
      (const void *)pkmn_gen1_battle_update, (const void *)pkmn_gen1_battle_choices};
  const size_t n = sizeof(syms) / sizeof(syms[0]);
  for (size_t i = 0; i < n && i < cap; i++) out[i] = syms[i];
  return n;
}

/* values a caller reads through the header's types: sizes of the enums the accessors return, the exported constants */
void tu_constants(size_t *out) {
  out[0] = PKMN_GEN1_MAX_CHOICES;
  out[1] = PKMN_GEN1_CHOICES_SIZE;
  out[2] = PKMN_GEN1_MAX_LOGS;
  out[3] = PKMN_GEN1_LOGS_SIZE;
  out[4] = PKMN_MAX_CHOICES;
  out[5] = PKMN_MAX_LOGS;
  out[6] = (size_t)PKMN_OPTIONS.showdown | (size_t)PKMN_OPTIONS.log << 1 | (size_t)PKMN_OPTIONS.chance << 2 | (size_t)PKMN_OPTIONS.calc << 3;
  out[7] = sizeof(pkmn_result_kind);
  out[8] = (size_t)pkmn_result_type((pkmn_result)0x93);  /* full-width enum return: 3, not 3 | garbage << 8 */
  out[9] = (size_t)pkmn_result_p1((pkmn_result)0x93);    /* 1 */
  out[10] = (size_t)pkmn_result_p2((pkmn_result)0x93);   /* 2 */
  out[11] = (size_t)pkmn_choice_init(PKMN_CHOICE_SWITCH, 5);
  out[12] = (size_t)pkmn_error((pkmn_result)0x0C) | (size_t)pkmn_error((pkmn_result)0x04) << 1;
  pkmn_psrng rng;
  pkmn_psrng_init(&rng, 0x0001000200030004ull); /* common/rng.zig:73-77: 121 then 226 after `* 256 >> 32` */
  out[13] = (size_t)(((uint64_t)pkmn_psrng_next(&rng) * 256) >> 32);
  out[14] = (size_t)(((uint64_t)pkmn_psrng_next(&rng) * 256) >> 32);
}

/* One trace record per update: c1, c2, result, log length (u16), battle[384], durations[8], actions[16], log[180] = 593 bytes */
#define TU_RECORD 593
size_t tu_playout(const uint8_t *battle_in, const uint8_t *durations_in, uint8_t result_in, uint64_t seed, size_t max_steps,
                  uint8_t *trace, size_t trace_cap, size_t log_len) {
  pkmn_gen1_battle battle;
  pkmn_gen1_battle_options options;
  pkmn_gen1_chance_options chance;
  pkmn_gen1_log_options log;
  uint8_t buf[256];
  pkmn_choice ch1[16], ch2[16];
  memcpy(battle.bytes, battle_in, PKMN_GEN1_BATTLE_SIZE);
  memset(&chance, 0, sizeof(chance));
  memcpy(chance.durations.bytes, durations_in, PKMN_GEN1_CHANCE_DURATIONS_SIZE);
  log.buf = buf;
  log.len = log_len;
  pkmn_gen1_battle_options_set(&options, &log, &chance, NULL);
  pkmn_result result = result_in;
  size_t step = 0;
  while (!pkmn_result_type(result) && step < max_steps && (step + 1) * TU_RECORD <= trace_cap) {
    const uint8_t m = pkmn_gen1_battle_choices(&battle, PKMN_PLAYER_P1, pkmn_result_p1(result), ch1, PKMN_GEN1_MAX_CHOICES);
    const uint8_t n = pkmn_gen1_battle_choices(&battle, PKMN_PLAYER_P2, pkmn_result_p2(result), ch2, PKMN_GEN1_MAX_CHOICES);
    const uint64_t pick = splitmix64(seed ^ (0xD1B54A32D192ED03ull * (uint64_t)step));
    const pkmn_choice c1 = ch1[pick % m], c2 = ch2[(pick >> 32) % n];
    pkmn_gen1_battle_options_set(&options, NULL, NULL, NULL); /* mcts.h:478: resets log position, actions, overrides */
    memset(buf, 0xFF, sizeof(buf));
    result = pkmn_gen1_battle_update(&battle, c1, c2, &options);
    uint8_t *r = trace + step * TU_RECORD;
    size_t used = 0;
    if (!pkmn_error(result)) { /* every update's log ends in an ArgType.None (0) terminator: the last byte that is not the 0xFF fill */
      used = log_len < sizeof(buf) ? log_len : sizeof(buf);
      while (used > 0 && buf[used - 1] == 0xFF) used--;
    }
    r[0] = c1;
    r[1] = c2;
    r[2] = result;
    r[3] = (uint8_t)(used & 255);
    r[4] = (uint8_t)(used >> 8);
    memcpy(r + 5, battle.bytes, 384);
    memcpy(r + 389, pkmn_gen1_battle_options_chance_durations(&options)->bytes, 8);
    memcpy(r + 397, pkmn_gen1_battle_options_chance_actions(&options)->bytes, 16);
    memcpy(r + 413, buf, 180);
    step++;
    if (pkmn_error(result)) break;
  }
  return step;
}
