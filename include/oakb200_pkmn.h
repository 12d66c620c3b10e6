/* libpkmn-compatible scalar entry points, served by the B200 engine.
 *
 * These are the libpkmn C symbols Oak itself references (reference cpp/extern/engine/src/include/pkmn.h; call sites
 * cpp/include/libpkmn/pkmn.h:67-151, cpp/include/search/mcts.h:161-617, cpp/src/dur.cc, cpp/src/log.cc), with the
 * reference's own names, argument meaning and return values, so that code written against pkmn.h links against
 * liboakb200.so unchanged. Each call runs the same CUDA kernels as the batched API (include/oakb200.h) on a batch of
 * one, on a process-wide context bound to device OAKB200_SCALAR_DEVICE (default 0): correct and bit-exact, but a
 * host<->device round trip per call — the batched entry points are the fast path. There is no CPU implementation:
 * without a CUDA device pkmn_gen1_battle_update returns the PKMN_RESULT_ERROR result byte (4) and
 * pkmn_gen1_battle_choices returns 0, states the reference never produces in showdown mode.
 *
 * Engine options are libpkmn's as Oak builds it (dev/libpkmn:9): -Dshowdown -Dchance -Dcalc, no -Dlog.
 * pkmn_gen1_battle_options is caller-owned opaque storage of 128 bytes exactly as in pkmn.h:151; the log facet is
 * accepted and ignored (Oak's build has no -Dlog; its generate/search programs pass NULL).
 */
#ifndef OAKB200_PKMN_H
#define OAKB200_PKMN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint8_t pkmn_choice;  /* type | data << 2, pkmn.h:78-95 */
typedef uint8_t pkmn_result;  /* type | p1 << 4 | p2 << 6, pkmn.h:97-112 */
typedef struct { uint8_t bytes[384]; } pkmn_gen1_battle;          /* pkmn.h:150,170 */
typedef struct { uint8_t bytes[128]; } pkmn_gen1_battle_options;  /* pkmn.h:151,171 */
typedef struct { uint8_t bytes[16]; } pkmn_gen1_chance_actions;   /* pkmn.h:152 */
typedef struct { uint8_t bytes[8]; } pkmn_gen1_chance_durations;  /* pkmn.h:153 */
typedef struct { uint8_t bytes[24]; } pkmn_gen1_calc_overrides;   /* pkmn.h:154 (the engine reads the first 16: c.zig:101-103) */
typedef struct { uint8_t *buf; size_t len; } pkmn_gen1_log_options;                                    /* pkmn.h:190-193 */
typedef struct { uint8_t probability[16]; pkmn_gen1_chance_actions actions; pkmn_gen1_chance_durations durations; }
    pkmn_gen1_chance_options;                                                                            /* pkmn.h:198-202 */
typedef struct { pkmn_gen1_calc_overrides overrides; } pkmn_gen1_calc_options;                         /* pkmn.h:207-209 */

/* pkmn.h:108-112 */
uint8_t pkmn_result_type(pkmn_result result);
uint8_t pkmn_result_p1(pkmn_result result);
uint8_t pkmn_result_p2(pkmn_result result);

/* pkmn.h:225-229; bindings/c.zig:116-148. NULL resets a facet: chance keeps its durations and clears the actions
 * (chance.zig:367-373), calc clears the overrides. */
void pkmn_gen1_battle_options_set(pkmn_gen1_battle_options *options, const pkmn_gen1_log_options *log,
                                  const pkmn_gen1_chance_options *chance, const pkmn_gen1_calc_options *calc);
/* pkmn.h:242-243, 249-250 */
pkmn_gen1_chance_actions *pkmn_gen1_battle_options_chance_actions(const pkmn_gen1_battle_options *options);
pkmn_gen1_chance_durations *pkmn_gen1_battle_options_chance_durations(const pkmn_gen1_battle_options *options);
/* pkmn.h:264-268; bindings/c.zig:174-184. options may be NULL (no chance tracking, no overrides). */
pkmn_result pkmn_gen1_battle_update(pkmn_gen1_battle *battle, pkmn_choice c1, pkmn_choice c2, pkmn_gen1_battle_options *options);
/* pkmn.h:283-288; bindings/c.zig:186-198. Returns the number of choices written (at most min(len, 9)). */
uint8_t pkmn_gen1_battle_choices(const pkmn_gen1_battle *battle, uint8_t player, uint8_t request, pkmn_choice out[], size_t len);

#ifdef __cplusplus
}
#endif
#endif
