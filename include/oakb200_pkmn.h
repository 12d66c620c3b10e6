/* libpkmn's C interface, served by the B200 engine.
 *
 * Every name libpkmn exports (reference cpp/extern/engine/src/lib/c.zig:4-46, declared in
 * cpp/extern/engine/src/include/pkmn.h) with the reference's types, argument meaning and return values, so that code
 * compiled against the reference's own pkmn.h links against liboakb200.so unchanged; tests/test_pkmn_link.py compiles a
 * translation unit against that header and does exactly this. Oak's call sites: cpp/include/libpkmn/pkmn.h:67-151,
 * cpp/include/search/mcts.h:161-617 (incl. the data symbol PKMN_GEN1_MAX_CHOICES at :163,166,338,...), cpp/src/dur.cc,
 * cpp/src/log.cc, cpp/include/util/debug-log.h:13-68.
 *
 * pkmn_gen1_battle_update / _choices run the same CUDA kernels as the batched API (include/oakb200.h) on a batch of one,
 * on a context owned by the CALLING THREAD (created on first use on device OAKB200_SCALAR_DEVICE, default 0), so threads
 * working on disjoint battles do not serialise, as the reference's contract has it. Correct and bit-exact, but one
 * host<->device round trip per call: the batched entry points are the fast path. There is no CPU implementation: without
 * a CUDA device update returns the PKMN_RESULT_ERROR byte (4) and choices returns 0, values the reference never produces
 * in showdown mode.
 *
 * Build options are Oak's (dev/libpkmn:9) plus the log: PKMN_OPTIONS = {showdown, log, chance, calc} all true.
 * Not computed: the chance probability (pkmn_gen1_battle_options_chance_probability keeps what options_set installed,
 * 1/1 after a reset; Oak never reads it) and the calc summaries (zeros).
 */
#ifndef OAKB200_PKMN_H
#define OAKB200_PKMN_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* pkmn.h:50-74 */
typedef struct { bool showdown; bool log; bool chance; bool calc; } pkmn_options;
extern const pkmn_options PKMN_OPTIONS;
extern const size_t PKMN_MAX_CHOICES;  /* 9 */
extern const size_t PKMN_CHOICES_SIZE; /* 16 */
extern const size_t PKMN_MAX_LOGS;     /* 180 */
extern const size_t PKMN_LOGS_SIZE;    /* 256 */

/* pkmn.h:77-131 */
typedef enum pkmn_player { PKMN_PLAYER_P1 = 0, PKMN_PLAYER_P2 = 1 } pkmn_player;
typedef uint8_t pkmn_choice; /* type | data << 2 */
typedef enum pkmn_choice_kind { PKMN_CHOICE_PASS = 0, PKMN_CHOICE_MOVE = 1, PKMN_CHOICE_SWITCH = 2 } pkmn_choice_kind;
pkmn_choice pkmn_choice_init(pkmn_choice_kind type, uint8_t data);
pkmn_choice_kind pkmn_choice_type(pkmn_choice choice);
uint8_t pkmn_choice_data(pkmn_choice choice);
typedef uint8_t pkmn_result; /* type | p1 << 4 | p2 << 6 */
typedef enum pkmn_result_kind {
  PKMN_RESULT_NONE = 0, PKMN_RESULT_WIN = 1, PKMN_RESULT_LOSE = 2, PKMN_RESULT_TIE = 3, PKMN_RESULT_ERROR = 4
} pkmn_result_kind;
pkmn_result_kind pkmn_result_type(pkmn_result result);
pkmn_choice_kind pkmn_result_p1(pkmn_result result);
pkmn_choice_kind pkmn_result_p2(pkmn_result result);
bool pkmn_error(pkmn_result result); /* result == 0b1100: the log buffer was too small */

/* pkmn.h:133-161 */
typedef double float64_t;
typedef struct { uint8_t bytes[8]; } pkmn_psrng;
void pkmn_psrng_init(pkmn_psrng *psrng, uint64_t seed);
uint32_t pkmn_psrng_next(pkmn_psrng *psrng);
typedef struct { uint8_t bytes[16]; } pkmn_rational;
void pkmn_rational_init(pkmn_rational *rational);
void pkmn_rational_reduce(pkmn_rational *rational);
float64_t pkmn_rational_numerator(const pkmn_rational *rational);
float64_t pkmn_rational_denominator(const pkmn_rational *rational);

/* pkmn.h:163-209 */
#define PKMN_GEN1_BATTLE_SIZE 384
#define PKMN_GEN1_BATTLE_OPTIONS_SIZE 128
#define PKMN_GEN1_CHANCE_ACTIONS_SIZE 16
#define PKMN_GEN1_CHANCE_DURATIONS_SIZE 8
#define PKMN_GEN1_CALC_OVERRIDES_SIZE 24 /* the engine reads the first 16 (bindings/c.zig:101-103) */
#define PKMN_GEN1_CALC_SUMMARIES_SIZE 8
extern const size_t PKMN_GEN1_MAX_CHOICES;  /* 9 */
extern const size_t PKMN_GEN1_CHOICES_SIZE; /* 16 */
extern const size_t PKMN_GEN1_MAX_LOGS;     /* 180 */
extern const size_t PKMN_GEN1_LOGS_SIZE;    /* 256 */
typedef struct { uint8_t bytes[PKMN_GEN1_BATTLE_SIZE]; } pkmn_gen1_battle;
typedef struct { uint8_t bytes[PKMN_GEN1_BATTLE_OPTIONS_SIZE]; } pkmn_gen1_battle_options;
typedef struct { uint8_t bytes[PKMN_GEN1_CHANCE_ACTIONS_SIZE]; } pkmn_gen1_chance_actions;
typedef struct { uint8_t bytes[PKMN_GEN1_CHANCE_DURATIONS_SIZE]; } pkmn_gen1_chance_durations;
typedef struct { uint8_t bytes[PKMN_GEN1_CALC_OVERRIDES_SIZE]; } pkmn_gen1_calc_overrides;
typedef struct { uint8_t bytes[PKMN_GEN1_CALC_SUMMARIES_SIZE]; } pkmn_gen1_calc_summaries;
typedef struct { uint8_t *buf; size_t len; } pkmn_gen1_log_options;
typedef struct {
  pkmn_rational probability;
  pkmn_gen1_chance_actions actions;
  pkmn_gen1_chance_durations durations;
} pkmn_gen1_chance_options;
typedef struct { pkmn_gen1_calc_overrides overrides; } pkmn_gen1_calc_options;

/* pkmn.h:225-229; bindings/c.zig:116-148. A non-NULL facet is installed; NULL resets it: the log writer goes back to the
 * start of the buffer it already has (`writer.reset()`), chance keeps its durations and clears actions and probability
 * (chance.zig:367-373), calc clears the overrides. */
void pkmn_gen1_battle_options_set(pkmn_gen1_battle_options *options, const pkmn_gen1_log_options *log,
                                  const pkmn_gen1_chance_options *chance, const pkmn_gen1_calc_options *calc);
/* pkmn.h:231-256 */
pkmn_rational *pkmn_gen1_battle_options_chance_probability(const pkmn_gen1_battle_options *options);
pkmn_gen1_chance_actions *pkmn_gen1_battle_options_chance_actions(const pkmn_gen1_battle_options *options);
pkmn_gen1_chance_durations *pkmn_gen1_battle_options_chance_durations(const pkmn_gen1_battle_options *options);
pkmn_gen1_calc_summaries *pkmn_gen1_battle_options_calc_summaries(const pkmn_gen1_battle_options *options);
/* pkmn.h:264-268; bindings/c.zig:174-184. options may be NULL (no log, no chance tracking, no overrides). With a log
 * buffer installed the update's protocol messages are appended at the writer's position; if they do not fit the result is
 * 0b1100 (pkmn_error). */
pkmn_result pkmn_gen1_battle_update(pkmn_gen1_battle *battle, pkmn_choice c1, pkmn_choice c2, pkmn_gen1_battle_options *options);
/* pkmn.h:283-288; bindings/c.zig:186-198. Writes at most len choices, returns how many it wrote. */
uint8_t pkmn_gen1_battle_choices(const pkmn_gen1_battle *battle, pkmn_player player, pkmn_choice_kind request, pkmn_choice out[],
                                 size_t len);

#ifdef __cplusplus
}
#endif
#endif
