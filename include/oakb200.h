/* oakb200 — B200-native batched Generation I battle engine: C ABI.
 *
 * Drop-in boundary for the data-parallel hot path of lab-oak/oak. Host buffers keep the reference's
 * own array-of-records formats so existing Oak buffers pass through untouched:
 *   battle    384 bytes  pkmn_gen1_battle            (reference engine/src/include/pkmn.h:150-160,
 *                                                     layout engine/src/lib/gen1/data.zig:54-68)
 *   durations   8 bytes  pkmn_gen1_chance_durations  (engine/src/lib/gen1/chance.zig:255-318)
 *   actions    16 bytes  pkmn_gen1_chance_actions    (chance.zig:30-38,160-206)
 *   overrides  16 bytes  pkmn_gen1_calc_options.overrides (bindings/c.zig:101-103)
 *   choice      1 byte   pkmn_choice  (type | data << 2, common/data.zig:56-72)
 *   result      1 byte   pkmn_result  (type | p1 << 4 | p2 << 6, common/data.zig:84-112)
 * Engine semantics are libpkmn's as Oak builds it (reference dev/libpkmn:9: -Dshowdown, miss=false,
 * advance=false, ebc=false, key=true, -Dchance, -Dcalc).
 *
 * Every entry point returns 0 on success or a negative OAKB200_E* code; oakb200_last_error() gives
 * the text. Nothing throws across the ABI. One context per GPU; calls on one context are serialised
 * by the caller (the reference's own contract is one search object per thread, cpp/include/search/
 * mcts.h:135-139). There is no CPU fallback: without a CUDA device oakb200_create fails.
 *
 * Entry points ending in _device take DEVICE pointers and enqueue on the context stream without
 * synchronising; the others take HOST pointers, copy, run and synchronise.
 */
#ifndef OAKB200_H
#define OAKB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OAKB200_BATTLE_SIZE 384    /* PKMN_GEN1_BATTLE_SIZE, pkmn.h:150 */
#define OAKB200_DURATIONS_SIZE 8   /* PKMN_GEN1_CHANCE_DURATIONS_SIZE */
#define OAKB200_ACTIONS_SIZE 16    /* PKMN_GEN1_CHANCE_ACTIONS_SIZE */
#define OAKB200_MAX_CHOICES 9      /* PKMN_GEN1_MAX_CHOICES, pkmn.h:163 */

#define OAKB200_OK 0
#define OAKB200_ENODEVICE (-1)
#define OAKB200_ECUDA (-2)
#define OAKB200_EINVAL (-3)
#define OAKB200_ECAPACITY (-4)

typedef struct oakb200_ctx oakb200_ctx;

/* rollout flags */
#define OAKB200_ROLLOUT_WRITE_STATE 1u /* write final battle/durations/result back (parity checks) */
#define OAKB200_ROLLOUT_TRACK_CHANCE 2u /* maintain chance durations during playouts, as the reference's
                                           options-enabled update does (mcts.h:462-479); the outputs
                                           (score, turns, steps) do not depend on it */

int oakb200_create(int device, size_t n_max, oakb200_ctx** out);
void oakb200_destroy(oakb200_ctx* ctx);
const char* oakb200_last_error(const oakb200_ctx* ctx);
int oakb200_device(const oakb200_ctx* ctx);
/* Use an existing cudaStream_t (e.g. the caller's framework stream) instead of the context's own. */
int oakb200_set_stream(oakb200_ctx* ctx, void* cuda_stream);
int oakb200_synchronize(oakb200_ctx* ctx);
/* Number of kernels this library has launched on the context since creation. */
uint64_t oakb200_launch_count(const oakb200_ctx* ctx);

/* ---- resident batch -------------------------------------------------------------------------
 * upload: make `n` battles resident (durations NULL = zeros, results NULL = 0x50 "both move").
 * Replaces: building a std::vector of MCTS::Input / Frames rows (cpp/include/search/mcts.h:62-66,
 * cpp/include/py/battle/frames.h:22-27). */
int oakb200_upload(oakb200_ctx* ctx, size_t n, const uint8_t* battles, const uint8_t* durations, const uint8_t* results);
int oakb200_download(oakb200_ctx* ctx, size_t n, uint8_t* battles, uint8_t* durations, uint8_t* results);
/* Device pointers of the resident arrays (battle[n][384], durations[n][8], result[n]). */
int oakb200_resident(oakb200_ctx* ctx, void** battles, void** durations, void** results, size_t* n);

/* ---- pkmn_gen1_battle_choices for a batch (pkmn.h:283-288; mechanics.zig:3139-3251) ------------
 * For every resident battle and both players, given the request kinds held in the resident result
 * byte: out[n][2][9] choices (switches first, then moves) and counts[n][2]. HOST outputs. */
int oakb200_choices(oakb200_ctx* ctx, uint8_t* out, uint8_t* counts);
int oakb200_choices_device(oakb200_ctx* ctx, size_t n, const void* battles, const void* results, void* out, void* counts);

/* ---- pkmn_gen1_battle_options_set + pkmn_gen1_battle_update for a batch (pkmn.h:225-229,264-268) -
 * c1/c2[n]: the players' choices. overrides[n][16] or NULL: calc overrides (Oak writes bytes 0 and 8,
 * mcts.h:569-590); NULL = options_set(calc = null). Resident battle/durations/result are updated in
 * place; results_out[n] and actions_out[n][16] (each nullable) receive copies. HOST pointers. */
int oakb200_step(oakb200_ctx* ctx, const uint8_t* c1, const uint8_t* c2, const uint8_t* overrides,
                 uint8_t* results_out, uint8_t* actions_out);
int oakb200_step_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* c1,
                        const void* c2, const void* overrides, void* actions_out);
/* The same update with libpkmn's binary protocol log, i.e. what a -Dlog build of libpkmn writes through
 * pkmn_gen1_log_options (pkmn.h:184-193; messages engine/src/lib/common/protocol.zig:313-824, docs/PROTOCOL.md; consumer
 * in Oak: cpp/include/util/debug-log.h:13-68, cpp/src/log.cc). logs_out[n][log_stride] receives battle i's messages for
 * this update at row i, log_lens_out[n] (u32) their byte count. log_stride >= OAKB200_MAX_LOGS always suffices
 * (PKMN_GEN1_MAX_LOGS, engine/src/lib/gen1/data.zig:22). A battle whose log does not fit keeps the message parts that did,
 * gets the result byte 0x0C (pkmn_error(), bindings/c.zig:6,181) and is not stepped again; unlike the reference, which
 * abandons the update half-way at that point, its state is that of the completed update. */
#define OAKB200_MAX_LOGS 180
int oakb200_step_log(oakb200_ctx* ctx, const uint8_t* c1, const uint8_t* c2, const uint8_t* overrides, uint8_t* results_out,
                     uint8_t* actions_out, uint8_t* logs_out, uint32_t log_stride, uint32_t* log_lens_out);
int oakb200_step_log_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* c1,
                            const void* c2, const void* overrides, void* actions_out, void* logs, uint32_t log_stride,
                            void* log_lens);
/* One random-playout step (body of the loop at mcts.h:467-479): choices for both players, then
 * c1 = choices1[pick % m], c2 = choices2[(pick >> 32) % n], options_set(null), update.
 * pick for battle i = splitmix64(seeds[i] ^ (0xD1B54A32D192ED03 * step_index)). */
int oakb200_step_random_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results,
                               const void* seeds, uint32_t step_index, void* actions_out, void* chosen_out);

/* ---- random playouts to a terminal result (MCTS::init_stats_and_rollout, mcts.h:448-496) ------
 * For every resident battle i play picks splitmix64(seeds[i] ^ (K * step)), step = 0, 1, ... until
 * the result type is non-zero or max_steps updates were made. score2_out[n]: 2 win / 1 tie / 0 lose
 * for player 1 (PKMN::score2, cpp/include/libpkmn/pkmn.h:192-208; 255 if max_steps hit);
 * turns_out[n] (u16): battle turn counter at the end; steps_out[n] (u32): updates made. Any of the
 * three may be NULL. The resident state is left untouched unless OAKB200_ROLLOUT_WRITE_STATE. */
int oakb200_rollout(oakb200_ctx* ctx, const uint64_t* seeds, uint32_t max_steps, uint32_t flags, uint8_t* score2_out,
                    uint16_t* turns_out, uint32_t* steps_out);
int oakb200_rollout_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* seeds,
                           uint32_t max_steps, uint32_t flags, void* score2_out, void* turns_out, void* steps_out);
/* oakb200_upload + oakb200_rollout in one call with the upload hidden behind the playouts (chunk copies on a copy stream,
 * one persistent kernel that consumes records as they land): HOST pointers, PINNED memory required for the overlap. durations / results / any output may be NULL.
 * The batch stays resident afterwards. No state write-back and no chance tracking (flags = 0). */
int oakb200_rollout_host(oakb200_ctx* ctx, size_t n, const uint8_t* battles, const uint8_t* durations, const uint8_t* results,
                         const uint64_t* seeds, uint32_t max_steps, uint8_t* score2_out, uint16_t* turns_out, uint32_t* steps_out);
/* Sum of updates performed by all rollout launches on this context (device counter, synchronises). */
int oakb200_total_steps(oakb200_ctx* ctx, uint64_t* out);

/* ---- MCTS::randomize_hidden_variables for a batch (cpp/include/search/durations.h:25-97) ------- */
int oakb200_randomize_hidden_device(oakb200_ctx* ctx, size_t n, void* battles, const void* durations);

/* ---- PKMN::battle + update(0,0) for a batch (pkmn.h:50-57, init.h:90-154, generate.cc:238-240) --
 * teams[n][2][6][5] = {species, move x4}; seeds[n]; builds level-100 max-stat battles on the device
 * into the resident arrays, sends out the leads, durations = 0. HOST pointers. */
int oakb200_build_battles(oakb200_ctx* ctx, size_t n, const uint8_t* teams, const uint64_t* seeds);

/* ---- leaf evaluation: battle encoding -> embedding nets -> main MLP ------------------------------
 * Replaces NN::Battle::NetworkImpl::{read_parameters, value_policy_inference} (reference
 * cpp/include/nn/battle/network.h:52-70,102-123; encoder cpp/include/encode/battle/battle.h; parameter
 * file = 8-byte header (byte 0: activation - 1; cpp/src/search.cc:127-147) + 12 Affine blocks
 * {u32 in, u32 out, f32 bias[out], f32 W[out][in]} (cpp/include/nn/affine.h:34-69), must end exactly at
 * the last block. Evaluation is cache-free (the reference's per-search embedding memo is an
 * optimisation of the same function, cpp/include/nn/battle/cache.h).
 * net_load: parse + upload a parameter file image (HOST pointer). Limits: hidden dims <= 128,
 * embedding dims <= 96, main hidden <= 64, head hidden <= 64; OAKB200_EINVAL otherwise. */
int oakb200_net_load(oakb200_ctx* ctx, const uint8_t* file_bytes, size_t len);
/* The same parameter file loaded as the reference's "discrete" agent (cpp/src/search.cc:105-123): float embedding nets,
 * embeddings and hp quantised to u8 (static_cast<uint8_t>(127 * f), nn/battle/cache.h:106,234, network.h:147,165), int8
 * main net (nn/battle/quantized/{affine,clipped_relu,main-net}.h: int8 weights = w * 64, int32 biases = b * 64 * 127,
 * ClippedReLU clamp(x >> 6, 0, 127), outputs / (127 * 64)). The integer part is exact. OAKB200_EINVAL unless the file's
 * header says clamped activations and every main-net weight is inside (-2, 2), as the reference requires. */
int oakb200_net_load_discrete(oakb200_ctx* ctx, const uint8_t* file_bytes, size_t len);
/* Which kernel evaluates: 0 = auto (the tcgen05 tensor-core kernel whenever the layer shapes fit: hidden <= 128,
 * pokemon/active embedding <= 63/95, main hidden <= 64, value hidden <= 32, policy hidden <= 64 and a multiple
 * of 4 -- the reference's defaults do), 1 = the fp32 CUDA-core kernel, 2 = tensor cores or OAKB200_EINVAL.
 * Both compute the same function; the tensor-core kernel uses split-bf16 operands (~1e-5 relative). */
int oakb200_eval_set_mode(oakb200_ctx* ctx, int mode);
/* For every battle: legal choices are derived from its result byte (as choices does), then
 * value[n] (after the sigmoid, network.h:118), logits[n][2][9] for those choices in choices order
 * (unused entries 0), counts[n][2]. */
int oakb200_eval(oakb200_ctx* ctx, float* value_out, float* logits_out, uint8_t* counts_out);
int oakb200_eval_device(oakb200_ctx* ctx, size_t n, const void* battles, const void* durations, const void* results,
                        void* value_out, void* logits_out, void* counts_out);
/* Evaluation of a HOST batch with the PCIe transfers hidden behind the kernels: chunked upload, evaluation and download
 * on three streams. Same outputs as oakb200_upload + oakb200_eval; durations / results may be NULL as in oakb200_upload.
 * Returns when the outputs are in the host arrays. */
int oakb200_eval_host(oakb200_ctx* ctx, size_t n, const uint8_t* battles, const uint8_t* durations, const uint8_t* results,
                      float* value_out, float* logits_out, uint8_t* counts_out);

/* Rebuild, in place, every battle of the batch whose result byte says it has ended (PKMN::battle + update(0,0) as
 * oakb200_build_battles; teams[n][2][6][5] and seeds[n] are DEVICE arrays, only the entries of ended slots are read).
 * Replaces the "next game" step of the generate loop (cpp/src/generate.cc:218-240). */
int oakb200_respawn_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* teams,
                           const void* seeds);

/* ---- policy post-processing + sampling (cpp/include/util/policy.h:76-107, util/random.h:41-50, search/util/softmax.h) ----
 * For every battle and both players: probs = softmax(logits over the legal choices) (fp32); then, in double,
 * p^temperature renormalised, entries below min_prob zeroed, renormalised; the choice is drawn as sample_pdf does with
 * u = (splitmix64(seeds[i] ^ 0xD1B54A32D192ED03 * step_index ^ player) >> 11) * 2^-53.
 * choices[n][2][9], counts[n][2] as oakb200_choices_device writes them; logits[n][2][9] as oakb200_eval_device writes them.
 * Out: probs_out[n][2][9] f32 (the softmax, unused entries 0), c1_out[n], c2_out[n]. DEVICE pointers. */
int oakb200_policy_sample_device(oakb200_ctx* ctx, size_t n, const void* choices, const void* counts, const void* logits,
                                 const void* seeds, uint32_t step_index, double temperature, double min_prob, void* probs_out,
                                 void* c1_out, void* c2_out);

/* ---- training-frame recorder / writer -----------------------------------------------------------------------------
 * Replaces Train::Battle::CompressedFrames (cpp/include/train/battle/compressed-frame.h:36-199) and the per-thread
 * frame buffer of generate (cpp/src/generate.cc:161-188,344-359; cpp/include/util/battle-frame-buffer.h) for `games`
 * concurrent games, slot i recording the batch's battle i. The stream oakb200_frames_drain returns is the concatenation
 * of the finished games' records, byte-identical to what the reference appends to an N.battle.data file:
 *   u32 n_bytes | u16 n_updates | battle[384] | result | { u8 (m-1)|(n-1)<<4, c1, c2, u32 iterations,
 *   u16 empirical_value, u16 nash_value, u16 p1_empirical[m], p1_nash[m], p2_empirical[n], p2_nash[n] }*
 * with probabilities and values stored as (uint16)(x * 65535) (compress_probs, compressed-frame.h:12-24).
 * A game longer than max_updates is dropped when it ends (the reference discards games over --max-battle-length,
 * generate.cc:268-271,333-336). */
int oakb200_frames_create(oakb200_ctx* ctx, size_t games, uint32_t max_updates);
/* Start recording in every idle slot whose battle is not over: snapshots battle[384] (DEVICE pointers). */
int oakb200_frames_begin_device(oakb200_ctx* ctx, size_t n, const void* battles, const void* results);
/* Append one update to every recording slot whose battle (result byte BEFORE the update) is still running.
 * c1/c2[n] the choices made; counts[n][2] the legal-choice counts (m, n); iterations[n] u32 or NULL (0);
 * empirical_value/nash_value[n] f32; empirical/nash[n][2][9] f32 probabilities in choices order. DEVICE pointers. */
int oakb200_frames_record_device(oakb200_ctx* ctx, size_t n, const void* results, const void* c1, const void* c2, const void* counts,
                                 const void* iterations, const void* empirical_value, const void* nash_value,
                                 const void* empirical, const void* nash);
/* Close every recording slot whose battle has ended (result byte AFTER the update): header + result. */
int oakb200_frames_finish_device(oakb200_ctx* ctx, size_t n, const void* results);
/* One turn of network-policy self-play for n games in one call (the battle loop of cpp/src/generate.cc:271-333 with a
 * budget-0 agent): choices, evaluation, RuntimePolicy temperature / floor / sample (counter-based stream from seeds and
 * step_index), frame record (empirical = nash = the network's policy and value, iterations 0), update, finish. All
 * arrays DEVICE; choices[n][2][9], counts[n][2], value[n], logits[n][2][9], eval_counts[n][2], probs[n][2][9], c1[n], c2[n],
 * before[n] are outputs / scratch the caller owns. Needs oakb200_net_load and oakb200_frames_create + _begin_device. */
int oakb200_selfplay_step_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* seeds,
                                 uint32_t step_index, double temperature, double min_prob, void* choices, void* counts, void* value,
                                 void* logits, void* eval_counts, void* probs, void* c1, void* c2, void* before);
/* Finished games waiting to be drained / games dropped for length so far (synchronises). */
int oakb200_frames_pending(oakb200_ctx* ctx, uint32_t* finished, uint32_t* dropped);
/* Pack the finished games in slot order into out[cap] (HOST or DEVICE memory: a device buffer keeps the stream on the GPU
 * for a gather over NCCL), free their slots. *bytes = stream length (also set when the call fails with
 * OAKB200_ECAPACITY, which consumes nothing), *games = how many games it holds. */
int oakb200_frames_drain(oakb200_ctx* ctx, uint8_t* out, size_t cap, size_t* bytes, uint32_t* games);

/* ---- batched-leaf tree search: host trees + bandits, GPU battle operations -----------------------------------------
 * Replaces MCTS::Search::run with MonteCarlo evaluation, the joint UCB bandit and a Node heap (reference
 * cpp/include/search/mcts.h:154-496,569-604; cpp/include/search/bandit/ucb.h; cpp/include/search/joint.h) for many
 * root positions at once: `roots` x `particles` descents run in lockstep, one kernel launch per tree level (legal
 * choices by index, clamped-roll update, chance observation), one for the hidden-variable draw and one for all leaf
 * playouts of a wave. particles = 1 reproduces the reference's sequential search selection for selection given the
 * same random streams (root seed s, n-th iteration: battle rng = splitmix64(s ^ K(2n+1)), playout seed =
 * splitmix64(s ^ K(2n+2)), K = 0xD1B54A32D192ED03); particles > 1 batches descents of one tree with virtual visits.
 * The Nash solve of the root matrix (LRSNash, mcts.h:645-649) stays with the caller: outputs are the matrices. */
typedef struct oakb200_search oakb200_search;
int oakb200_search_create(oakb200_ctx* ctx, size_t roots, uint32_t particles, oakb200_search** out);
void oakb200_search_destroy(oakb200_search* search);
/* MCTS::Input rows (mcts.h:62-66) + one stream seed per root; resets the trees. HOST pointers. */
int oakb200_search_set_roots(oakb200_search* search, size_t n, const uint8_t* battles, const uint8_t* durations,
                             const uint8_t* results, const uint64_t* seeds);
/* Exactly `iterations` more iterations per root (the last wave spawns only what is left). ucb_c: UCB::Bandit::Params::c;
 * root_rolls / other_rolls: SearchOptions (1, 2, 3, 20, or 39 = unclamped; the reference default is 3 / 1);
 * host_threads <= 0: all cores. */
int oakb200_search_run(oakb200_search* search, uint32_t iterations, float ucb_c, uint32_t root_rolls, uint32_t other_rolls,
                       int host_threads);
/* MCTS::Output per root (mcts.h:68-89): k[n][2], choices[n][2][9], visit_matrix[n][9][9] u32, value_matrix[n][9][9] f64,
 * iterations[n], empirical_value[n] f64, empirical[n][2][9] f64 (process_output, mcts.h:623-661), nodes[n]. Any may be NULL. */
int oakb200_search_output(oakb200_search* search, uint8_t* k, uint8_t* choices, uint32_t* visit_matrix, double* value_matrix,
                          uint64_t* iterations, double* empirical_value, double* empirical, uint64_t* nodes);
/* The bandit of every tree node (cpp/include/search/bandit/*.h; strings of cpp/src/search.cc:236-300): kind 0 "ucb-<c>"
 * (the default, with the c passed to oakb200_search_run), 1 "ucb1-<c>", 2 "pucb-<c>", 3 "exp3-<gamma>-<alpha>", 4
 * "pexp3-<gamma>-<alpha>"; f1 = c or gamma, f2 = alpha (ignored by the UCB family; the reference's default is 0.05). pucb and
 * pexp3 are primed with the network's policy logits at the root and at every new leaf (mcts.h:190-209,409-421): they need a
 * loaded network and leaf mode 1. Resets the trees of roots already set. Not built: root matrix-UCB (mcts.h:498-566). */
int oakb200_search_set_bandit(oakb200_search* search, int kind, float f1, float f2);
/* Leaf evaluation of the search: 0 = MonteCarlo playout (default), 1 = the loaded network's value (value_inference at a new
 * leaf, mcts.h:409-424, with the chance durations tracked down the path; a battle that ended on the way scores 1 / 0 /
 * 0.5). Mode 1 needs oakb200_net_load[_discrete] first; all leaves of a wave are one batched evaluation. */
int oakb200_search_set_leaf(oakb200_search* search, int mode);
int oakb200_search_stats(oakb200_search* search, uint64_t* waves, uint64_t* levels);
/* Zero-sum matrix game, the job of LRSNash::solve_fast (mcts.h:643-649; pyoak solve_matrix, cpp/src/pyoak.cc:394-425):
 * the row player receives payoff[i * cols + j] / denom, the column player 1 - that; rows, cols <= 9. -> both players'
 * equilibrium strategies and the game value. An LP solved by simplex: the value is the game's, the strategies one of its
 * equilibria (LRSNash may return another vertex of the same solution set). Pure host code, no GPU needed. */
int oakb200_solve_matrix(int rows, int cols, const int* payoff, int denom, float* p1, float* p2, float* value);
/* process_output's Nash half for every root (mcts.h:628-656: average values discretised to 1/256, then solved):
 * nash[n][2][9] f32 (zero past k), nash_value[n] f32. Either may be NULL. */
int oakb200_search_nash(oakb200_search* search, float* nash, float* nash_value);
/* Where the wall time of the last oakb200_search_run went (no counterpart in the reference; a measurement aid):
 * out4 = {total ms, ms the calling thread spent in host phases (tree work of the pool + copies and launches),
 * ms it spent with every root group's work on the device, waves}. */
int oakb200_search_timing(oakb200_search* search, double* out4);

/* ---- training rows: stored games -> the dense tensors the learner trains on ------------------------------------------
 * Replaces pyoak's EncodedBattleFrames.from_bytes / sample (reference cpp/include/py/battle/encoded-frames.h:24-133,
 * cpp/include/py/battle/target.h:17-70, cpp/src/pyoak.cc:111-260,655-673): replay every stored game of a .battle.data
 * stream (cpp/include/train/battle/compressed-frame.h:173-224) from its first battle with its stored choices and,
 * before each update, write one row. Row-major arrays of n_rows rows, in DEVICE memory (they feed the learner there):
 *   pokemon[n][2][6][198] f32, active[n][2][1][229] f32, hp[n][2][6][1] f32   dense encodings (encode/battle/battle.h),
 *                                                                              zero for fainted / empty slots
 *   choice_indices[n][2][9] i64   policy index of each legal choice (encode/battle/policy.h:29-58), 315 past the k-th
 *   k[n][2] u8, choice[n][2] u8, iterations[n] u32, empirical_policies / nash_policies[n][2][9] f32 (u16 / 65535, zero
 *   past k), empirical_value[n], nash_value[n] f32, score[n] f32 (1 win, 0 lose, 0.5 tie of the game's stored result)
 * Any pointer may be NULL (that array is skipped). Rows of game g start at the running sum of the games' n_updates. */
typedef struct oakb200_training_rows {
  float* pokemon;
  float* active;
  float* hp;
  int64_t* choice_indices;
  uint8_t* k;
  uint8_t* choice;
  uint32_t* iterations;
  float* empirical_policies;
  float* nash_policies;
  float* empirical_value;
  float* nash_value;
  float* score;
  /* the uncompressed frame of the same row (pyoak BattleFrames, py/battle/frames.h:17-51): the battle, durations and
   * result the row was built from; legal_choices[n][2][9] u8 = the legal choices of both players, zero past their count
   * (the reference allocates `choices` and leaves it zero, frames.h:47-49 — this array is what it was meant to hold) */
  uint8_t* battle;
  uint8_t* durations;
  uint8_t* result;
  uint8_t* legal_choices;
} oakb200_training_rows;

/* Host walk over the game headers of a stream (pyoak read_battle_data, pyoak.cc:43-71): byte offset and first row of
 * every game. game_off / row_base may be NULL to count only; otherwise they hold `capacity` entries. OAKB200_EINVAL for a
 * stream that does not split into whole games, OAKB200_ECAPACITY if there are more games than capacity. No GPU needed. */
int oakb200_rows_index_host(const uint8_t* stream, uint64_t n_bytes, uint64_t* game_off, uint32_t* row_base, uint64_t capacity,
                            uint64_t* n_games, uint64_t* n_rows);
/* Stream and index already in device memory (n_bytes < 4 GiB). only_update NULL: every stored update becomes a row
 * (from_bytes). only_update[g] >= 0: game g contributes the single row g, built before its update only_update[g]
 * (`sample`: the caller draws the game and the update); row_base is then ignored. Asynchronous on the context's stream. */
int oakb200_rows_build_device(oakb200_ctx* ctx, const uint8_t* stream, uint64_t n_bytes, const uint64_t* game_off,
                              const uint32_t* row_base, const int32_t* only_update, uint64_t n_games, uint64_t n_rows,
                              const oakb200_training_rows* out);
/* Waits for the build and reports the lowest index of a game whose replay did not reproduce its stored final result or
 * byte count (-1 if none; the reference asserts this, encoded-frames.h:132). Its rows are written but not trustworthy. */
int oakb200_rows_status(oakb200_ctx* ctx, int64_t* first_bad_game);
/* The whole path from a HOST stream: index, upload, build, status. out arrays (device) hold capacity_rows rows.
 * only_update is a host array or NULL. */
int oakb200_rows_build_host(oakb200_ctx* ctx, const uint8_t* stream, uint64_t n_bytes, const int32_t* only_update,
                            const oakb200_training_rows* out, uint64_t capacity_rows, uint64_t* n_games, uint64_t* n_rows,
                            int64_t* first_bad_game);


/* oakb200_rows_build_device without any allocation. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). The staged row states and the control words live in a file-scope `static __device__` array of the library
 * (256 B + OAKB200_ROWS_STATIC_MAX x 432 B = 6.75 MiB, held by every GPU that loads the library until the process exits);
 * the call itself makes no cudaMalloc / cudaFree / allocating library call: it enqueues two 4-byte memsets, the replay
 * kernel, the encode kernel and one 4-byte device-to-device copy on the context stream, and does not synchronise.
 * Same arguments, same rows, bit for bit, as oakb200_rows_build_device; all pointers are DEVICE pointers the caller owns.
 * n_rows > OAKB200_ROWS_STATIC_MAX returns OAKB200_ECAPACITY (oakb200_last_error says so) and enqueues nothing.
 * first_bad_game_out (device u32, nullable) receives what oakb200_rows_status reports for the other entry points: the
 * lowest index of a game whose replay missed its stored result or byte count, 0xFFFFFFFF if none.
 * The array is shared by every call on that GPU, from any context or thread, and keeps the previous call's contents (this
 * call resets what it needs): calls that can overlap on the GPU — different streams, different contexts — race on it;
 * issue them on one stream or order them with events. The first call on each device looks the array's address up
 * (cudaGetSymbolAddress); make that one call outside graph capture. */
#define OAKB200_ROWS_STATIC_MAX 16384
int oakb200_rows_build_static(oakb200_ctx* ctx, const uint8_t* stream, uint64_t n_bytes, const uint64_t* game_off,
                              const uint32_t* row_base, const int32_t* only_update, uint64_t n_games, uint64_t n_rows,
                              const oakb200_training_rows* out, uint32_t* first_bad_game_out);

#ifdef __cplusplus
}
#endif
#endif
