// Private host-side context shared by the translation units of liboakb200.so.
#pragma once

#include <cuda_runtime.h>

#include <cstdint>
#include <string>

#include "../../include/oakb200.h"

struct oakb200_net;     // device copy of a loaded network (value_net.cu)
struct oakb200_frames;  // training-frame recorder (frame_log.cu)
struct oakb200_rows;    // training-row builder scratch (train_rows.cu)

struct oakb200_ctx {
  int device = -1;
  int sm_count = 0;
  size_t n_max = 0, n = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  uint8_t *battles = nullptr, *durations = nullptr, *results = nullptr;
  // scratch for the host-pointer entry points
  uint8_t *c1 = nullptr, *c2 = nullptr, *overrides = nullptr, *actions = nullptr, *choices = nullptr, *counts = nullptr;
  uint8_t* scalar_pin = nullptr;  // pinned staging of the scalar pkmn_* symbols (1 KiB)
  uint8_t* logs = nullptr;  // scratch of oakb200_step_log: [n] u32 lengths, then [n][log_stride] bytes
  size_t log_bytes = 0;
  uint64_t* seeds = nullptr;
  uint8_t* score2 = nullptr;
  uint16_t* turns = nullptr;
  uint32_t* steps = nullptr;
  uint8_t* teams = nullptr;
  uint32_t* ticket = nullptr;
  int rollout_cta_cap = 0;  // > 0: at most this many rollout CTAs per SM (search.cu leaves room for its level kernels)
  int ticket_slot = 0;  // which of the 64 ticket words a rollout launch uses (search.cu runs rollouts on several streams)
  unsigned long long* total_steps = nullptr;
  uint64_t launches = 0;
  oakb200_net* net = nullptr;
  oakb200_frames* frames = nullptr;
  oakb200_rows* rows = nullptr;
  int eval_mode = 0;  // 0 auto (tensor cores when the shape fits), 1 fp32 CUDA-core kernel, 2 tensor cores or fail
  float* eval_values = nullptr;
  float* eval_logits = nullptr;
  uint8_t* eval_counts = nullptr;
  // transfer/compute pipeline of oakb200_rollout_host
  bool pipe_ready = false;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t pipe_start = nullptr;
  uint32_t* pipe_marks = nullptr;  // pinned: records available after each chunk
  // transfer/compute pipeline of oakb200_eval_host
  bool eval_pipe_ready = false;
  cudaStream_t eval_h2d = nullptr, eval_d2h = nullptr;
  cudaEvent_t eval_ev[130] = {};  // [0] start, then {uploaded, evaluated} per chunk
  std::string error;
};

// pkmn_compat.cu -> oakb200.cu: a batch of one through pinned staging, every copy and the kernel queued on the context
// stream, ONE synchronisation. `log` non-null selects the LOG kernel (log_cap bytes at most, *log_len written).
int oakb200_scalar_update(oakb200_ctx* ctx, uint8_t* battle, uint8_t c1, uint8_t c2, uint8_t* durations, const uint8_t* overrides,
                          uint8_t* actions, uint8_t* result, uint8_t* log, uint32_t log_cap, uint32_t* log_len);
int oakb200_scalar_choices(oakb200_ctx* ctx, const uint8_t* battle, uint8_t result, uint8_t* choices18, uint8_t* counts2);

// search.cu -> oakb200.cu
cudaError_t oakb200_launch_search_spawn(oakb200_ctx* ctx, const uint8_t* root_battles, const uint8_t* root_durations,
                                        const uint8_t* root_results, const uint64_t* rng_values, uint32_t per_root,
                                        uint8_t* battles, uint8_t* durations, uint8_t* results, uint32_t n);
cudaError_t oakb200_launch_search_step(oakb200_ctx* ctx, uint8_t* battles, uint8_t* durations, uint8_t* results, const uint8_t* idx1,
                                       const uint8_t* idx2, const uint8_t* rolls, uint8_t* actions, uint8_t* counts, uint32_t n);

// train_rows.cu -> oakb200.cu
cudaError_t oakb200_launch_replay(oakb200_ctx* ctx, const uint8_t* stream, const uint64_t* game_off, const uint32_t* row_base,
                                  const int32_t* only_update, uint32_t n_games, uint32_t* words, uint8_t* rows);
cudaError_t oakb200_launch_replay_encode(oakb200_ctx* ctx, const uint8_t* stream, const uint64_t* game_off, const uint32_t* row_base,
                                         const int32_t* only_update, uint32_t n_games, uint32_t* words, uint8_t* rows,
                                         const oakb200_training_rows* out);
cudaError_t oakb200_launch_replay_split(oakb200_ctx* ctx, const uint8_t* stream, const uint64_t* game_off, const uint32_t* row_base,
                                        const int32_t* only_update, uint32_t n_games, uint32_t n_rows, uint32_t* control, uint32_t* queue,
                                        uint8_t* rows, const oakb200_training_rows* out);
static const int OAKB200_ROW_STATE_BYTES = 432;  // = ROW_STATE_BYTES of battle_kernels.cuh

namespace {

inline int fail(oakb200_ctx* ctx, int code, const char* what, cudaError_t e = cudaSuccess) {
  if (ctx) {
    ctx->error = what;
    if (e != cudaSuccess) {
      ctx->error += ": ";
      ctx->error += cudaGetErrorString(e);
    }
  }
  return code;
}

#define CU(ctx, call)                                              \
  do {                                                             \
    cudaError_t e__ = (call);                                      \
    if (e__ != cudaSuccess) return fail(ctx, OAKB200_ECUDA, #call, e__); \
  } while (0)

}  // namespace
