// C ABI of the B200 battle engine (include/oakb200.h). Host side: context, resident buffers,
// launches. No CPU compute path: every entry point either launches sm_100a kernels or fails.
#include "../../include/oakb200.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>

#include "battle_kernels.cuh"

using namespace oakb200;

namespace {
const int BLOCK = 128;                                  // battles per CTA
const size_t TILE_BYTES = (size_t)BATTLE_WORDS * BLOCK * 4;  // 48 KB shared-memory tile
}  // namespace

#include "context.h"

void oakb200_net_release(oakb200_ctx* ctx);     // value_net.cu
void oakb200_frames_release(oakb200_ctx* ctx);  // frame_log.cu
void oakb200_rows_release(oakb200_ctx* ctx);    // train_rows.cu

namespace {

template <class K>
cudaError_t allow_smem(K kernel) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TILE_BYTES);
}

unsigned grid_for(size_t n) { return (unsigned)((n + BLOCK - 1) / BLOCK); }

template <bool RANDOM, bool CALC>
cudaError_t launch_step(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* c1,
                        const void* c2, const void* seeds, uint32_t step_index, const void* overrides, void* actions,
                        void* chosen) {
  auto k = step_kernel<BLOCK, RANDOM, CALC>;
  cudaError_t e = allow_smem(k);
  if (e != cudaSuccess) return e;
  k<<<grid_for(n), BLOCK, TILE_BYTES, ctx->stream>>>((uint8_t*)battles, (uint8_t*)durations, (uint8_t*)results,
                                                     (const uint8_t*)c1, (const uint8_t*)c2, (const uint64_t*)seeds,
                                                     step_index, (const uint8_t*)overrides, (uint8_t*)actions,
                                                     (uint8_t*)chosen, (uint32_t)n);
  ctx->launches++;
  return cudaGetLastError();
}

template <bool CHANCE, bool WRITE>
cudaError_t launch_rollout(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* seeds,
                           uint32_t max_steps, void* score2, void* turns, void* steps, uint32_t* available = nullptr) {
  cudaStream_t stream = ctx->stream;
  uint32_t* ticket = ctx->ticket + ctx->ticket_slot;
  // OAKB200_ROLLOUT_GROUPED=1 selects the variant that regroups a CTA's battles by action class before every update.
  // Measured on B200 (profiles/r1h_rollout_grouping.txt): 16 % fewer warp instructions and 9.8 -> 12.8 active lanes, but
  // the CTA-wide barriers cost more issue slots than that saves (2.08 vs 2.35 G battle-steps/s), so it is not the default.
  static const bool grouped = [] {
    const char* g = getenv("OAKB200_ROLLOUT_GROUPED");
    return g && g[0] == '1';
  }();
  // OAKB200_ROLLOUT_SPLIT=0 selects the full 96-word on-chip layout (16 warps per SM) instead of the default 48-word one
  static const bool split = [] {
    const char* g = getenv("OAKB200_ROLLOUT_SPLIT");
    return !(g && g[0] == '0');
  }();
  auto k = grouped ? rollout_grouped_kernel<BLOCK, CHANCE, WRITE>
                   : (split ? rollout_kernel<BLOCK, CHANCE, WRITE, true> : rollout_kernel<BLOCK, CHANCE, WRITE, false>);
  size_t tile_bytes = grouped ? TILE_BYTES + (size_t)(5 * BLOCK + 64) * 4 : (split ? TILE_BYTES / 2 : TILE_BYTES);
  // tuning/diagnostic knob: pad the dynamic shared memory to lower the resident CTAs per SM
  if (const char* pad = getenv("OAKB200_ROLLOUT_SMEM_PAD")) tile_bytes += (size_t)atol(pad);
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_bytes);
  if (e != cudaSuccess) return e;
  int per_sm = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, BLOCK, tile_bytes);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  if (ctx->rollout_cta_cap > 0 && per_sm > ctx->rollout_cta_cap) per_sm = ctx->rollout_cta_cap;
  // persistent grid: exactly the resident CTAs of the chip (a multiple of the SM count), never more
  // CTAs than there is work
  unsigned grid = (unsigned)(ctx->sm_count * per_sm);
  const unsigned need = grid_for(n);
  if (grid > need) grid = need;
  e = cudaMemsetAsync(ticket, 0, sizeof(uint32_t), stream);
  if (e != cudaSuccess) return e;
  k<<<grid, BLOCK, tile_bytes, stream>>>((uint8_t*)battles, (uint8_t*)durations, (uint8_t*)results,
                                              (const uint64_t*)seeds, max_steps, (uint32_t)n, ticket,
                                              (uint8_t*)score2, (uint16_t*)turns, (uint32_t*)steps, ctx->total_steps, available);
  ctx->launches++;
  return cudaGetLastError();
}

}  // namespace

// kernel launches used by the tree-search driver (search.cu); battle_kernels.cuh is compiled in this unit only
cudaError_t oakb200_launch_search_spawn(oakb200_ctx* ctx, const uint8_t* root_battles, const uint8_t* root_durations,
                                        const uint8_t* root_results, const uint64_t* rng_values, uint32_t per_root,
                                        uint8_t* battles, uint8_t* durations, uint8_t* results, uint32_t n) {
  search_spawn_kernel<<<(unsigned)(((size_t)n * 32 + 255) / 256), 256, 0, ctx->stream>>>(root_battles, root_durations, root_results,
                                                                                        rng_values, per_root, battles, durations,
                                                                                        results, n);
  ctx->launches++;
  return cudaGetLastError();
}
cudaError_t oakb200_launch_search_step(oakb200_ctx* ctx, uint8_t* battles, uint8_t* durations, uint8_t* results, const uint8_t* idx1,
                                       const uint8_t* idx2, const uint8_t* rolls, uint8_t* actions, uint8_t* counts, uint32_t n) {
  auto k = search_step_kernel<BLOCK>;
  cudaError_t e = allow_smem(k);
  if (e != cudaSuccess) return e;
  k<<<grid_for(n), BLOCK, TILE_BYTES, ctx->stream>>>(battles, durations, results, idx1, idx2, rolls, actions, counts, n);
  ctx->launches++;
  return cudaGetLastError();
}

// training rows (train_rows.cu): stage the pre-update state of every stored update. `words` = two zeroed uint32:
// [0] the game ticket, [1] receives the lowest index of a game whose replay does not reproduce its stored result
cudaError_t oakb200_launch_replay(oakb200_ctx* ctx, const uint8_t* stream, const uint64_t* game_off, const uint32_t* row_base,
                                  const int32_t* only_update, uint32_t n_games, uint32_t* words, uint8_t* rows) {
  static_assert(ROW_STATE_BYTES == OAKB200_ROW_STATE_BYTES, "staged row layout");
  auto k = replay_kernel<BLOCK>;
  cudaError_t e = allow_smem(k);
  if (e != cudaSuccess) return e;
  int per_sm = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, BLOCK, TILE_BYTES);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  unsigned grid = (unsigned)(ctx->sm_count * per_sm);
  const unsigned need = grid_for(n_games);
  if (grid > need) grid = need;
  k<<<grid, BLOCK, TILE_BYTES, ctx->stream>>>(stream, game_off, row_base, only_update, n_games, words, rows, words + 1);
  ctx->launches++;
  return cudaGetLastError();
}

// the same build in one launch (replay_encode_kernel): each warp stages and then encodes its own rows
cudaError_t oakb200_launch_replay_encode(oakb200_ctx* ctx, const uint8_t* stream, const uint64_t* game_off, const uint32_t* row_base,
                                         const int32_t* only_update, uint32_t n_games, uint32_t* words, uint8_t* rows,
                                         const oakb200_training_rows* out) {
  auto k = replay_encode_kernel<BLOCK>;
  const size_t smem = TILE_BYTES + (BLOCK / 32) * ROW_STATE_BYTES;
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int per_sm = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, BLOCK, smem);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  unsigned grid = (unsigned)(ctx->sm_count * per_sm);
  const unsigned need = grid_for(n_games);
  if (grid > need) grid = need;
  k<<<grid, BLOCK, smem, ctx->stream>>>(stream, game_off, row_base, only_update, n_games, words, rows, words + 1, *out);
  ctx->launches++;
  return cudaGetLastError();
}

// one cooperative launch, two roles (replay_split_kernel): `control` = 5 zeroed words except [1] = 0xFFFFFFFF, `queue` =
// n_rows zeroed words
cudaError_t oakb200_launch_replay_split(oakb200_ctx* ctx, const uint8_t* stream, const uint64_t* game_off, const uint32_t* row_base,
                                        const int32_t* only_update, uint32_t n_games, uint32_t n_rows, uint32_t* control, uint32_t* queue,
                                        uint8_t* rows, const oakb200_training_rows* out) {
  auto k = replay_split_kernel<BLOCK>;
  cudaError_t e = allow_smem(k);
  if (e != cudaSuccess) return e;
  int per_sm = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, BLOCK, TILE_BYTES);
  if (e != cudaSuccess) return e;
  if (per_sm < 2) return cudaErrorLaunchOutOfResources;  // needs a producer and a consumer CTA per SM
  const unsigned grid = (unsigned)(ctx->sm_count * per_sm);  // every CTA resident: consumers may wait on producers
  unsigned producers = grid_for(n_games);
  int prod_per_sm = 1;  // producer CTAs per SM (the first CTAs placed); OAKB200_ROWS_PRODUCERS overrides, < per_sm
  if (const char* pp = getenv("OAKB200_ROWS_PRODUCERS")) prod_per_sm = atoi(pp);
  if (prod_per_sm < 1) prod_per_sm = 1;
  if (prod_per_sm > per_sm - 1) prod_per_sm = per_sm - 1;
  if (producers > (unsigned)(ctx->sm_count * prod_per_sm)) producers = (unsigned)(ctx->sm_count * prod_per_sm);
  uint32_t n_prod = producers;
  oakb200_training_rows o = *out;
  void* args[] = {(void*)&stream, (void*)&game_off, (void*)&row_base, (void*)&only_update, (void*)&n_games, (void*)&n_rows,
                  (void*)&n_prod, (void*)&control, (void*)&queue, (void*)&rows, (void*)&o};
  e = cudaLaunchCooperativeKernel((const void*)k, dim3(grid), dim3(BLOCK), args, TILE_BYTES, ctx->stream);
  if (e != cudaSuccess) return e;
  ctx->launches++;
  return cudaGetLastError();
}

// ---- batch of one for the scalar libpkmn symbols (pkmn_compat.cu). Staging layout in the pinned block:
//   [0,384) battle  [384,392) durations  392 c1  393 c2  394 result  [400,416) overrides  [416,432) actions
//   [432,436) log length  [448,448+256) log   [704,722) choices  [722,724) counts
static int scalar_pin(oakb200_ctx* ctx) {
  if (ctx->scalar_pin) return OAKB200_OK;
  CU(ctx, cudaMallocHost(&ctx->scalar_pin, 1024));
  return OAKB200_OK;
}

int oakb200_scalar_update(oakb200_ctx* ctx, uint8_t* battle, uint8_t c1, uint8_t c2, uint8_t* durations, const uint8_t* overrides,
                          uint8_t* actions, uint8_t* result, uint8_t* log, uint32_t log_cap, uint32_t* log_len) {
  if (!ctx || !battle || !result) return OAKB200_EINVAL;
  CU(ctx, cudaSetDevice(ctx->device));
  if (int rc = scalar_pin(ctx)) return rc;
  uint8_t* pin = ctx->scalar_pin;
  cudaStream_t st = ctx->stream;
  memcpy(pin, battle, 384);
  if (durations) memcpy(pin + 384, durations, 8); else memset(pin + 384, 0, 8);
  pin[392] = c1;
  pin[393] = c2;
  pin[394] = RESULT_DEFAULT;  // the stored result byte only gates "battle already over"; the caller decides that
  if (overrides) memcpy(pin + 400, overrides, 16);
  CU(ctx, cudaMemcpyAsync(ctx->battles, pin, 384, cudaMemcpyHostToDevice, st));
  CU(ctx, cudaMemcpyAsync(ctx->durations, pin + 384, 8, cudaMemcpyHostToDevice, st));
  CU(ctx, cudaMemcpyAsync(ctx->c1, pin + 392, 1, cudaMemcpyHostToDevice, st));
  CU(ctx, cudaMemcpyAsync(ctx->c2, pin + 393, 1, cudaMemcpyHostToDevice, st));
  CU(ctx, cudaMemcpyAsync(ctx->results, pin + 394, 1, cudaMemcpyHostToDevice, st));
  if (overrides) CU(ctx, cudaMemcpyAsync(ctx->overrides, pin + 400, 16, cudaMemcpyHostToDevice, st));
  if (log) {
    if (log_cap > 256) log_cap = 256;  // one update writes at most PKMN_GEN1_MAX_LOGS = 180 bytes
    if (ctx->log_bytes < 512) {
      if (ctx->logs) CU(ctx, cudaFree(ctx->logs));
      ctx->logs = nullptr;
      ctx->log_bytes = 0;
      CU(ctx, cudaMalloc(&ctx->logs, 512));
      ctx->log_bytes = 512;
    }
    int rc = oakb200_step_log_device(ctx, 1, ctx->battles, ctx->durations, ctx->results, ctx->c1, ctx->c2,
                                     overrides ? ctx->overrides : nullptr, ctx->actions, ctx->logs + 16, log_cap, ctx->logs);
    if (rc) return rc;
    CU(ctx, cudaMemcpyAsync(pin + 432, ctx->logs, 4, cudaMemcpyDeviceToHost, st));
    if (log_cap) CU(ctx, cudaMemcpyAsync(pin + 448, ctx->logs + 16, log_cap, cudaMemcpyDeviceToHost, st));
  } else {
    int rc = oakb200_step_device(ctx, 1, ctx->battles, ctx->durations, ctx->results, ctx->c1, ctx->c2,
                                 overrides ? ctx->overrides : nullptr, ctx->actions);
    if (rc) return rc;
  }
  CU(ctx, cudaMemcpyAsync(pin, ctx->battles, 384, cudaMemcpyDeviceToHost, st));
  CU(ctx, cudaMemcpyAsync(pin + 384, ctx->durations, 8, cudaMemcpyDeviceToHost, st));
  CU(ctx, cudaMemcpyAsync(pin + 394, ctx->results, 1, cudaMemcpyDeviceToHost, st));
  CU(ctx, cudaMemcpyAsync(pin + 416, ctx->actions, 16, cudaMemcpyDeviceToHost, st));
  CU(ctx, cudaStreamSynchronize(st));
  ctx->n = 1;
  memcpy(battle, pin, 384);
  if (durations) memcpy(durations, pin + 384, 8);
  if (actions) memcpy(actions, pin + 416, 16);
  *result = pin[394];
  if (log) {
    uint32_t len;
    memcpy(&len, pin + 432, 4);
    if (len > log_cap) len = log_cap;
    memcpy(log, pin + 448, len);
    if (log_len) *log_len = len;
  }
  return OAKB200_OK;
}

int oakb200_scalar_choices(oakb200_ctx* ctx, const uint8_t* battle, uint8_t result, uint8_t* choices18, uint8_t* counts2) {
  if (!ctx || !battle) return OAKB200_EINVAL;
  CU(ctx, cudaSetDevice(ctx->device));
  if (int rc = scalar_pin(ctx)) return rc;
  uint8_t* pin = ctx->scalar_pin;
  cudaStream_t st = ctx->stream;
  memcpy(pin, battle, 384);
  pin[394] = result;
  CU(ctx, cudaMemcpyAsync(ctx->battles, pin, 384, cudaMemcpyHostToDevice, st));
  CU(ctx, cudaMemcpyAsync(ctx->results, pin + 394, 1, cudaMemcpyHostToDevice, st));
  int rc = oakb200_choices_device(ctx, 1, ctx->battles, ctx->results, ctx->choices, ctx->counts);
  if (rc) return rc;
  CU(ctx, cudaMemcpyAsync(pin + 704, ctx->choices, 18, cudaMemcpyDeviceToHost, st));
  CU(ctx, cudaMemcpyAsync(pin + 722, ctx->counts, 2, cudaMemcpyDeviceToHost, st));
  CU(ctx, cudaStreamSynchronize(st));
  ctx->n = 1;
  memcpy(choices18, pin + 704, 18);
  memcpy(counts2, pin + 722, 2);
  return OAKB200_OK;
}

extern "C" {

int oakb200_create(int device, size_t n_max, oakb200_ctx** out) {
  if (!out || n_max == 0 || n_max > 0x7FFFFFFFull) return OAKB200_EINVAL;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0 || device < 0 || device >= count) return OAKB200_ENODEVICE;
  oakb200_ctx* ctx = new (std::nothrow) oakb200_ctx();
  if (!ctx) return OAKB200_ECAPACITY;
  ctx->device = device;
  ctx->n_max = n_max;
  *out = ctx;
  CU(ctx, cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(ctx, cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return fail(ctx, OAKB200_ENODEVICE, "device is not sm_100 class");
  ctx->sm_count = prop.multiProcessorCount;
  CU(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  ctx->own_stream = true;
  CU(ctx, cudaMalloc(&ctx->battles, n_max * 384));
  CU(ctx, cudaMalloc(&ctx->durations, n_max * 8));
  CU(ctx, cudaMalloc(&ctx->results, n_max));
  CU(ctx, cudaMalloc(&ctx->c1, n_max));
  CU(ctx, cudaMalloc(&ctx->c2, n_max));
  CU(ctx, cudaMalloc(&ctx->overrides, n_max * 16));
  CU(ctx, cudaMalloc(&ctx->actions, n_max * 16));
  CU(ctx, cudaMalloc(&ctx->choices, n_max * 18));
  CU(ctx, cudaMalloc(&ctx->counts, n_max * 2));
  CU(ctx, cudaMalloc(&ctx->seeds, n_max * 8));
  CU(ctx, cudaMalloc(&ctx->score2, n_max));
  CU(ctx, cudaMalloc(&ctx->turns, n_max * 2));
  CU(ctx, cudaMalloc(&ctx->steps, n_max * 4));
  CU(ctx, cudaMalloc(&ctx->teams, n_max * 60));
  CU(ctx, cudaMalloc(&ctx->eval_values, n_max * 4));
  CU(ctx, cudaMalloc(&ctx->eval_logits, n_max * 18 * 4));
  CU(ctx, cudaMalloc(&ctx->eval_counts, n_max * 2));
  CU(ctx, cudaMalloc(&ctx->ticket, 64 * sizeof(uint32_t)));  // [0] the context stream's, [1..] one per pipelined chunk
  CU(ctx, cudaMalloc(&ctx->total_steps, sizeof(unsigned long long)));
  CU(ctx, cudaMemset(ctx->total_steps, 0, sizeof(unsigned long long)));
  return OAKB200_OK;
}

void oakb200_destroy(oakb200_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  void* bufs[] = {ctx->logs, ctx->battles, ctx->durations, ctx->results, ctx->c1,    ctx->c2,    ctx->overrides, ctx->actions, ctx->choices,
                  ctx->counts,  ctx->seeds,     ctx->score2,  ctx->turns, ctx->steps, ctx->teams,     ctx->ticket,  ctx->total_steps,
                  ctx->eval_values, ctx->eval_logits, ctx->eval_counts};
  oakb200_net_release(ctx);
  oakb200_frames_release(ctx);
  oakb200_rows_release(ctx);
  for (void* b : bufs)
    if (b) cudaFree(b);
  if (ctx->eval_pipe_ready) {
    cudaStreamDestroy(ctx->eval_h2d);
    cudaStreamDestroy(ctx->eval_d2h);
    for (cudaEvent_t ev : ctx->eval_ev)
      if (ev) cudaEventDestroy(ev);
  }
  if (ctx->scalar_pin) cudaFreeHost(ctx->scalar_pin);
  if (ctx->pipe_ready) {
    cudaStreamDestroy(ctx->copy_stream);
    cudaEventDestroy(ctx->pipe_start);
    cudaFreeHost(ctx->pipe_marks);
  }
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char* oakb200_last_error(const oakb200_ctx* ctx) { return ctx ? ctx->error.c_str() : "null context"; }
int oakb200_device(const oakb200_ctx* ctx) { return ctx ? ctx->device : -1; }
uint64_t oakb200_launch_count(const oakb200_ctx* ctx) { return ctx ? ctx->launches : 0; }

int oakb200_set_stream(oakb200_ctx* ctx, void* cuda_stream) {
  if (!ctx) return OAKB200_EINVAL;
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  ctx->stream = (cudaStream_t)cuda_stream;
  ctx->own_stream = false;
  return OAKB200_OK;
}

int oakb200_synchronize(oakb200_ctx* ctx) {
  if (!ctx) return OAKB200_EINVAL;
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}

int oakb200_upload(oakb200_ctx* ctx, size_t n, const uint8_t* battles, const uint8_t* durations, const uint8_t* results) {
  if (!ctx || !battles) return fail(ctx, OAKB200_EINVAL, "upload: null argument");
  if (n > ctx->n_max) return fail(ctx, OAKB200_ECAPACITY, "upload: n exceeds n_max");
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemcpyAsync(ctx->battles, battles, n * 384, cudaMemcpyHostToDevice, ctx->stream));
  if (durations) {
    CU(ctx, cudaMemcpyAsync(ctx->durations, durations, n * 8, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    CU(ctx, cudaMemsetAsync(ctx->durations, 0, n * 8, ctx->stream));
  }
  if (results) {
    CU(ctx, cudaMemcpyAsync(ctx->results, results, n, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    CU(ctx, cudaMemsetAsync(ctx->results, RESULT_DEFAULT, n, ctx->stream));
  }
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->n = n;
  return OAKB200_OK;
}

int oakb200_download(oakb200_ctx* ctx, size_t n, uint8_t* battles, uint8_t* durations, uint8_t* results) {
  if (!ctx) return OAKB200_EINVAL;
  if (n > ctx->n) return fail(ctx, OAKB200_EINVAL, "download: n exceeds resident count");
  CU(ctx, cudaSetDevice(ctx->device));
  if (battles) CU(ctx, cudaMemcpyAsync(battles, ctx->battles, n * 384, cudaMemcpyDeviceToHost, ctx->stream));
  if (durations) CU(ctx, cudaMemcpyAsync(durations, ctx->durations, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (results) CU(ctx, cudaMemcpyAsync(results, ctx->results, n, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}

int oakb200_resident(oakb200_ctx* ctx, void** battles, void** durations, void** results, size_t* n) {
  if (!ctx) return OAKB200_EINVAL;
  if (battles) *battles = ctx->battles;
  if (durations) *durations = ctx->durations;
  if (results) *results = ctx->results;
  if (n) *n = ctx->n;
  return OAKB200_OK;
}

int oakb200_choices_device(oakb200_ctx* ctx, size_t n, const void* battles, const void* results, void* out, void* counts) {
  if (!ctx || !battles || !results || !out || !counts) return fail(ctx, OAKB200_EINVAL, "choices: null argument");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  auto k = choices_kernel<BLOCK>;
  CU(ctx, allow_smem(k));
  k<<<grid_for(n), BLOCK, TILE_BYTES, ctx->stream>>>((const uint8_t*)battles, (const uint8_t*)results, (uint8_t*)out,
                                                     (uint8_t*)counts, (uint32_t)n);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

int oakb200_choices(oakb200_ctx* ctx, uint8_t* out, uint8_t* counts) {
  if (!ctx || !out || !counts) return fail(ctx, OAKB200_EINVAL, "choices: null argument");
  const size_t n = ctx->n;
  int rc = oakb200_choices_device(ctx, n, ctx->battles, ctx->results, ctx->choices, ctx->counts);
  if (rc) return rc;
  CU(ctx, cudaMemcpyAsync(out, ctx->choices, n * 18, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaMemcpyAsync(counts, ctx->counts, n * 2, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}

int oakb200_step_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* c1,
                        const void* c2, const void* overrides, void* actions_out) {
  if (!ctx || !battles || !durations || !results || !c1 || !c2) return fail(ctx, OAKB200_EINVAL, "step: null argument");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  if (overrides) {
    CU(ctx, (launch_step<false, true>(ctx, n, battles, durations, results, c1, c2, nullptr, 0, overrides, actions_out, nullptr)));
  } else {
    CU(ctx, (launch_step<false, false>(ctx, n, battles, durations, results, c1, c2, nullptr, 0, nullptr, actions_out, nullptr)));
  }
  return OAKB200_OK;
}

int oakb200_step(oakb200_ctx* ctx, const uint8_t* c1, const uint8_t* c2, const uint8_t* overrides, uint8_t* results_out,
                 uint8_t* actions_out) {
  if (!ctx || !c1 || !c2) return fail(ctx, OAKB200_EINVAL, "step: null argument");
  const size_t n = ctx->n;
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemcpyAsync(ctx->c1, c1, n, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(ctx->c2, c2, n, cudaMemcpyHostToDevice, ctx->stream));
  if (overrides) CU(ctx, cudaMemcpyAsync(ctx->overrides, overrides, n * 16, cudaMemcpyHostToDevice, ctx->stream));
  int rc = oakb200_step_device(ctx, n, ctx->battles, ctx->durations, ctx->results, ctx->c1, ctx->c2,
                               overrides ? ctx->overrides : nullptr, actions_out ? ctx->actions : nullptr);
  if (rc) return rc;
  if (results_out) CU(ctx, cudaMemcpyAsync(results_out, ctx->results, n, cudaMemcpyDeviceToHost, ctx->stream));
  if (actions_out) CU(ctx, cudaMemcpyAsync(actions_out, ctx->actions, n * 16, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}

int oakb200_step_log_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* c1,
                            const void* c2, const void* overrides, void* actions_out, void* logs, uint32_t log_stride,
                            void* log_lens) {
  if (!ctx || !battles || !durations || !results || !c1 || !c2 || !logs || !log_lens)
    return fail(ctx, OAKB200_EINVAL, "step_log: null argument");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  auto k = step_log_kernel<BLOCK>;
  CU(ctx, allow_smem(k));
  k<<<grid_for(n), BLOCK, TILE_BYTES, ctx->stream>>>((uint8_t*)battles, (uint8_t*)durations, (uint8_t*)results, (const uint8_t*)c1,
                                                     (const uint8_t*)c2, (const uint8_t*)overrides, (uint8_t*)actions_out,
                                                     (uint8_t*)logs, log_stride, (uint32_t*)log_lens, (uint32_t)n);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

int oakb200_step_log(oakb200_ctx* ctx, const uint8_t* c1, const uint8_t* c2, const uint8_t* overrides, uint8_t* results_out,
                     uint8_t* actions_out, uint8_t* logs_out, uint32_t log_stride, uint32_t* log_lens_out) {
  if (!ctx || !c1 || !c2 || !logs_out || !log_lens_out) return fail(ctx, OAKB200_EINVAL, "step_log: null argument");
  const size_t n = ctx->n;
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  const size_t need = n * (size_t)log_stride + n * 4;
  if (ctx->log_bytes < need) {
    if (ctx->logs) CU(ctx, cudaFree(ctx->logs));
    ctx->logs = nullptr;
    ctx->log_bytes = 0;
    CU(ctx, cudaMalloc(&ctx->logs, need));
    ctx->log_bytes = need;
  }
  uint8_t* lens = ctx->logs;  // [n] u32 first (aligned), then the log rows
  uint8_t* rows = ctx->logs + n * 4;
  CU(ctx, cudaMemcpyAsync(ctx->c1, c1, n, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(ctx->c2, c2, n, cudaMemcpyHostToDevice, ctx->stream));
  if (overrides) CU(ctx, cudaMemcpyAsync(ctx->overrides, overrides, n * 16, cudaMemcpyHostToDevice, ctx->stream));
  int rc = oakb200_step_log_device(ctx, n, ctx->battles, ctx->durations, ctx->results, ctx->c1, ctx->c2,
                                   overrides ? ctx->overrides : nullptr, actions_out ? ctx->actions : nullptr, rows, log_stride, lens);
  if (rc) return rc;
  if (results_out) CU(ctx, cudaMemcpyAsync(results_out, ctx->results, n, cudaMemcpyDeviceToHost, ctx->stream));
  if (actions_out) CU(ctx, cudaMemcpyAsync(actions_out, ctx->actions, n * 16, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaMemcpyAsync(logs_out, rows, n * (size_t)log_stride, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaMemcpyAsync(log_lens_out, lens, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}

int oakb200_step_random_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* seeds,
                               uint32_t step_index, void* actions_out, void* chosen_out) {
  if (!ctx || !battles || !durations || !results || !seeds) return fail(ctx, OAKB200_EINVAL, "step_random: null argument");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, (launch_step<true, false>(ctx, n, battles, durations, results, nullptr, nullptr, seeds, step_index, nullptr,
                                    actions_out, chosen_out)));
  return OAKB200_OK;
}

int oakb200_rollout_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* seeds,
                           uint32_t max_steps, uint32_t flags, void* score2_out, void* turns_out, void* steps_out) {
  if (!ctx || !battles || !durations || !results || !seeds) return fail(ctx, OAKB200_EINVAL, "rollout: null argument");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  const bool write = (flags & OAKB200_ROLLOUT_WRITE_STATE) != 0;
  const bool chance = (flags & OAKB200_ROLLOUT_TRACK_CHANCE) != 0;
  cudaError_t e;
  if (write) {
    e = chance ? launch_rollout<true, true>(ctx, n, battles, durations, results, seeds, max_steps, score2_out, turns_out, steps_out)
               : launch_rollout<false, true>(ctx, n, battles, durations, results, seeds, max_steps, score2_out, turns_out, steps_out);
  } else {
    e = chance ? launch_rollout<true, false>(ctx, n, battles, durations, results, seeds, max_steps, score2_out, turns_out, steps_out)
               : launch_rollout<false, false>(ctx, n, battles, durations, results, seeds, max_steps, score2_out, turns_out, steps_out);
  }
  CU(ctx, e);
  return OAKB200_OK;
}

int oakb200_rollout(oakb200_ctx* ctx, const uint64_t* seeds, uint32_t max_steps, uint32_t flags, uint8_t* score2_out,
                    uint16_t* turns_out, uint32_t* steps_out) {
  if (!ctx || !seeds) return fail(ctx, OAKB200_EINVAL, "rollout: null argument");
  const size_t n = ctx->n;
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemcpyAsync(ctx->seeds, seeds, n * 8, cudaMemcpyHostToDevice, ctx->stream));
  int rc = oakb200_rollout_device(ctx, n, ctx->battles, ctx->durations, ctx->results, ctx->seeds, max_steps, flags,
                                  ctx->score2, ctx->turns, ctx->steps);
  if (rc) return rc;
  if (score2_out) CU(ctx, cudaMemcpyAsync(score2_out, ctx->score2, n, cudaMemcpyDeviceToHost, ctx->stream));
  if (turns_out) CU(ctx, cudaMemcpyAsync(turns_out, ctx->turns, n * 2, cudaMemcpyDeviceToHost, ctx->stream));
  if (steps_out) CU(ctx, cudaMemcpyAsync(steps_out, ctx->steps, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}

// Host buffers in, host buffers out, with the upload hidden behind the playouts: every chunk copy is queued on a copy
// stream, each followed by a 4-byte copy that raises the device-side "records available" counter; ONE persistent rollout
// launch then consumes records as they land (wait_available in battle_kernels.cuh). Same results as
// oakb200_upload + oakb200_rollout.
int oakb200_rollout_host(oakb200_ctx* ctx, size_t n, const uint8_t* battles, const uint8_t* durations, const uint8_t* results,
                         const uint64_t* seeds, uint32_t max_steps, uint8_t* score2_out, uint16_t* turns_out, uint32_t* steps_out) {
  if (!ctx || !battles || !seeds) return fail(ctx, OAKB200_EINVAL, "rollout_host: null argument");
  if (n > ctx->n_max) return fail(ctx, OAKB200_ECAPACITY, "rollout_host: n exceeds n_max");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  const int MAX_CHUNKS = 32;
  if (!ctx->pipe_ready) {
    CU(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    CU(ctx, cudaEventCreateWithFlags(&ctx->pipe_start, cudaEventDisableTiming));
    CU(ctx, cudaMallocHost(&ctx->pipe_marks, (MAX_CHUNKS + 1) * sizeof(uint32_t)));
    ctx->pipe_ready = true;
  }
  size_t chunk = (size_t)1 << 16;
  if (const char* c = getenv("OAKB200_PIPE_CHUNK")) chunk = (size_t)atol(c) > 0 ? (size_t)atol(c) : chunk;  // tuning knob
  if ((n + chunk - 1) / chunk > (size_t)MAX_CHUNKS) chunk = (n + MAX_CHUNKS - 1) / MAX_CHUNKS;
  // chunk boundaries on 128 records: no 32-byte sector of results / seeds / durations ever spans two copies
  chunk = (chunk + 127) / 128 * 128;
  const int chunks = (int)((n + chunk - 1) / chunk);
  uint32_t* avail = ctx->ticket + 2;  // [0] records available, [1] "a lane gave up waiting"
  // the counter is cleared in context-stream order; the copies start after that and after everything queued before
  CU(ctx, cudaMemsetAsync(avail, 0, 2 * sizeof(uint32_t), ctx->stream));
  // defaults are filled here, before the launch: the copy stream must carry DMA copies only, nothing that needs an SM
  if (!durations) CU(ctx, cudaMemsetAsync(ctx->durations, 0, n * 8, ctx->stream));
  if (!results) CU(ctx, cudaMemsetAsync(ctx->results, RESULT_DEFAULT, n, ctx->stream));
  CU(ctx, cudaEventRecord(ctx->pipe_start, ctx->stream));
  CU(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->pipe_start, 0));
  cudaStream_t cs = ctx->copy_stream;
  for (int k = 0; k < chunks; k++) {
    const size_t at = (size_t)k * chunk, m = n - at < chunk ? n - at : chunk;
    CU(ctx, cudaMemcpyAsync(ctx->battles + at * 384, battles + at * 384, m * 384, cudaMemcpyHostToDevice, cs));
    if (durations) CU(ctx, cudaMemcpyAsync(ctx->durations + at * 8, durations + at * 8, m * 8, cudaMemcpyHostToDevice, cs));
    if (results) CU(ctx, cudaMemcpyAsync(ctx->results + at, results + at, m, cudaMemcpyHostToDevice, cs));
    CU(ctx, cudaMemcpyAsync(ctx->seeds + at, seeds + at, m * 8, cudaMemcpyHostToDevice, cs));
    ctx->pipe_marks[k] = (uint32_t)(at + m);
    CU(ctx, cudaMemcpyAsync(avail, ctx->pipe_marks + k, sizeof(uint32_t), cudaMemcpyHostToDevice, cs));
  }
  CU(ctx, (launch_rollout<false, false>(ctx, n, ctx->battles, ctx->durations, ctx->results, ctx->seeds, max_steps, ctx->score2,
                                        ctx->turns, ctx->steps, avail)));
  if (score2_out) CU(ctx, cudaMemcpyAsync(score2_out, ctx->score2, n, cudaMemcpyDeviceToHost, ctx->stream));
  if (turns_out) CU(ctx, cudaMemcpyAsync(turns_out, ctx->turns, n * 2, cudaMemcpyDeviceToHost, ctx->stream));
  if (steps_out) CU(ctx, cudaMemcpyAsync(steps_out, ctx->steps, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  uint32_t* gave_up = ctx->pipe_marks + MAX_CHUNKS;
  CU(ctx, cudaMemcpyAsync(gave_up, avail + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->copy_stream));
  ctx->n = n;
  if (*gave_up) return fail(ctx, OAKB200_ECUDA, "rollout_host: the upload stalled; playouts abandoned");
  return OAKB200_OK;
}

int oakb200_total_steps(oakb200_ctx* ctx, uint64_t* out) {
  if (!ctx || !out) return OAKB200_EINVAL;
  CU(ctx, cudaSetDevice(ctx->device));
  unsigned long long v = 0;
  CU(ctx, cudaMemcpyAsync(&v, ctx->total_steps, sizeof(v), cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  *out = v;
  return OAKB200_OK;
}

int oakb200_randomize_hidden_device(oakb200_ctx* ctx, size_t n, void* battles, const void* durations) {
  if (!ctx || !battles || !durations) return fail(ctx, OAKB200_EINVAL, "randomize_hidden: null argument");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  randomize_hidden_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>((uint8_t*)battles, (const uint8_t*)durations, (uint32_t)n);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

int oakb200_build_battles(oakb200_ctx* ctx, size_t n, const uint8_t* teams, const uint64_t* seeds) {
  if (!ctx || !teams || !seeds) return fail(ctx, OAKB200_EINVAL, "build_battles: null argument");
  if (n > ctx->n_max) return fail(ctx, OAKB200_ECAPACITY, "build_battles: n exceeds n_max");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemcpyAsync(ctx->teams, teams, n * 60, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(ctx->seeds, seeds, n * 8, cudaMemcpyHostToDevice, ctx->stream));
  auto k = build_kernel<BLOCK>;
  CU(ctx, allow_smem(k));
  k<<<grid_for(n), BLOCK, TILE_BYTES, ctx->stream>>>(ctx->teams, ctx->seeds, ctx->battles, ctx->durations, ctx->results, (uint32_t)n, 0u);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->n = n;
  return OAKB200_OK;
}

int oakb200_respawn_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* teams,
                           const void* seeds) {
  if (!ctx || !battles || !durations || !results || !teams || !seeds) return fail(ctx, OAKB200_EINVAL, "respawn: null argument");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  auto k = build_kernel<BLOCK>;
  CU(ctx, allow_smem(k));
  k<<<grid_for(n), BLOCK, TILE_BYTES, ctx->stream>>>((const uint8_t*)teams, (const uint64_t*)seeds, (uint8_t*)battles,
                                                     (uint8_t*)durations, (uint8_t*)results, (uint32_t)n, 1u);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

}  // extern "C"
