// Training rows on the GPU: stored games -> the dense tensors the learner trains on.
//
// Reference behaviour: Py::Battle::EncodedFrames::{write, uncompress_from_bytes, from_bytes}
// (cpp/include/py/battle/encoded-frames.h:53-138), Target::write / Update::write_to_tensor (py/battle/target.h:56-70,
// train/battle/compressed-frame.h:141-163), the dense encoders (cpp/include/encode/battle/battle.h:22-29,64-71,146-151,
// 171-178,201-206,251-264,297-326,375-400,465-478), Policy::get_index (encode/battle/policy.h:29-58), PKMN::score
// (libpkmn/pkmn.h:174-190), pyoak `sample` (cpp/src/pyoak.cc:111-260).
//
// Two kernels. replay_kernel (battle_kernels.cuh, launched through oakb200.cu) advances every stored game with its
// stored choices and stages the pre-update state of each row (432 B). rows_encode_kernel below turns one staged state
// into one row: 11.7 KB of floats of which ~190 are non-zero. It is an HBM-write stream and is built as one: a warp per
// row clears the row with 16-byte stores, then scatters the non-zeros over it with every lane busy on an equal task
// (one stat, one move slot, one volatile bit ...) rather than one lane walking a whole Pokemon. The row lives nowhere on
// chip — the first version built it in a shared-memory image and sent it out with a bulk copy, which capped the SM at
// 16 warps and left each of them latency-bound on two busy lanes (profiles/r1k_training_rows.txt) — so occupancy is
// bounded by registers only, and the next row's state is prefetched while this one is written.
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "context.h"
#include "gen1_state.cuh"
#include "train_rows_encode.cuh"

using namespace oakb200;

using namespace oakb200::rows;

namespace {

const int OAKB200_ROWS_DEFAULT_MODE = 0;
const int ENC_WARPS = 8;
static_assert(STATE_BYTES == OAKB200_ROW_STATE_BYTES, "staged row layout");

__global__ void __launch_bounds__(ENC_WARPS * 32, 4)
rows_encode_kernel(const uint8_t* __restrict__ states, const uint8_t* __restrict__ stream, oakb200_training_rows out, uint32_t n_rows) {
  __shared__ __align__(16) uint8_t s_state[ENC_WARPS][STATE_BYTES];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t* st = s_state[warp];
  const uint32_t stride = gridDim.x * ENC_WARPS;
  uint32_t row = blockIdx.x * ENC_WARPS + warp;
  uint4 next = make_uint4(0, 0, 0, 0);
  if (row < n_rows && lane < STATE_BYTES / 16) next = reinterpret_cast<const uint4*>(states + (size_t)row * STATE_BYTES)[lane];

  for (; row < n_rows; row += stride) {
    if (lane < STATE_BYTES / 16) reinterpret_cast<uint4*>(st)[lane] = next;
    if (row + stride < n_rows && lane < STATE_BYTES / 16)
      next = reinterpret_cast<const uint4*>(states + (size_t)(row + stride) * STATE_BYTES)[lane];
    __syncwarp();
    encode_row(st, stream, out, row, lane);
    __syncwarp();  // every lane is done with this state before the next one overwrites it
  }
}

}  // namespace

// device scratch kept between calls (grown on demand)
struct oakb200_rows {
  uint8_t* states = nullptr;  // [rows_cap][432]
  size_t rows_cap = 0;
  uint32_t* words = nullptr;  // [0] ticket, [1] first bad game, [2] queue tail, [3] queue head, [4] a consumer gave up
  uint32_t* queue = nullptr;  // [rows_cap] work queue of the one-launch form
  // host entry point only
  uint8_t* stream = nullptr;
  size_t stream_cap = 0;
  uint64_t* game_off = nullptr;
  uint32_t* row_base = nullptr;
  int32_t* only = nullptr;
  size_t games_cap = 0;
  std::vector<uint64_t> h_off;
  std::vector<uint32_t> h_base;
};

void oakb200_rows_release(oakb200_ctx* ctx) {
  oakb200_rows* r = ctx->rows;
  if (!r) return;
  cudaFree(r->states);
  cudaFree(r->words);
  cudaFree(r->queue);
  cudaFree(r->stream);
  cudaFree(r->game_off);
  cudaFree(r->row_base);
  cudaFree(r->only);
  delete r;
  ctx->rows = nullptr;
}

namespace {

int rows_scratch(oakb200_ctx* ctx, size_t n_rows) {
  if (!ctx->rows) {
    ctx->rows = new (std::nothrow) oakb200_rows();
    if (!ctx->rows) return fail(ctx, OAKB200_ECAPACITY, "rows: out of host memory");
  }
  oakb200_rows* r = ctx->rows;
  if (!r->words) CU(ctx, cudaMalloc(&r->words, 8 * sizeof(uint32_t)));
  if (n_rows > r->rows_cap) {
    // the previous build may still be reading the old buffer
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(r->states);
    cudaFree(r->queue);
    r->states = nullptr;
    r->queue = nullptr;
    r->rows_cap = 0;
    CU(ctx, cudaMalloc(&r->states, n_rows * (size_t)OAKB200_ROW_STATE_BYTES));
    CU(ctx, cudaMalloc(&r->queue, n_rows * sizeof(uint32_t)));
    r->rows_cap = n_rows;
  }
  return OAKB200_OK;
}

}  // namespace

extern "C" {

int oakb200_rows_index_host(const uint8_t* stream, uint64_t n_bytes, uint64_t* game_off, uint32_t* row_base, uint64_t capacity,
                            uint64_t* n_games, uint64_t* n_rows) {
  if (!stream && n_bytes) return OAKB200_EINVAL;
  uint64_t at = 0, games = 0, rows = 0;
  while (at < n_bytes) {
    if (at + 391 > n_bytes) return OAKB200_EINVAL;
    uint32_t nb;
    uint16_t nu;
    memcpy(&nb, stream + at, 4);
    memcpy(&nu, stream + at + 4, 2);
    // an update is 11 + 4 (m + n) bytes with 1 <= m, n <= 9 (compressed-frame.h:58-75)
    if (nb < 391 + 19ull * nu || nb > 391 + 83ull * nu || at + nb > n_bytes) return OAKB200_EINVAL;
    if (game_off || row_base) {
      if (games >= capacity) return OAKB200_ECAPACITY;
      if (game_off) game_off[games] = at;
      if (row_base) row_base[games] = (uint32_t)rows;
    }
    games++;
    rows += nu;
    at += nb;
  }
  if (rows > 0xFFFFFFFFull) return OAKB200_ECAPACITY;
  if (n_games) *n_games = games;
  if (n_rows) *n_rows = rows;
  return OAKB200_OK;
}

int oakb200_rows_build_device(oakb200_ctx* ctx, const uint8_t* stream, uint64_t n_bytes, const uint64_t* game_off,
                              const uint32_t* row_base, const int32_t* only_update, uint64_t n_games, uint64_t n_rows,
                              const oakb200_training_rows* out) {
  if (!ctx) return OAKB200_EINVAL;
  if (!out || (!stream && n_games) || (!game_off && n_games) || (!row_base && n_games) || n_bytes >= (1ull << 32) ||
      n_games > 0x7FFFFFFFull || n_rows > 0x7FFFFFFFull)
    return fail(ctx, OAKB200_EINVAL, "rows_build_device: bad arguments");
  CU(ctx, cudaSetDevice(ctx->device));
  if (int rc = rows_scratch(ctx, n_rows ? n_rows : 1)) return rc;
  oakb200_rows* r = ctx->rows;
  const uint32_t init[8] = {0u, 0xFFFFFFFFu, 0u, 0u, 0u, 0u, 0u, 0u};
  CU(ctx, cudaMemcpyAsync(r->words, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
  if (n_games == 0 || n_rows == 0) return OAKB200_OK;
  // Three forms of the same build, all bit-exact (OAKB200_ROWS_MODE): 0 = replay_kernel then rows_encode_kernel; 1 = one
  // launch in which every warp alternates replaying and encoding its own rows (replay_encode_kernel: measured slower,
  // 33.1 ms against 17.0 ms for 3.98 M rows); 2 = one cooperative launch of producer CTAs (replay) and consumer CTAs
  // (encode) joined by a work queue (replay_split_kernel). Numbers in profiles/r1k_training_rows.txt.
  static const int mode = [] {
    const char* f = getenv("OAKB200_ROWS_MODE");
    return f ? atoi(f) : OAKB200_ROWS_DEFAULT_MODE;
  }();
  if (mode == 2) {
    CU(ctx, cudaMemsetAsync(r->queue, 0, n_rows * sizeof(uint32_t), ctx->stream));
    CU(ctx, oakb200_launch_replay_split(ctx, stream, game_off, row_base, only_update, (uint32_t)n_games, (uint32_t)n_rows, r->words, r->queue,
                                        r->states, out));
    return OAKB200_OK;
  }
  if (mode == 1) {
    CU(ctx, oakb200_launch_replay_encode(ctx, stream, game_off, row_base, only_update, (uint32_t)n_games, r->words, r->states, out));
    return OAKB200_OK;
  }
  CU(ctx, oakb200_launch_replay(ctx, stream, game_off, row_base, only_update, (uint32_t)n_games, r->words, r->states));
  // persistent: the resident CTAs of the chip, rows interleaved over the warps of the grid so that neighbouring rows are
  // written together
  static int per_sm = 0;
  if (per_sm == 0) {
    CU(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rows_encode_kernel, ENC_WARPS * 32, 0));
    if (per_sm < 1) per_sm = 1;
  }
  unsigned grid = (unsigned)(per_sm * ctx->sm_count);
  const unsigned need = (unsigned)((n_rows + ENC_WARPS - 1) / ENC_WARPS);
  if (grid > need) grid = need;
  rows_encode_kernel<<<grid, ENC_WARPS * 32, 0, ctx->stream>>>(r->states, stream, *out, (uint32_t)n_rows);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

int oakb200_rows_status(oakb200_ctx* ctx, int64_t* first_bad_game) {
  if (!ctx) return OAKB200_EINVAL;
  if (!ctx->rows || !ctx->rows->words) return fail(ctx, OAKB200_EINVAL, "rows_status: nothing was built");
  CU(ctx, cudaSetDevice(ctx->device));
  uint32_t w[8];
  CU(ctx, cudaMemcpyAsync(w, ctx->rows->words, sizeof(w), cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  if (w[4]) return fail(ctx, OAKB200_ECUDA, "rows: an encode warp gave up waiting for its row (internal error)");
  if (first_bad_game) *first_bad_game = w[1] == 0xFFFFFFFFu ? -1 : (int64_t)w[1];
  return OAKB200_OK;
}

int oakb200_rows_build_host(oakb200_ctx* ctx, const uint8_t* stream, uint64_t n_bytes, const int32_t* only_update,
                            const oakb200_training_rows* out, uint64_t capacity_rows, uint64_t* n_games, uint64_t* n_rows,
                            int64_t* first_bad_game) {
  if (!ctx) return OAKB200_EINVAL;
  if (!out || (!stream && n_bytes) || n_bytes >= (1ull << 32)) return fail(ctx, OAKB200_EINVAL, "rows_build_host: bad arguments");
  CU(ctx, cudaSetDevice(ctx->device));
  if (int rc = rows_scratch(ctx, 1)) return rc;
  oakb200_rows* r = ctx->rows;
  uint64_t games = 0, rows = 0;
  int rc = oakb200_rows_index_host(stream, n_bytes, nullptr, nullptr, 0, &games, &rows);
  if (rc != OAKB200_OK) return fail(ctx, rc, "rows_build_host: the stream does not split into whole games");
  const uint64_t out_rows = only_update ? games : rows;
  if (out_rows > capacity_rows) return fail(ctx, OAKB200_ECAPACITY, "rows_build_host: more rows than the output arrays hold");
  r->h_off.resize(games);
  r->h_base.resize(games);
  oakb200_rows_index_host(stream, n_bytes, r->h_off.data(), r->h_base.data(), games, nullptr, nullptr);
  if (only_update)
    for (uint64_t g = 0; g < games; g++) r->h_base[g] = (uint32_t)g;
  if (n_bytes > r->stream_cap || games > r->games_cap) CU(ctx, cudaStreamSynchronize(ctx->stream));
  if (n_bytes > r->stream_cap) {
    cudaFree(r->stream);
    r->stream = nullptr;
    r->stream_cap = 0;
    CU(ctx, cudaMalloc(&r->stream, n_bytes));
    r->stream_cap = n_bytes;
  }
  if (games > r->games_cap) {
    cudaFree(r->game_off);
    cudaFree(r->row_base);
    cudaFree(r->only);
    r->game_off = nullptr;
    r->row_base = nullptr;
    r->only = nullptr;
    r->games_cap = 0;
    CU(ctx, cudaMalloc(&r->game_off, games * sizeof(uint64_t)));
    CU(ctx, cudaMalloc(&r->row_base, games * sizeof(uint32_t)));
    CU(ctx, cudaMalloc(&r->only, games * sizeof(int32_t)));
    r->games_cap = games;
  }
  if (games) {
    CU(ctx, cudaMemcpyAsync(r->stream, stream, n_bytes, cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaMemcpyAsync(r->game_off, r->h_off.data(), games * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaMemcpyAsync(r->row_base, r->h_base.data(), games * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    if (only_update) CU(ctx, cudaMemcpyAsync(r->only, only_update, games * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
  }
  rc = oakb200_rows_build_device(ctx, r->stream, n_bytes, r->game_off, r->row_base, only_update ? r->only : nullptr, games, out_rows, out);
  if (rc != OAKB200_OK) return rc;
  if (n_games) *n_games = games;
  if (n_rows) *n_rows = out_rows;
  return oakb200_rows_status(ctx, first_bad_game);
}

// ---- oakb200_rows_build_static: the two-kernel build above with its scratch in a file-scope device array -------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc). Every GPU that loads
// this library holds the array until the process exits: 256 control bytes + OAKB200_ROWS_STATIC_MAX staged states of 432 B
// = 6.75 MiB. In HOST code `g_rows_static` is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, looked up once per device with that device current (rows_static_base below).
static __device__ __align__(256) uint8_t g_rows_static[256 + (size_t)OAKB200_ROWS_STATIC_MAX * OAKB200_ROW_STATE_BYTES];

namespace {

const int STATIC_MAX_DEVICES = 64;
uint8_t* g_rows_static_addr[STATIC_MAX_DEVICES];  // per device, 0 = not looked up yet (benign race: same value every time)

int rows_static_base(oakb200_ctx* ctx, uint8_t** base) {
  if (ctx->device < 0 || ctx->device >= STATIC_MAX_DEVICES) return fail(ctx, OAKB200_EINVAL, "rows_build_static: device index out of range");
  uint8_t* p = g_rows_static_addr[ctx->device];
  if (!p) {
    CU(ctx, cudaGetSymbolAddress((void**)&p, g_rows_static));  // the context's device is current (caller)
    if (!p) return fail(ctx, OAKB200_ECUDA, "rows_build_static: cudaGetSymbolAddress gave a null address");
    g_rows_static_addr[ctx->device] = p;
  }
  *base = p;
  return OAKB200_OK;
}

}  // namespace

int oakb200_rows_build_static(oakb200_ctx* ctx, const uint8_t* stream, uint64_t n_bytes, const uint64_t* game_off,
                              const uint32_t* row_base, const int32_t* only_update, uint64_t n_games, uint64_t n_rows,
                              const oakb200_training_rows* out, uint32_t* first_bad_game_out) {
  if (!ctx) return OAKB200_EINVAL;
  if (!out || (!stream && n_games) || (!game_off && n_games) || (!row_base && n_games) || n_bytes >= (1ull << 32) ||
      n_games > 0x7FFFFFFFull)
    return fail(ctx, OAKB200_EINVAL, "rows_build_static: bad arguments");
  if (n_rows > (uint64_t)OAKB200_ROWS_STATIC_MAX)
    return fail(ctx, OAKB200_ECAPACITY, "rows_build_static: n_rows exceeds OAKB200_ROWS_STATIC_MAX (16384); use oakb200_rows_build_device");
  CU(ctx, cudaSetDevice(ctx->device));
  uint8_t* base = nullptr;
  if (int rc = rows_static_base(ctx, &base)) return rc;
  uint32_t* words = reinterpret_cast<uint32_t*>(base);  // [0] game ticket, [1] lowest bad game; the array keeps the last call's
  uint8_t* states = base + 256;                          // contents, so both are reset here, in stream order
  CU(ctx, cudaMemsetAsync(words, 0, sizeof(uint32_t), ctx->stream));
  CU(ctx, cudaMemsetAsync(words + 1, 0xFF, sizeof(uint32_t), ctx->stream));
  if (n_games != 0 && n_rows != 0) {
    // every staged state the encoder reads is written by this call's replay first: nothing stale is consumed
    CU(ctx, oakb200_launch_replay(ctx, stream, game_off, row_base, only_update, (uint32_t)n_games, words, states));
    int per_sm = 0;
    CU(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rows_encode_kernel, ENC_WARPS * 32, 0));
    if (per_sm < 1) per_sm = 1;
    unsigned grid = (unsigned)(per_sm * ctx->sm_count);
    const unsigned need = (unsigned)((n_rows + ENC_WARPS - 1) / ENC_WARPS);
    if (grid > need) grid = need;
    rows_encode_kernel<<<grid, ENC_WARPS * 32, 0, ctx->stream>>>(states, stream, *out, (uint32_t)n_rows);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
  }
  if (first_bad_game_out)
    CU(ctx, cudaMemcpyAsync(first_bad_game_out, words + 1, sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
  return OAKB200_OK;
}

}  // extern "C"
