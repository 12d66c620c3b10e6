// Training-frame recorder and writer on the device, plus the policy post-processing / sampling step that
// feeds it. Replaces, for many concurrent games:
//   Train::Battle::CompressedFramesImpl<>::{Update::write, write, n_bytes} + compress_probs
//       (reference cpp/include/train/battle/compressed-frame.h:12-24,70-112,173-199)
//   BattleFrameBuffer::write_frames (cpp/include/util/battle-frame-buffer.h:41-46)
//   RuntimePolicy::get_policy temperature / floor / renormalise + mt19937::sample_pdf
//       (cpp/include/util/policy.h:76-107, cpp/include/util/random.h:41-50) and softmax (search/util/softmax.h:5-15)
//
// Every game slot owns a byte buffer that always holds the reference's serialisation of the game so far:
//   u32 n_bytes | u16 n_updates | battle[384] | result u8 | updates...
//   update = u8 (m-1)|(n-1)<<4 | c1 | c2 | u32 iterations | u16 empirical value | u16 nash value |
//            u16 p1 empirical[m] | u16 p1 nash[m] | u16 p2 empirical[n] | u16 p2 nash[n]        (11 + 4(m+n) bytes)
// record appends one update per live game, finish fills the header, drain concatenates the finished games
// in slot order (size scan + one copy CTA per game) into a stream that is byte-identical to what the
// reference writes into an N.battle.data file, and frees their slots.
#include <cuda_runtime.h>

#include "context.h"

struct oakb200_frames {
  size_t games = 0;
  uint32_t cap = 0, max_updates = 0;
  uint8_t* buf = nullptr;     // games * cap
  uint32_t* used = nullptr;   // bytes used per slot
  uint16_t* nupd = nullptr;
  uint8_t* state = nullptr;   // 0 idle, 1 recording, 2 finished, 3 overflowed (dropped when it ends)
  uint64_t* offsets = nullptr;  // games + 1 (exclusive scan of finished sizes, total last)
  uint32_t* counters = nullptr; // [0] finished games waiting, [1] dropped games
  uint8_t* pack = nullptr;
  size_t pack_cap = 0;
};

namespace {

enum : uint8_t { FS_IDLE = 0, FS_RECORDING = 1, FS_FINISHED = 2, FS_OVERFLOW = 3 };
const uint32_t HEADER = 4 + 2 + 384 + 1;

__device__ __forceinline__ uint16_t compress_prob(float x) {  // compress_probs<double, uint16_t>: x * 65535, truncated
  const double v = (double)x * 65535.0;
  return (uint16_t)(v <= 0.0 ? 0.0 : (v >= 65535.0 ? 65535.0 : v));
}

// one warp per slot: start recording every idle slot whose battle is not over
__global__ void frames_begin_kernel(uint8_t* buf, uint32_t* used, uint16_t* nupd, uint8_t* state, uint32_t cap,
                                    const uint8_t* __restrict__ battles, const uint8_t* __restrict__ results, uint32_t n) {
  const uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (g >= n) return;
  if (state[g] != FS_IDLE || (results[g] & 15u) != 0) return;
  uint8_t* dst = buf + (size_t)g * cap;
  const uint16_t* src = reinterpret_cast<const uint16_t*>(battles + (size_t)g * 384);
  for (int i = lane; i < 192; i += 32) *reinterpret_cast<uint16_t*>(dst + 6 + 2 * i) = src[i];  // header is 6 bytes: 2-byte aligned
  __syncwarp();
  if (lane == 0) {
    used[g] = HEADER;
    nupd[g] = 0;
    state[g] = FS_RECORDING;
  }
}

// one WARP per slot: append the update of this step (before the battle is advanced). Lane l writes bytes l, l + 32, l + 64
// of the record (11 + 4 (m + k) <= 83 bytes), so a record is three coalesced byte stores of the warp; a thread per slot
// made ~83 scattered byte stores each (56 us per 65 536 games, a seventh of a self-play step).
__global__ void frames_record_kernel(uint8_t* buf, uint32_t* used, uint16_t* nupd, uint8_t* state, uint32_t cap, uint32_t max_updates,
                                     const uint8_t* __restrict__ results, const uint8_t* __restrict__ c1,
                                     const uint8_t* __restrict__ c2, const uint8_t* __restrict__ counts,
                                     const uint32_t* __restrict__ iterations, const float* __restrict__ emp_value,
                                     const float* __restrict__ nash_value, const float* __restrict__ emp,
                                     const float* __restrict__ nash, uint32_t n) {
  const uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (g >= n) return;
  // every input is asked for before the first decision: the kernel is a chain of dependent loads per warp, and a load
  // behind a branch waits for the one the branch tests
  const float* e = emp + (size_t)g * 18;
  const float* q = nash + (size_t)g * 18;
  const uint32_t st = state[g], res = results[g], m = counts[2 * g], k = counts[2 * g + 1], at = used[g], nu = nupd[g];
  const uint32_t it = iterations ? iterations[g] : 0;
  const uint32_t ca = c1[g], cb = c2[g];
  const float fev = emp_value[g], fnv = nash_value[g];
  const float xe = lane < 18 ? e[lane] : 0.0f, xq = lane < 18 ? q[lane] : 0.0f;  // lane j holds entry j of both tables
  if (st != FS_RECORDING || (res & 15u) != 0) return;  // uniform over the warp
  const uint32_t sz = 11 + 4 * (m + k);
  if (m < 1 || k < 1 || m > 9 || k > 9 || nu >= max_updates || at + sz > cap) {
    __syncwarp();
    if (lane == 0) state[g] = FS_OVERFLOW;
    return;
  }
  uint8_t* p = buf + (size_t)g * cap + at;
  const uint32_t ce = compress_prob(xe), cq = compress_prob(xq);  // 16-bit forms of this lane's two entries
#pragma unroll
  for (uint32_t r = 0; r < 3; r++) {
    const uint32_t b = lane + 32 * r;
    // 16-bit entry j of the record: m empirical p1, m nash p1, k empirical p2, k nash p2 -> (table, index) -> owning lane
    const uint32_t j = b >= 11 ? (b - 11) >> 1 : 0;
    const bool from_q = (j >= m && j < 2 * m) || j >= 2 * m + k;
    const uint32_t src = j < m ? j : j < 2 * m ? j - m : j < 2 * m + k ? 9 + j - 2 * m : 9 + j - 2 * m - k;
    const uint32_t ve = __shfl_sync(0xFFFFFFFFu, ce, (int)(src & 31u)), vq = __shfl_sync(0xFFFFFFFFu, cq, (int)(src & 31u));
    if (b >= sz) continue;
    uint32_t v;
    if (b < 11) {
      const uint32_t ev = compress_prob(fev), nv = compress_prob(fnv);
      v = b == 0 ? ((m - 1) | ((k - 1) << 4)) : b == 1 ? ca : b == 2 ? cb : b < 7 ? it >> (8 * (b - 3))
          : b < 9 ? ev >> (8 * (b - 7)) : nv >> (8 * (b - 9));
    } else {
      v = (from_q ? vq : ve) >> (8 * ((b - 11) & 1));
    }
    p[b] = (uint8_t)v;
  }
  __syncwarp();
  if (lane == 0) {
    used[g] = at + sz;
    nupd[g] = (uint16_t)(nupd[g] + 1);
  }
}

// one thread per slot: a recording game whose battle has ended gets its header and result
__global__ void frames_finish_kernel(uint8_t* buf, uint32_t* used, uint16_t* nupd, uint8_t* state, uint32_t cap,
                                     const uint8_t* __restrict__ results, uint32_t* counters, uint32_t n) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n || (results[g] & 15u) == 0) return;
  if (state[g] == FS_OVERFLOW) {
    state[g] = FS_IDLE;
    atomicAdd(counters + 1, 1u);
    return;
  }
  if (state[g] != FS_RECORDING) return;
  uint8_t* p = buf + (size_t)g * cap;
  const uint32_t nb = used[g];
  p[0] = (uint8_t)nb; p[1] = (uint8_t)(nb >> 8); p[2] = (uint8_t)(nb >> 16); p[3] = (uint8_t)(nb >> 24);
  p[4] = (uint8_t)nupd[g]; p[5] = (uint8_t)(nupd[g] >> 8);
  p[390] = results[g];
  state[g] = FS_FINISHED;
  atomicAdd(counters, 1u);
}

// single CTA: exclusive scan of the finished games' sizes, total in offsets[n]
__global__ void frames_scan_kernel(const uint32_t* used, const uint8_t* state, uint64_t* offsets, uint32_t n) {
  __shared__ uint64_t warp_sums[32];
  __shared__ uint64_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < n; base += blockDim.x) {
    const uint32_t g = base + threadIdx.x;
    const uint64_t v = (g < n && state[g] == FS_FINISHED) ? used[g] : 0;
    uint64_t x = v;
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
      if ((int)(threadIdx.x & 31) >= o) x += y;
    }
    if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      uint64_t w = threadIdx.x < (blockDim.x >> 5) ? warp_sums[threadIdx.x] : 0;
      for (int o = 1; o < 32; o <<= 1) {
        const uint64_t y = __shfl_up_sync(0xFFFFFFFFu, w, o);
        if ((int)threadIdx.x >= o) w += y;
      }
      warp_sums[threadIdx.x] = w;  // inclusive over warps
    }
    __syncthreads();
    const uint64_t before = carry + ((threadIdx.x >> 5) ? warp_sums[(threadIdx.x >> 5) - 1] : 0) + (x - v);
    if (g < n) offsets[g] = before;
    __syncthreads();
    if (threadIdx.x == 0) carry += warp_sums[(blockDim.x >> 5) - 1];
    __syncthreads();
  }
  if (threadIdx.x == 0) offsets[n] = carry;
}

// one CTA per slot: copy a finished game to its place in the stream and free the slot
__global__ void frames_copy_kernel(const uint8_t* buf, const uint32_t* used, uint8_t* state, const uint64_t* offsets, uint32_t cap,
                                   uint8_t* out, uint64_t out_cap) {
  const uint32_t g = blockIdx.x;
  if (state[g] != FS_FINISHED) return;
  const uint32_t nb = used[g];
  const uint64_t at = offsets[g];
  if (at + nb > out_cap) return;  // the host checks the total before trusting the stream
  const uint8_t* src = buf + (size_t)g * cap;
  uint8_t* dst = out + at;
  for (uint32_t i = threadIdx.x; i < nb; i += blockDim.x) dst[i] = src[i];
  __syncthreads();
  if (threadIdx.x == 0) state[g] = FS_IDLE;
}

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// one thread per (battle, player): softmax over the legal logits (softmax.h:5-15, fp32), then get_policy's
// temperature / floor / renormalise in double (policy.h:76-97) and sample_pdf (random.h:41-50) with
// u = (mix64(seed ^ K * step ^ player) >> 11) * 2^-53.
__global__ void policy_sample_kernel(const uint8_t* __restrict__ choices, const uint8_t* __restrict__ counts,
                                     const float* __restrict__ logits, const uint64_t* __restrict__ seeds, uint32_t step,
                                     double temp, double floor_p, float* __restrict__ probs, uint8_t* __restrict__ c1,
                                     uint8_t* __restrict__ c2, uint32_t n) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 2 * n) return;
  const uint32_t g = t >> 1, s = t & 1;
  const uint32_t k = counts[2 * g + s];
  const float* lg = logits + (size_t)g * 18 + 9 * s;
  float* pr = probs + (size_t)g * 18 + 9 * s;
  float y[9];
  float sum = 0;
  for (uint32_t i = 0; i < 9; i++) {
    y[i] = i < k ? expf(lg[i]) : 0.0f;
    if (i < k) sum += y[i];
  }
  double pol[9];
  for (uint32_t i = 0; i < 9; i++) {
    y[i] = i < k ? y[i] / sum : 0.0f;
    pr[i] = y[i];
    pol[i] = (double)y[i];
  }
  if (temp != 1.0) {
    double ps = 0;
    for (uint32_t i = 0; i < 9; i++) {
      pol[i] = pow(pol[i], temp);
      ps += pol[i];
    }
    for (uint32_t i = 0; i < 9; i++) pol[i] /= ps;
  }
  double ps = 0;
  for (uint32_t i = 0; i < 9; i++) {
    if (pol[i] < floor_p) pol[i] = 0;
    ps += pol[i];
  }
  uint32_t pick = 0;
  if (ps > 0) {
    double u = (double)(mix64(seeds[g] ^ (0xD1B54A32D192ED03ull * (uint64_t)step) ^ (uint64_t)s) >> 11) * (1.0 / 9007199254740992.0);
    for (uint32_t i = 0; i < 9; i++) {
      u -= pol[i] / ps;
      if (u <= 0) {
        pick = i;
        break;
      }
    }
  }
  if (pick >= k) pick = 0;
  const uint8_t c = k ? choices[(size_t)g * 18 + 9 * s + pick] : 0;
  (s ? c2 : c1)[g] = c;
}

}  // namespace

void oakb200_frames_release(oakb200_ctx* ctx) {
  if (!ctx || !ctx->frames) return;
  oakb200_frames* f = ctx->frames;
  void* bufs[] = {f->buf, f->used, f->nupd, f->state, f->offsets, f->counters, f->pack};
  for (void* b : bufs)
    if (b) cudaFree(b);
  delete f;
  ctx->frames = nullptr;
}

extern "C" {

int oakb200_frames_create(oakb200_ctx* ctx, size_t games, uint32_t max_updates) {
  if (!ctx || games == 0 || games > ctx->n_max || max_updates == 0 || max_updates > 65535)
    return fail(ctx, OAKB200_EINVAL, "frames_create: games must be 1..n_max and max_updates 1..65535");
  CU(ctx, cudaSetDevice(ctx->device));
  oakb200_frames_release(ctx);
  oakb200_frames* f = new oakb200_frames();
  ctx->frames = f;
  f->games = games;
  f->max_updates = max_updates;
  f->cap = (HEADER + max_updates * 83u + 15u) & ~15u;
  // the pack buffer holds every slot at once, so a drain can never fail for lack of room however long the caller waited
  // between drains (a clamp here once made every later drain fail for good: nothing else shrinks the backlog)
  f->pack_cap = games * (size_t)f->cap;
  CU(ctx, cudaMalloc(&f->buf, games * (size_t)f->cap));
  CU(ctx, cudaMalloc(&f->used, games * 4));
  CU(ctx, cudaMalloc(&f->nupd, games * 2));
  CU(ctx, cudaMalloc(&f->state, games));
  CU(ctx, cudaMalloc(&f->offsets, (games + 1) * 8));
  CU(ctx, cudaMalloc(&f->counters, 8));
  CU(ctx, cudaMalloc(&f->pack, f->pack_cap));
  CU(ctx, cudaMemsetAsync(f->state, 0, games, ctx->stream));
  CU(ctx, cudaMemsetAsync(f->counters, 0, 8, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}

int oakb200_frames_begin_device(oakb200_ctx* ctx, size_t n, const void* battles, const void* results) {
  if (!ctx || !ctx->frames || !battles || !results || n > ctx->frames->games) return fail(ctx, OAKB200_EINVAL, "frames_begin: bad argument");
  if (n == 0) return OAKB200_OK;
  oakb200_frames* f = ctx->frames;
  CU(ctx, cudaSetDevice(ctx->device));
  frames_begin_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, ctx->stream>>>(f->buf, f->used, f->nupd, f->state, f->cap,
                                                                                (const uint8_t*)battles, (const uint8_t*)results, (uint32_t)n);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

int oakb200_frames_record_device(oakb200_ctx* ctx, size_t n, const void* results, const void* c1, const void* c2, const void* counts,
                                 const void* iterations, const void* empirical_value, const void* nash_value,
                                 const void* empirical, const void* nash) {
  if (!ctx || !ctx->frames || !results || !c1 || !c2 || !counts || !empirical_value || !nash_value || !empirical || !nash ||
      n > ctx->frames->games)
    return fail(ctx, OAKB200_EINVAL, "frames_record: bad argument");
  if (n == 0) return OAKB200_OK;
  oakb200_frames* f = ctx->frames;
  CU(ctx, cudaSetDevice(ctx->device));
  frames_record_kernel<<<(unsigned)((n * 32 + 127) / 128), 128, 0, ctx->stream>>>(
      f->buf, f->used, f->nupd, f->state, f->cap, f->max_updates, (const uint8_t*)results, (const uint8_t*)c1, (const uint8_t*)c2,
      (const uint8_t*)counts, (const uint32_t*)iterations, (const float*)empirical_value, (const float*)nash_value,
      (const float*)empirical, (const float*)nash, (uint32_t)n);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

int oakb200_frames_finish_device(oakb200_ctx* ctx, size_t n, const void* results) {
  if (!ctx || !ctx->frames || !results || n > ctx->frames->games) return fail(ctx, OAKB200_EINVAL, "frames_finish: bad argument");
  if (n == 0) return OAKB200_OK;
  oakb200_frames* f = ctx->frames;
  CU(ctx, cudaSetDevice(ctx->device));
  frames_finish_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(f->buf, f->used, f->nupd, f->state, f->cap,
                                                                            (const uint8_t*)results, f->counters, (uint32_t)n);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

int oakb200_frames_pending(oakb200_ctx* ctx, uint32_t* finished, uint32_t* dropped) {
  if (!ctx || !ctx->frames) return fail(ctx, OAKB200_EINVAL, "frames_pending: no recorder");
  uint32_t c[2];
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemcpyAsync(c, ctx->frames->counters, 8, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  if (finished) *finished = c[0];
  if (dropped) *dropped = c[1];
  return OAKB200_OK;
}

int oakb200_frames_drain(oakb200_ctx* ctx, uint8_t* out, size_t cap, size_t* bytes, uint32_t* games) {
  if (!ctx || !ctx->frames || !bytes) return fail(ctx, OAKB200_EINVAL, "frames_drain: bad argument");
  oakb200_frames* f = ctx->frames;
  CU(ctx, cudaSetDevice(ctx->device));
  frames_scan_kernel<<<1, 1024, 0, ctx->stream>>>(f->used, f->state, f->offsets, (uint32_t)f->games);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  uint64_t total = 0;
  uint32_t c[2];
  CU(ctx, cudaMemcpyAsync(&total, f->offsets + f->games, 8, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaMemcpyAsync(c, f->counters, 8, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  *bytes = (size_t)total;
  if (games) *games = c[0];
  if (total == 0) return OAKB200_OK;
  if (!out || total > cap || total > f->pack_cap) return fail(ctx, OAKB200_ECAPACITY, "frames_drain: output buffer too small (bytes holds the size needed)");
  frames_copy_kernel<<<(unsigned)f->games, 128, 0, ctx->stream>>>(f->buf, f->used, f->state, f->offsets, f->cap, f->pack, f->pack_cap);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  CU(ctx, cudaMemsetAsync(f->counters, 0, 4, ctx->stream));
  CU(ctx, cudaMemcpyAsync(out, f->pack, total, cudaMemcpyDefault, ctx->stream));  // `out` may be host or device memory
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return OAKB200_OK;
}

int oakb200_policy_sample_device(oakb200_ctx* ctx, size_t n, const void* choices, const void* counts, const void* logits,
                                 const void* seeds, uint32_t step_index, double temperature, double min_prob, void* probs_out,
                                 void* c1_out, void* c2_out) {
  if (!ctx || !choices || !counts || !logits || !seeds || !probs_out || !c1_out || !c2_out)
    return fail(ctx, OAKB200_EINVAL, "policy_sample: null argument");
  if (n == 0) return OAKB200_OK;
  CU(ctx, cudaSetDevice(ctx->device));
  policy_sample_kernel<<<(unsigned)((2 * n + 127) / 128), 128, 0, ctx->stream>>>(
      (const uint8_t*)choices, (const uint8_t*)counts, (const float*)logits, (const uint64_t*)seeds, step_index, temperature, min_prob,
      (float*)probs_out, (uint8_t*)c1_out, (uint8_t*)c2_out, (uint32_t)n);
  ctx->launches++;
  CU(ctx, cudaGetLastError());
  return OAKB200_OK;
}

}  // extern "C"

// One turn of the network-policy self-play loop as ONE call (generate.cc:271-333 for every running game at once):
// legal choices, leaf evaluation, RuntimePolicy + sample, frame record, update, finish. The same six launches
// oak_b200/selfplay.py used to make one by one; issuing them from here removes the per-launch host round trips that
// dominated a step at 16 Ki games. `before` is scratch for the results as they were before the update.
extern "C" int oakb200_selfplay_step_device(oakb200_ctx* ctx, size_t n, void* battles, void* durations, void* results, const void* seeds,
                                            uint32_t step_index, double temperature, double min_prob, void* choices, void* counts,
                                            void* value, void* logits, void* eval_counts, void* probs, void* c1, void* c2, void* before) {
  if (!ctx || !battles || !durations || !results || !seeds || !choices || !counts || !value || !logits || !eval_counts || !probs || !c1 ||
      !c2 || !before)
    return fail(ctx, OAKB200_EINVAL, "selfplay_step: null argument");
  int rc = oakb200_choices_device(ctx, n, battles, results, choices, counts);
  if (rc) return rc;
  rc = oakb200_eval_device(ctx, n, battles, durations, results, value, logits, eval_counts);
  if (rc) return rc;
  rc = oakb200_policy_sample_device(ctx, n, choices, counts, logits, seeds, step_index, temperature, min_prob, probs, c1, c2);
  if (rc) return rc;
  CU(ctx, cudaMemcpyAsync(before, results, n, cudaMemcpyDeviceToDevice, ctx->stream));
  rc = oakb200_frames_record_device(ctx, n, before, c1, c2, counts, nullptr, value, value, probs, probs);
  if (rc) return rc;
  rc = oakb200_step_device(ctx, n, battles, durations, results, c1, c2, nullptr, nullptr);
  if (rc) return rc;
  return oakb200_frames_finish_device(ctx, n, results);
}
