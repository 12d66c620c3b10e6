// sm_100a kernels for batched Generation I battles. One battle per thread; each CTA stages its
// battles as columns of a word-transposed shared-memory tile (gen1_state.cuh) and runs the engine
// (gen1_engine.cuh) on them. Global memory holds the reference's AoS records.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "gen1_engine.cuh"
#include "train_rows_encode.cuh"

namespace oakb200 {

static const int BATTLE_WORDS = 96;

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// pick word of playout step `step`: the shape of mcts.h:456-461 with a counter-based stream
__device__ __forceinline__ uint64_t pick_word(uint64_t seed, uint32_t step) {
  return splitmix64(seed ^ (0xD1B54A32D192ED03ull * (uint64_t)step));
}
// Remainders by a small divisor (1..9: the number of legal choices) without a division instruction sequence: the
// compiler's run-time `%` is ~25 instructions for 32 bits and a ~55-instruction subroutine for 64 bits, and the playout
// pick needs four of them per update (3.6 % of the playout kernel's instructions, profiles/r2p_rollout_hot_code.txt).
// q = umulhi(x, floor(2^32 / d)) is floor(x / d) or one less, so one conditional subtraction finishes it (d = 1: M = 2^32 - 1).
__constant__ uint32_t SMALL_RECIP[10] = {0u, 0xFFFFFFFFu, 0x80000000u, 0x55555555u, 0x40000000u, 0x33333333u,
                                         0x2AAAAAAAu, 0x24924924u, 0x20000000u, 0x1C71C71Cu};
__constant__ uint32_t SMALL_TWO32[10] = {0u, 0u, 0u, 1u, 0u, 1u, 4u, 4u, 0u, 4u};  // 2^32 mod d
__device__ __forceinline__ uint32_t rem_small(uint32_t x, uint32_t d) {
  const uint32_t r = x - __umulhi(x, SMALL_RECIP[d]) * d;
  return r >= d ? r - d : r;
}
// 64-bit value mod a small divisor (1..9)
__device__ __forceinline__ uint32_t mod_small(uint64_t v, uint32_t d) {
  const uint32_t hi = rem_small((uint32_t)(v >> 32), d), lo = rem_small((uint32_t)v, d);
  return rem_small(hi * SMALL_TWO32[d] + lo, d);
}

// AoS record (384 B, 16-byte aligned) -> this thread's column. A 32-byte-aligned record (every record of a 32-byte-aligned
// array: 384 = 12 x 32) is read with twelve 256-bit loads, one whole 32-byte sector each: with 128-bit loads at a 384-byte
// lane stride the two halves of a sector are asked for by two instructions and the second mostly misses L1 again
// (ncu on step_kernel: 871 MB of sectors from L2 for 420 MB of records).
template <int STRIDE>
__device__ __forceinline__ void load_column(uint32_t* col, const uint8_t* record) {
  if ((reinterpret_cast<uintptr_t>(record) & 31u) == 0) {
#pragma unroll 4
    for (int k = 0; k < BATTLE_WORDS / 8; k++) {
      uint32_t v[8];
      asm volatile("ld.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                   : "l"(record + 32 * k)
                   : "memory");
#pragma unroll
      for (int j = 0; j < 8; j++) col[(8 * k + j) * STRIDE] = v[j];
    }
    return;
  }
  const uint4* src = reinterpret_cast<const uint4*>(record);
#pragma unroll 6
  for (int k = 0; k < BATTLE_WORDS / 4; k++) {
    const uint4 v = src[k];
    col[(4 * k + 0) * STRIDE] = v.x;
    col[(4 * k + 1) * STRIDE] = v.y;
    col[(4 * k + 2) * STRIDE] = v.z;
    col[(4 * k + 3) * STRIDE] = v.w;
  }
}
// The same through L2 only (ld.global.cg): for records that a copy engine may have written while this kernel was already
// running (oakb200_rollout_host). The read-only path (ld.global.nc) is defined only for data that does not change during
// the kernel; L2 is the point of coherence with the copy engines.
template <int STRIDE>
__device__ __forceinline__ void load_column_cg(uint32_t* col, const uint8_t* record) {
  const uint4* src = reinterpret_cast<const uint4*>(record);
#pragma unroll 6
  for (int k = 0; k < BATTLE_WORDS / 4; k++) {
    const uint4 v = __ldcg(src + k);
    col[(4 * k + 0) * STRIDE] = v.x;
    col[(4 * k + 1) * STRIDE] = v.y;
    col[(4 * k + 2) * STRIDE] = v.z;
    col[(4 * k + 3) * STRIDE] = v.w;
  }
}
template <int STRIDE>
__device__ __forceinline__ void store_column(const uint32_t* col, uint8_t* record) {
  if ((reinterpret_cast<uintptr_t>(record) & 31u) == 0) {  // whole sectors, as load_column
#pragma unroll 4
    for (int k = 0; k < BATTLE_WORDS / 8; k++) {
      uint32_t v[8];
#pragma unroll
      for (int j = 0; j < 8; j++) v[j] = col[(8 * k + j) * STRIDE];
      asm volatile("st.global.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(record + 32 * k), "r"(v[0]), "r"(v[1]), "r"(v[2]),
                   "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                   : "memory");
    }
    return;
  }
  uint4* dst = reinterpret_cast<uint4*>(record);
#pragma unroll 6
  for (int k = 0; k < BATTLE_WORDS / 4; k++) {
    uint4 v;
    v.x = col[(4 * k + 0) * STRIDE];
    v.y = col[(4 * k + 1) * STRIDE];
    v.z = col[(4 * k + 2) * STRIDE];
    v.w = col[(4 * k + 3) * STRIDE];
    dst[k] = v;
  }
}

// split_fill (gen1_state.cuh) with vector loads through L2: per party Pokémon the 16 bytes it needs as two 64-bit loads
// (the 24-byte records are 8-byte aligned), active block + order as two 128-bit and one 64-bit load per side, the tail as
// one 128-bit load: 31 loads instead of 72. A lane refills alone while the other lanes of its warp wait.
template <int STRIDE>
__device__ __forceinline__ void split_fill_vec(uint32_t* col, const uint8_t* rec) {
#pragma unroll
  for (int p = 0; p < 2; p++) {
#pragma unroll
    for (int idx = 0; idx < 6; idx++) {
      const uint8_t* r = rec + p * SIDE_SIZE + idx * PK_SIZE;
      const uint2 a = __ldcg(reinterpret_cast<const uint2*>(r + 8)), b = __ldcg(reinterpret_cast<const uint2*>(r + 16));
      // a.x a.y b.x b.y = record words 2..5: spc | id0 pp0 ; id1 pp1 id2 pp2 ; id3 pp3 hp ; status species types level
      uint32_t* c = col + (p * 22 + idx * 2) * STRIDE;
      c[0] = (a.x >> 24) | ((a.y >> 8) & 0xFFu) << 8 | (a.y >> 24) << 16 | ((b.x >> 8) & 0xFFu) << 24;
      c[STRIDE] = (b.x >> 16) | (b.y & 0xFFu) << 16;
    }
    const uint8_t* a = rec + p * SIDE_SIZE + S_ACTIVE;  // 144 / 328: 8-byte aligned, 40 bytes
    uint32_t* c = col + (p * 22 + 12) * STRIDE;
#pragma unroll
    for (int k = 0; k < 5; k++) {
      const uint2 v = __ldcg(reinterpret_cast<const uint2*>(a + 8 * k));
      c[(2 * k) * STRIDE] = v.x;
      c[(2 * k + 1) * STRIDE] = v.y;
    }
  }
  const uint4 t = __ldcg(reinterpret_cast<const uint4*>(rec + B_TURN));
  col[44 * STRIDE] = t.x;
  col[45 * STRIDE] = t.y;
  col[46 * STRIDE] = t.z;
  col[47 * STRIDE] = t.w;
}

struct CgLoad {  // record words through L2 (see load_column_cg)
  static __device__ __forceinline__ uint32_t ld(const uint32_t* p) { return __ldcg(p); }
};

// Both players' legal choices for `result`'s request kinds, then the playout pick (mcts.h:467-477)
template <class E>
__device__ __forceinline__ void pick_choices(const E& e, uint32_t result, uint64_t pick, uint32_t& c1, uint32_t& c2) {
  int m, n;
  const uint64_t l1 = e.choices(0, (result >> 4) & 3, &m);
  const uint64_t l2 = e.choices(1, (result >> 6) & 3, &n);
  c1 = E::choice_at(l1, (int)mod_small(pick, (uint32_t)m));
  c2 = E::choice_at(l2, (int)rem_small((uint32_t)(pick >> 32), (uint32_t)n));
}

// ---------------------------------------------------------------------------------------------
// choices_kernel: pkmn_gen1_battle_choices for both players of every battle
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK) choices_kernel(const uint8_t* __restrict__ battles, const uint8_t* __restrict__ results,
                                                         uint8_t* __restrict__ out, uint8_t* __restrict__ counts, uint32_t n) {
  extern __shared__ uint32_t tile[];
  const uint32_t i = blockIdx.x * BLOCK + threadIdx.x;
  if (i >= n) return;
  uint32_t* col = tile + threadIdx.x;
  load_column<BLOCK>(col, battles + (size_t)i * 384);
  Engine<BLOCK, false, false> e;
  e.begin(reinterpret_cast<uint8_t*>(col), 0, 0);
  const uint32_t r = results[i];
#pragma unroll
  for (int p = 0; p < 2; p++) {
    int k;
    const uint64_t list = e.choices(p, (r >> (4 + 2 * p)) & 3, &k);
    for (int j = 0; j < 9; j++) out[(size_t)i * 18 + p * 9 + j] = j < k ? (uint8_t)Engine<BLOCK, false, false>::choice_at(list, j) : 0;
    counts[(size_t)i * 2 + p] = (uint8_t)k;
  }
}

// ---------------------------------------------------------------------------------------------
// step_kernel: one update per battle, in place.
//   RANDOM = false: choices from c1[]/c2[];  RANDOM = true: playout pick from seeds[] and step_index
//   CALC: per-battle 16-byte calc overrides
template <int BLOCK, bool RANDOM, bool CALC>
__global__ void __launch_bounds__(BLOCK, 3)
step_kernel(uint8_t* __restrict__ battles, uint8_t* __restrict__ durations, uint8_t* __restrict__ results,
            const uint8_t* __restrict__ c1s, const uint8_t* __restrict__ c2s, const uint64_t* __restrict__ seeds,
            uint32_t step_index, const uint8_t* __restrict__ overrides, uint8_t* __restrict__ actions_out,
            uint8_t* __restrict__ chosen_out, uint32_t n) {
  extern __shared__ uint32_t tile[];
  const uint32_t i = blockIdx.x * BLOCK + threadIdx.x;
  if (i >= n) return;
  uint32_t* col = tile + threadIdx.x;
  load_column<BLOCK>(col, battles + (size_t)i * 384);
  const uint2 d = *reinterpret_cast<const uint2*>(durations + (size_t)i * 8);
  Engine<BLOCK, CALC, true> e;
  e.begin(reinterpret_cast<uint8_t*>(col), d.x, d.y);
  const uint32_t prev = results[i];
  uint32_t c1, c2;
  if (RANDOM) {
    pick_choices(e, prev, pick_word(seeds[i], step_index), c1, c2);
    if (chosen_out) {
      chosen_out[(size_t)i * 2] = (uint8_t)c1;
      chosen_out[(size_t)i * 2 + 1] = (uint8_t)c2;
    }
  } else {
    c1 = c1s[i];
    c2 = c2s[i];
  }
  if (CALC) {
    const uint4 o = *reinterpret_cast<const uint4*>(overrides + (size_t)i * 16);
    e.ovr[0] = (uint64_t)o.x | ((uint64_t)o.y << 32);
    e.ovr[1] = (uint64_t)o.z | ((uint64_t)o.w << 32);
  }
  uint32_t r = prev;
  if ((prev & 15u) == 0) {  // a finished battle is frozen
    r = e.update(c1, c2);
    e.end();
    store_column<BLOCK>(col, battles + (size_t)i * 384);
    *reinterpret_cast<uint2*>(durations + (size_t)i * 8) = make_uint2(e.dur[0], e.dur[1]);
    results[i] = (uint8_t)r;
  }
  if (actions_out) {
    uint4 a;
    a.x = (uint32_t)e.act[0];
    a.y = (uint32_t)(e.act[0] >> 32);
    a.z = (uint32_t)e.act[1];
    a.w = (uint32_t)(e.act[1] >> 32);
    *reinterpret_cast<uint4*>(actions_out + (size_t)i * 16) = a;
  }
}

// step_log_kernel: step_kernel with libpkmn's binary protocol log (the -Dlog build; common/protocol.zig:313-863,
// call sites = every `options.log.*` of mechanics.zig). Battle i's messages for this update go to logs[i * log_stride ..],
// their byte count to log_lens[i]; a log that does not fit in log_stride bytes ends at the last message part that did and
// the battle's result byte becomes 0b1100 (bindings/c.zig:6,181). A separate instantiation of the engine: the kernels
// above carry no trace of it.
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK)
step_log_kernel(uint8_t* __restrict__ battles, uint8_t* __restrict__ durations, uint8_t* __restrict__ results,
                const uint8_t* __restrict__ c1s, const uint8_t* __restrict__ c2s, const uint8_t* __restrict__ overrides,
                uint8_t* __restrict__ actions_out, uint8_t* __restrict__ logs, uint32_t log_stride, uint32_t* __restrict__ log_lens,
                uint32_t n) {
  extern __shared__ uint32_t tile[];
  const uint32_t i = blockIdx.x * BLOCK + threadIdx.x;
  if (i >= n) return;
  uint32_t* col = tile + threadIdx.x;
  load_column<BLOCK>(col, battles + (size_t)i * 384);
  const uint2 d = *reinterpret_cast<const uint2*>(durations + (size_t)i * 8);
  Engine<BLOCK, true, true, true> e;
  e.begin(reinterpret_cast<uint8_t*>(col), d.x, d.y);
  e.sink.buf = logs + (size_t)i * log_stride;
  e.sink.cap = log_stride;
  e.sink.pos = 0;
  e.sink.overflow = 0;
  if (overrides) {
    const uint4 o = *reinterpret_cast<const uint4*>(overrides + (size_t)i * 16);
    e.ovr[0] = (uint64_t)o.x | ((uint64_t)o.y << 32);
    e.ovr[1] = (uint64_t)o.z | ((uint64_t)o.w << 32);
  }
  const uint32_t prev = results[i];
  if ((prev & 15u) == 0) {  // a finished battle is frozen
    uint32_t r = e.update(c1s[i], c2s[i]);
    if (e.sink.overflow) r = RESULT_LOG_OVERFLOW;
    e.end();
    store_column<BLOCK>(col, battles + (size_t)i * 384);
    *reinterpret_cast<uint2*>(durations + (size_t)i * 8) = make_uint2(e.dur[0], e.dur[1]);
    results[i] = (uint8_t)r;
  }
  log_lens[i] = e.sink.pos;
  if (actions_out) {
    uint4 a;
    a.x = (uint32_t)e.act[0];
    a.y = (uint32_t)(e.act[0] >> 32);
    a.z = (uint32_t)e.act[1];
    a.w = (uint32_t)(e.act[1] >> 32);
    *reinterpret_cast<uint4*>(actions_out + (size_t)i * 16) = a;
  }
}

// Streaming input (oakb200_rollout_host): the batch is still arriving over PCIe while the kernel runs. The copy stream
// raises *available (the number of leading records that have landed) after each chunk; a lane whose ticket is beyond it
// waits. Tickets and chunks both advance in index order and every copy was queued before the launch, so the wait ends.
// available[1] is an error word: a lane that has polled for about a minute (a failed or blocked copy) raises it and gives
// up instead of holding the SMs forever; the host turns it into OAKB200_ECUDA.
__device__ __forceinline__ bool wait_available(uint32_t* available, uint32_t i) {
  if (!available) return true;
  uint32_t polls = 0;
  while (i >= *reinterpret_cast<const volatile uint32_t*>(available)) {
    __nanosleep(500);
    if (++polls > (1u << 27)) {
      atomicExch(available + 1, 1u);
      return false;
    }
  }
  __threadfence();
  return true;
}

// ---------------------------------------------------------------------------------------------
// rollout_kernel: random playouts to a terminal result. Persistent CTAs; every lane pulls its next
// battle from a global ticket counter the moment its current one ends (lane refill), so a warp
// stays full although playout lengths differ by an order of magnitude. The loop body is
// "maybe refill, then exactly one update" so that all lanes re-converge once per update.
// SPLIT: the on-chip column holds only the 48 words an update can write (gen1_state.cuh); stored stats, move ids, species,
// types and level are read from the battle's record in global memory. 24 KB per CTA instead of 48: 8 CTAs = 32 warps per SM.
#ifndef OAKB200_ROLLOUT_MINB
#define OAKB200_ROLLOUT_MINB 8  // resident CTAs per SM the register budget is set for (A/B knob: 7 -> 72 registers, 6 -> 80)
#endif
template <int BLOCK, bool CHANCE, bool WRITE_STATE, bool SPLIT>
__global__ void __launch_bounds__(BLOCK, SPLIT ? OAKB200_ROLLOUT_MINB : 3)
rollout_kernel(uint8_t* battles, uint8_t* __restrict__ durations, uint8_t* __restrict__ results,
               const uint64_t* __restrict__ seeds, uint32_t max_steps, uint32_t n, uint32_t* __restrict__ ticket,
               uint8_t* __restrict__ score2_out, uint16_t* __restrict__ turns_out, uint32_t* __restrict__ steps_out,
               unsigned long long* __restrict__ total_steps, uint32_t* available) {
  extern __shared__ uint32_t tile[];
  uint32_t* col = tile + threadIdx.x;
  Engine<BLOCK, false, CHANCE, false, SPLIT> e;
  e.begin(reinterpret_cast<uint8_t*>(col), 0, 0);

  uint32_t i = 0, result = 0, step = 0;
  uint64_t seed = 0;
  bool have = false, exhausted = false;
  uint32_t my_steps = 0;

  for (;;) {
    if (!have && !exhausted) {
      i = atomicAdd(ticket, 1u);
      if (i < n && wait_available(available, i)) {
        if (SPLIT) {
          split_fill_vec<BLOCK>(col, battles + (size_t)i * 384);
        } else {
          load_column_cg<BLOCK>(col, battles + (size_t)i * 384);
        }
        const uint2 d = __ldcg(reinterpret_cast<const uint2*>(durations + (size_t)i * 8));
        e.begin(reinterpret_cast<uint8_t*>(col), d.x, d.y);
        e.m.set_const(battles + (size_t)i * 384);
        result = __ldcg(results + i);
        seed = __ldcg(seeds + i);
        step = 0;
        have = true;
      } else {
        exhausted = true;
      }
    }
    if (__all_sync(0xFFFFFFFFu, !have)) break;
    if (have) {
      const bool live = (result & 15u) == 0 && step < max_steps;
      if (live) {
        uint32_t c1, c2;
        pick_choices(e, result, pick_word(seed, step), c1, c2);
        e.act[0] = e.act[1] = 0;  // pkmn_gen1_battle_options_set(null, null, null) (mcts.h:478)
        result = e.update(c1, c2);
        step++;
      }
      if (!live || (result & 15u) != 0 || step >= max_steps) {
        const uint32_t t = result & 15u;
        if (score2_out) score2_out[i] = t == RES_WIN ? 2 : (t == RES_TIE ? 1 : (t == RES_LOSE ? 0 : 255));
        if (turns_out) turns_out[i] = (uint16_t)e.m.ld16(B_TURN);
        if (steps_out) steps_out[i] = step;
        if (WRITE_STATE) {
          e.end();
          if (SPLIT) {
            split_merge<BLOCK, CgLoad>(col, reinterpret_cast<uint32_t*>(battles + (size_t)i * 384));
          } else {
            store_column<BLOCK>(col, battles + (size_t)i * 384);
          }
          *reinterpret_cast<uint2*>(durations + (size_t)i * 8) = make_uint2(e.dur[0], e.dur[1]);
          results[i] = (uint8_t)result;
        }
        my_steps += step;
        have = false;
      }
    }
  }
  // one atomic per warp
  uint32_t s = my_steps;
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
  if ((threadIdx.x & 31) == 0 && total_steps) atomicAdd(total_steps, (unsigned long long)s);
}

// ---------------------------------------------------------------------------------------------
// rollout_grouped_kernel: the same playouts as rollout_kernel, with the CTA's 128 battles REGROUPED by action
// class before every update. A battle's state is a shared-memory column, not a thread's registers, so any
// thread of the CTA can advance any column: each iteration the owner thread t enumerates column t's legal
// choices and makes the playout pick, the CTA counting-sorts its columns by (class of P1's action, class of
// P2's action) — pass / switch / move by effect group — and thread u then runs the turn transition of the
// u-th column in sorted order. Warps therefore execute mostly one kind of turn (switch-switch, move-move of
// one effect family, ...) instead of the union of all of them, and idle columns collect in whole warps.
// The first profile of rollout_kernel showed 6.2 of 32 lanes active in executeMove and ~50 % of stall samples
// waiting on instruction fetch (profiles/r1b_rollout_kernel_by_function.txt).
__device__ __forceinline__ uint32_t action_class(const Mem<128>& m, int p, uint32_t choice) {
  const uint32_t type = choice & 3u, data = choice >> 2;
  if (type == CH_PASS) return 0;
  if (type == CH_SWITCH) return 1;
  if (data == 0) return 2;  // Struggle / forced continuation
  const uint32_t id = m.ld8(Mem<128>::sb(p), A_MOVES + 2 * ((int)data - 1));
  const uint32_t eff = mv_effect(move_word(id));
  if (eff == 0) return 2;          // plain damage
  if (eff <= 16) return 3;         // onBegin family: status / screens / heal / substitute ...
  if (eff <= 31) return 4;         // stat down, onEnd family
  if (eff <= 47) return 5;         // always-happens / special / multi-hit
  return 6;                        // secondary-chance families
}

template <int BLOCK, bool CHANCE, bool WRITE_STATE>
__global__ void __launch_bounds__(BLOCK, 3)
rollout_grouped_kernel(uint8_t* __restrict__ battles, uint8_t* __restrict__ durations, uint8_t* __restrict__ results,
                       const uint64_t* __restrict__ seeds, uint32_t max_steps, uint32_t n, uint32_t* __restrict__ ticket,
                       uint8_t* __restrict__ score2_out, uint16_t* __restrict__ turns_out, uint32_t* __restrict__ steps_out,
                       unsigned long long* __restrict__ total_steps, uint32_t* available) {
  static_assert(BLOCK == 128, "action_class and the sort assume 128 columns");
  extern __shared__ uint32_t tile[];
  uint32_t* s_meta = tile + BATTLE_WORDS * BLOCK;  // c1 | c2 << 8 | live << 16
  uint32_t* s_res = s_meta + BLOCK;
  uint32_t* s_perm = s_res + BLOCK;
  uint32_t* s_dur = s_perm + BLOCK;                // 2 per column
  uint32_t* s_cnt = s_dur + 2 * BLOCK;             // 64 buckets
  const int t = threadIdx.x;
  uint32_t* col = tile + t;
  Mem<BLOCK> m;
  m.set(reinterpret_cast<uint8_t*>(col));

  uint32_t i = 0, result = 0, step = 0, my_steps = 0;
  uint64_t seed = 0;
  bool have = false, exhausted = false;

  for (;;) {
    // ---- owner: refill, legal choices, playout pick, sort key
    if (!have && !exhausted) {
      i = atomicAdd(ticket, 1u);
      if (i < n && wait_available(available, i)) {
        load_column_cg<BLOCK>(col, battles + (size_t)i * 384);
        const uint2 d = __ldcg(reinterpret_cast<const uint2*>(durations + (size_t)i * 8));
        s_dur[2 * t] = d.x;
        s_dur[2 * t + 1] = d.y;
        result = __ldcg(results + i);
        seed = __ldcg(seeds + i);
        step = 0;
        have = true;
      } else {
        exhausted = true;
      }
    }
    if (t < 64) s_cnt[t] = 0;
    const bool live = have && (result & 15u) == 0 && step < max_steps;
    uint32_t key = 63;
    if (live) {
      const uint64_t pick = pick_word(seed, step);
      const uint64_t l1 = mem_choices(m, 0, (result >> 4) & 3);
      const uint64_t l2 = mem_choices(m, 1, (result >> 6) & 3);
      const uint32_t c1 = choice_at(l1, (int)mod_small(pick, (uint32_t)choice_count(l1)));
      const uint32_t c2 = choice_at(l2, (int)rem_small((uint32_t)(pick >> 32), (uint32_t)choice_count(l2)));
      s_meta[t] = c1 | (c2 << 8) | (1u << 16);
      key = action_class(m, 0, c1) * 8 + action_class(m, 1, c2);
    } else {
      s_meta[t] = 0;
    }
    if (__syncthreads_count(have) == 0) break;
    // ---- counting sort of the columns by key
    const uint32_t rank = atomicAdd(&s_cnt[key], 1u);
    __syncthreads();
    if (t < 32) {
      const uint32_t a = s_cnt[2 * t], b = s_cnt[2 * t + 1];
      uint32_t x = a + b;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
        if (t >= o) x += y;
      }
      s_cnt[2 * t] = x - a - b;
      s_cnt[2 * t + 1] = x - b;
    }
    __syncthreads();
    s_perm[s_cnt[key] + rank] = (uint32_t)t;
    __syncthreads();
    // ---- worker: the turn transition of the t-th column in class order
    {
      const uint32_t p = s_perm[t];
      const uint32_t meta = s_meta[p];
      if (meta >> 16) {
        Engine<BLOCK, false, CHANCE> e;
        e.begin(reinterpret_cast<uint8_t*>(tile + p), s_dur[2 * p], s_dur[2 * p + 1]);
        const uint32_t r = e.update(meta & 255u, (meta >> 8) & 255u);  // begin() cleared the actions (mcts.h:478)
        e.end();
        if (CHANCE) {
          s_dur[2 * p] = e.dur[0];
          s_dur[2 * p + 1] = e.dur[1];
        }
        s_res[p] = r;
      }
    }
    __syncthreads();
    // ---- owner: bookkeeping, outputs of a finished playout
    if (live) {
      result = s_res[t];
      step++;
    }
    if (have && (!live || (result & 15u) != 0 || step >= max_steps)) {
      const uint32_t ty = result & 15u;
      if (score2_out) score2_out[i] = ty == RES_WIN ? 2 : (ty == RES_TIE ? 1 : (ty == RES_LOSE ? 0 : 255));
      if (turns_out) turns_out[i] = (uint16_t)m.ld16(B_TURN);
      if (steps_out) steps_out[i] = step;
      if (WRITE_STATE) {
        store_column<BLOCK>(col, battles + (size_t)i * 384);
        *reinterpret_cast<uint2*>(durations + (size_t)i * 8) = make_uint2(s_dur[2 * t], s_dur[2 * t + 1]);
        results[i] = (uint8_t)result;
      }
      my_steps += step;
      have = false;
    }
  }
  uint32_t s = my_steps;
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
  if ((threadIdx.x & 31) == 0 && total_steps) atomicAdd(total_steps, (unsigned long long)s);
}

// ---------------------------------------------------------------------------------------------
// Tree-search support (reference cpp/include/search/mcts.h:232-246,304-361,569-604).
// search_spawn_kernel: particle b of root b / per_root starts from a copy of its root's battle whose rng is replaced by
// a fresh 64-bit value (run_root_iteration: `rng[0] = device.uniform_64()`); randomize_hidden_kernel then redraws the
// hidden counters from the root's observed durations.
__global__ void search_spawn_kernel(const uint8_t* __restrict__ root_battles, const uint8_t* __restrict__ root_durations,
                                    const uint8_t* __restrict__ root_results, const uint64_t* __restrict__ rng_values,
                                    uint32_t per_root, uint8_t* __restrict__ battles, uint8_t* __restrict__ durations,
                                    uint8_t* __restrict__ results, uint32_t n) {
  const uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (g >= n) return;
  const uint32_t r = g / per_root;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(root_battles + (size_t)r * 384);
  uint32_t* dst = reinterpret_cast<uint32_t*>(battles + (size_t)g * 384);
  for (int w = lane; w < BATTLE_WORDS; w += 32) {
    uint32_t v = src[w];
    if (w == B_RNG / 4) v = (uint32_t)rng_values[g];
    if (w == B_RNG / 4 + 1) v = (uint32_t)(rng_values[g] >> 32);
    dst[w] = v;
  }
  if (lane < 2) reinterpret_cast<uint32_t*>(durations + (size_t)g * 8)[lane] = reinterpret_cast<const uint32_t*>(root_durations + (size_t)r * 8)[lane];
  if (lane == 2) results[g] = root_results[r];
}

// search_step_kernel: one tree step of every active particle. The host's bandits choose INDICES into the legal-choice
// lists (mcts.h:334-342); the kernel enumerates the lists, applies Search::battle_options_set's clamped damage rolls
// (rolls[b] in {1, 2, 3, 20}: both players' damage overrides from bytes 6 and 7 of the battle rng, roll_byte
// mcts.h:593-604; 39 = no clamping), updates with chance tracking, and returns the result, the 16-byte observation that
// keys the child node (mcts.h:355-358) and the legal-choice counts of the new state (the child's bandit sizes).
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK, 3)
search_step_kernel(uint8_t* __restrict__ battles, uint8_t* __restrict__ durations, uint8_t* __restrict__ results,
                   const uint8_t* __restrict__ idx1, const uint8_t* __restrict__ idx2, const uint8_t* __restrict__ rolls,
                   uint8_t* __restrict__ actions_out, uint8_t* __restrict__ counts_out, uint32_t n) {
  extern __shared__ uint32_t tile[];
  const uint32_t i = blockIdx.x * BLOCK + threadIdx.x;
  if (i >= n) return;
  const uint32_t a = idx1[i], b = idx2[i], prev = results[i];
  if (a == 255u || (prev & 15u) != 0) return;
  uint32_t* col = tile + threadIdx.x;
  load_column<BLOCK>(col, battles + (size_t)i * 384);
  const uint2 d = *reinterpret_cast<const uint2*>(durations + (size_t)i * 8);
  Engine<BLOCK, true, true> e;
  e.begin(reinterpret_cast<uint8_t*>(col), d.x, d.y);
  int m, k;
  const uint64_t l1 = e.choices(0, (prev >> 4) & 3, &m);
  const uint64_t l2 = e.choices(1, (prev >> 6) & 3, &k);
  const uint32_t c1 = Engine<BLOCK, true, true>::choice_at(l1, a < (uint32_t)m ? (int)a : 0);
  const uint32_t c2 = Engine<BLOCK, true, true>::choice_at(l2, b < (uint32_t)k ? (int)b : 0);
  const uint32_t nr = rolls[i];
  if (nr != 39u) {
    const uint32_t r0 = (uint32_t)(e.rng >> 48) & 255u, r1 = (uint32_t)(e.rng >> 56);
    const uint32_t step = nr > 1 ? 38u / (nr - 1) : 0u;
    e.ovr[0] = nr == 1 ? 236u : 217u + step * (r0 % nr);
    e.ovr[1] = nr == 1 ? 236u : 217u + step * (r1 % nr);
  }
  const uint32_t r = e.update(c1, c2);
  e.end();
  store_column<BLOCK>(col, battles + (size_t)i * 384);
  *reinterpret_cast<uint2*>(durations + (size_t)i * 8) = make_uint2(e.dur[0], e.dur[1]);
  results[i] = (uint8_t)r;
  uint4 act;
  act.x = (uint32_t)e.act[0];
  act.y = (uint32_t)(e.act[0] >> 32);
  act.z = (uint32_t)e.act[1];
  act.w = (uint32_t)(e.act[1] >> 32);
  *reinterpret_cast<uint4*>(actions_out + (size_t)i * 16) = act;
  int m2 = 0, k2 = 0;
  if ((r & 15u) == 0) {
    e.choices(0, (r >> 4) & 3, &m2);
    e.choices(1, (r >> 6) & 3, &k2);
  }
  counts_out[(size_t)i * 2] = (uint8_t)m2;
  counts_out[(size_t)i * 2 + 1] = (uint8_t)k2;
}

// ---------------------------------------------------------------------------------------------
// Training rows, step 1 of 2 (reference cpp/include/py/battle/encoded-frames.h:112-133, uncompress_from_bytes): replay
// every stored game of a .battle.data stream from its first battle with its stored choices. Before each update the
// state the learner's row is built from is staged as one ROW_STATE_BYTES record:
//   [0,384) battle   [384,392) durations   392 result before the update   393 the game's stored final result
//   394, 395 legal-choice counts   [396,400) byte offset of the stored update in the stream
//   [400,409) P1's legal choices   [409,418) P2's   (pkmn_gen1_battle_choices order)
// rows_encode_kernel (train_rows.cu) turns these into the dense tensors. Persistent CTAs with lane refill from a ticket
// counter, as rollout_kernel: stored games differ in length by an order of magnitude.
// only_update: null = every update of game g goes to rows row_base[g] + u (from_bytes); else game g stages only update
// only_update[g] into row row_base[g] and stops there (pyoak `sample`, cpp/src/pyoak.cc:111-260).
static const int ROW_STATE_BYTES = 432;

template <int BLOCK>
__global__ void __launch_bounds__(BLOCK, 3)
replay_kernel(const uint8_t* __restrict__ stream, const uint64_t* __restrict__ game_off, const uint32_t* __restrict__ row_base,
              const int32_t* __restrict__ only_update, uint32_t n_games, uint32_t* __restrict__ ticket, uint8_t* __restrict__ rows,
              uint32_t* __restrict__ first_bad) {
  extern __shared__ uint32_t tile[];
  uint32_t* col = tile + threadIdx.x;
  typedef Engine<BLOCK, false, true> E;
  E e;
  e.begin(reinterpret_cast<uint8_t*>(col), 0, 0);

  uint32_t g = 0, u = 0, n_updates = 0, result = 0, stored = 0, base = 0;
  int32_t only = -1;
  uint64_t p = 0, end = 0;
  bool have = false, exhausted = false;

  for (;;) {
    if (!have && !exhausted) {
      g = atomicAdd(ticket, 1u);
      if (g < n_games) {
        const uint8_t* game = stream + game_off[g];  // games are packed back to back: no alignment to rely on
        const uint32_t n_bytes = game[0] | (game[1] << 8) | (game[2] << 16) | ((uint32_t)game[3] << 24);
        n_updates = game[4] | (game[5] << 8);
#pragma unroll 4
        for (int w = 0; w < BATTLE_WORDS; w++) {
          const uint8_t* b = game + 6 + 4 * w;
          col[w * BLOCK] = b[0] | (b[1] << 8) | (b[2] << 16) | ((uint32_t)b[3] << 24);
        }
        e.begin(reinterpret_cast<uint8_t*>(col), 0, 0);
        stored = game[390];
        result = RESULT_DEFAULT;  // PKMN::result(): both players to move (cpp/include/libpkmn/pkmn.h:228-233)
        p = game_off[g] + 391;
        end = game_off[g] + n_bytes;
        base = row_base[g];
        only = only_update ? max(only_update[g], 0) : -1;
        if (only >= 0 && n_updates != 0) {  // clamp to the game's last stored update
          if ((uint32_t)only >= n_updates) only = (int32_t)n_updates - 1;
          n_updates = (uint32_t)only + 1;
        }
        u = 0;
        have = n_updates != 0;
      } else {
        exhausted = true;
      }
    }
    if (__all_sync(0xFFFFFFFFu, !have)) break;
    // a stored update must lie inside its game's bytes (a corrupt header must not walk the reads out of the stream)
    if (have && (p + 11 > end || p + 11 + 4 * ((stream[p] & 15u) + (stream[p] >> 4) + 2) > end)) {
      atomicMin(first_bad, g);
      have = false;
    }
    if (have) {
      if (only < 0 || u == (uint32_t)only) {
        e.end();
        uint8_t* row = rows + (size_t)(only < 0 ? base + u : base) * ROW_STATE_BYTES;
        store_column<BLOCK>(col, row);
        int m, n;
        const uint64_t l1 = e.choices(0, (result >> 4) & 3, &m);
        const uint64_t l2 = e.choices(1, (result >> 6) & 3, &n);
        uint32_t cw[5] = {0, 0, 0, 0, 0};
        for (int i = 0; i < 9; i++) {
          if (i < m) cw[i >> 2] |= E::choice_at(l1, i) << (8 * (i & 3));
          if (i < n) cw[(9 + i) >> 2] |= E::choice_at(l2, i) << (8 * ((9 + i) & 3));
        }
        uint4* tail = reinterpret_cast<uint4*>(row + 384);
        tail[0] = make_uint4(e.dur[0], e.dur[1], result | (stored << 8) | ((uint32_t)m << 16) | ((uint32_t)n << 24), (uint32_t)p);
        tail[1] = make_uint4(cw[0], cw[1], cw[2], cw[3]);
        tail[2] = make_uint4(cw[4], 0, 0, 0);
      }
      const uint32_t hdr = stream[p], c1 = stream[p + 1], c2 = stream[p + 2];
      e.act[0] = e.act[1] = 0;  // PKMN::update: options_set(null) before every update (pkmn.h:134-138)
      result = e.update(c1, c2);
      p += 11 + 4 * ((hdr & 15u) + (hdr >> 4) + 2);
      u++;
      if (u == n_updates) {
        // a stored game must end on its stored result and fill its byte count (encoded-frames.h:132 asserts the first)
        if (only < 0 && (result != stored || p != end)) atomicMin(first_bad, g);
        have = false;
      }
    }
  }
}

// replay_encode_kernel: the two steps of the training-row build in ONE launch. replay_kernel is bound by instruction
// issue under divergence, rows_encode_kernel by DRAM writes; back to back they add up (profiles/r1k_training_rows.txt).
// Here every warp alternates: each lane advances its own game by one stored update and stages the state it had before
// (as replay_kernel), then the whole warp encodes the <= 32 rows just staged (rows::encode_row, the body of
// rows_encode_kernel). The staged state goes through global memory — written by its lane, read back by the warp out
// of L2 with ld.global.cg — because a game's state is one bank-private COLUMN of the tile: no other lane can read it
// from shared memory without a 32-way conflict. Warps of an SM are in different phases at any time, so the copy-out
// of some overlaps the simulation of others.
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK, 3)
replay_encode_kernel(const uint8_t* __restrict__ stream, const uint64_t* __restrict__ game_off, const uint32_t* __restrict__ row_base,
                     const int32_t* __restrict__ only_update, uint32_t n_games, uint32_t* __restrict__ ticket, uint8_t* rows,
                     uint32_t* __restrict__ first_bad, oakb200_training_rows out) {
  extern __shared__ uint32_t tile[];
  uint32_t* col = tile + threadIdx.x;
  uint8_t* st = reinterpret_cast<uint8_t*>(tile + BATTLE_WORDS * BLOCK) + (threadIdx.x >> 5) * ROW_STATE_BYTES;
  const int lane = threadIdx.x & 31;
  typedef Engine<BLOCK, false, true> E;
  E e;
  e.begin(reinterpret_cast<uint8_t*>(col), 0, 0);

  uint32_t g = 0, u = 0, n_updates = 0, result = 0, stored = 0, base = 0;
  int32_t only = -1;
  uint64_t p = 0, end = 0;
  bool have = false, exhausted = false;

  for (;;) {
    if (!have && !exhausted) {
      g = atomicAdd(ticket, 1u);
      if (g < n_games) {
        const uint8_t* game = stream + game_off[g];
        const uint32_t n_bytes = game[0] | (game[1] << 8) | (game[2] << 16) | ((uint32_t)game[3] << 24);
        n_updates = game[4] | (game[5] << 8);
#pragma unroll 4
        for (int w = 0; w < BATTLE_WORDS; w++) {
          const uint8_t* b = game + 6 + 4 * w;
          col[w * BLOCK] = b[0] | (b[1] << 8) | (b[2] << 16) | ((uint32_t)b[3] << 24);
        }
        e.begin(reinterpret_cast<uint8_t*>(col), 0, 0);
        stored = game[390];
        result = RESULT_DEFAULT;
        p = game_off[g] + 391;
        end = game_off[g] + n_bytes;
        base = row_base[g];
        only = only_update ? max(only_update[g], 0) : -1;
        if (only >= 0 && n_updates != 0) {
          if ((uint32_t)only >= n_updates) only = (int32_t)n_updates - 1;
          n_updates = (uint32_t)only + 1;
        }
        u = 0;
        have = n_updates != 0;
      } else {
        exhausted = true;
      }
    }
    if (__all_sync(0xFFFFFFFFu, !have)) break;
    bool emit = false;
    uint32_t my_row = 0;
    if (have && (p + 11 > end || p + 11 + 4 * ((stream[p] & 15u) + (stream[p] >> 4) + 2) > end)) {
      atomicMin(first_bad, g);
      have = false;
    }
    if (have) {
      if (only < 0 || u == (uint32_t)only) {
        e.end();
        my_row = only < 0 ? base + u : base;
        emit = true;
        uint8_t* row = rows + (size_t)my_row * ROW_STATE_BYTES;
        store_column<BLOCK>(col, row);
        int m, n;
        const uint64_t l1 = e.choices(0, (result >> 4) & 3, &m);
        const uint64_t l2 = e.choices(1, (result >> 6) & 3, &n);
        uint32_t cw[5] = {0, 0, 0, 0, 0};
        for (int i = 0; i < 9; i++) {
          if (i < m) cw[i >> 2] |= E::choice_at(l1, i) << (8 * (i & 3));
          if (i < n) cw[(9 + i) >> 2] |= E::choice_at(l2, i) << (8 * ((9 + i) & 3));
        }
        uint4* tail = reinterpret_cast<uint4*>(row + 384);
        tail[0] = make_uint4(e.dur[0], e.dur[1], result | (stored << 8) | ((uint32_t)m << 16) | ((uint32_t)n << 24), (uint32_t)p);
        tail[1] = make_uint4(cw[0], cw[1], cw[2], cw[3]);
        tail[2] = make_uint4(cw[4], 0, 0, 0);
      }
      const uint32_t hdr = stream[p], c1 = stream[p + 1], c2 = stream[p + 2];
      e.act[0] = e.act[1] = 0;
      result = e.update(c1, c2);
      p += 11 + 4 * ((hdr & 15u) + (hdr >> 4) + 2);
      u++;
      if (u == n_updates) {
        if (only < 0 && (result != stored || p != end)) atomicMin(first_bad, g);
        have = false;
      }
    }
    // ---- the warp encodes what its lanes staged (the staged bytes are ordered before these reads by the barrier)
    __threadfence_block();
    __syncwarp();
    uint32_t todo = __ballot_sync(0xFFFFFFFFu, emit);
    while (todo) {
      const int j = __ffs(todo) - 1;
      todo &= todo - 1;
      const uint32_t r = __shfl_sync(0xFFFFFFFFu, my_row, j);
      if (lane < ROW_STATE_BYTES / 16)
        reinterpret_cast<uint4*>(st)[lane] = __ldcg(reinterpret_cast<const uint4*>(rows + (size_t)r * ROW_STATE_BYTES) + lane);
      __syncwarp();
      rows::encode_row(st, stream, out, r, lane);
      __syncwarp();
    }
  }
}

// replay_split_kernel: the training-row build as ONE launch of co-resident CTAs with two roles. CTAs [0, n_producers) run
// the replay loop of replay_kernel and, after staging a row, append its index to a global work queue; every other CTA's
// warps pop queue tickets and encode one row each (rows::encode_row) as soon as its entry appears. The two halves have
// complementary limits — the replay is latency-bound on the longest game and leaves the SM's issue slots and all of
// DRAM idle, the encode is a DRAM-write stream — so running them side by side costs about the larger of the two instead
// of their sum. Producers never wait; consumers wait only on producers, and every CTA of the grid is resident
// (cooperative launch), so the wait ends; a consumer that polls for ~a minute gives up and flags the build.
//   queue[t] = 0 (not produced yet), row + 1, or 0xFFFFFFFF (a ticket without a row: a game with no updates in
//   only_update mode); control[0] game ticket, [1] first bad game, [2] queue tail, [3] queue head, [4] consumer gave up
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK, 3)
replay_split_kernel(const uint8_t* __restrict__ stream, const uint64_t* __restrict__ game_off, const uint32_t* __restrict__ row_base,
                    const int32_t* __restrict__ only_update, uint32_t n_games, uint32_t n_rows, uint32_t n_producers,
                    uint32_t* __restrict__ control, uint32_t* queue, uint8_t* rows, oakb200_training_rows out) {
  extern __shared__ uint32_t tile[];
  const int lane = threadIdx.x & 31;
  uint32_t* ticket = control;
  uint32_t* first_bad = control + 1;
  uint32_t* q_tail = control + 2;
  uint32_t* q_head = control + 3;

  if (blockIdx.x >= n_producers) {
    // ===== consumer warps =====
    uint8_t* st = reinterpret_cast<uint8_t*>(tile) + (threadIdx.x >> 5) * ROW_STATE_BYTES;
    for (;;) {
      uint32_t t = 0, token = 0;
      if (lane == 0) {
        t = atomicAdd(q_head, 1u);
        if (t < n_rows) {
          const volatile uint32_t* slot = queue + t;
          uint32_t polls = 0;
          while ((token = *slot) == 0) {
            __nanosleep(200);
            if (++polls > (1u << 28)) {  // never in a correct build; do not hang the device if the producers failed
              atomicExch(control + 4, 1u);
              token = 0xFFFFFFFFu;
              break;
            }
          }
          __threadfence();  // the staged state was written before the queue entry
        }
      }
      t = __shfl_sync(0xFFFFFFFFu, t, 0);
      if (t >= n_rows) break;
      token = __shfl_sync(0xFFFFFFFFu, token, 0);
      if (token == 0xFFFFFFFFu) continue;
      const uint32_t r = token - 1;
      if (lane < ROW_STATE_BYTES / 16)
        reinterpret_cast<uint4*>(st)[lane] = __ldcg(reinterpret_cast<const uint4*>(rows + (size_t)r * ROW_STATE_BYTES) + lane);
      __syncwarp();
      rows::encode_row(st, stream, out, r, lane);
      __syncwarp();
    }
    return;
  }

  // ===== producer threads: replay_kernel's loop + publication =====
  uint32_t* col = tile + threadIdx.x;
  typedef Engine<BLOCK, false, true> E;
  E e;
  e.begin(reinterpret_cast<uint8_t*>(col), 0, 0);
  uint32_t g = 0, u = 0, n_updates = 0, result = 0, stored = 0, base = 0;
  int32_t only = -1;
  uint64_t p = 0, end = 0;
  bool have = false, exhausted = false;
  for (;;) {
    uint32_t publish = 0;  // 0 nothing, row + 1, or 0xFFFFFFFF
    if (!have && !exhausted) {
      g = atomicAdd(ticket, 1u);
      if (g < n_games) {
        const uint8_t* game = stream + game_off[g];
        const uint32_t n_bytes = game[0] | (game[1] << 8) | (game[2] << 16) | ((uint32_t)game[3] << 24);
        n_updates = game[4] | (game[5] << 8);
#pragma unroll 4
        for (int w = 0; w < BATTLE_WORDS; w++) {
          const uint8_t* b = game + 6 + 4 * w;
          col[w * BLOCK] = b[0] | (b[1] << 8) | (b[2] << 16) | ((uint32_t)b[3] << 24);
        }
        e.begin(reinterpret_cast<uint8_t*>(col), 0, 0);
        stored = game[390];
        result = RESULT_DEFAULT;
        p = game_off[g] + 391;
        end = game_off[g] + n_bytes;
        base = row_base[g];
        only = only_update ? max(only_update[g], 0) : -1;
        if (only >= 0 && n_updates != 0) {
          if ((uint32_t)only >= n_updates) only = (int32_t)n_updates - 1;
          n_updates = (uint32_t)only + 1;
        }
        u = 0;
        have = n_updates != 0;
        if (!have && only_update) publish = 0xFFFFFFFFu;  // this game's ticket carries no row
      } else {
        exhausted = true;
      }
    }
    if (__all_sync(0xFFFFFFFFu, !have && publish == 0)) break;
    if (have && (p + 11 > end || p + 11 + 4 * ((stream[p] & 15u) + (stream[p] >> 4) + 2) > end)) {
      atomicMin(first_bad, g);
      // the rows this game still owes are published as empty tickets so that the consumers' count comes out
      const uint32_t owed = only < 0 ? n_updates - u : 1u;
      const uint32_t at = atomicAdd(q_tail, owed);
      for (uint32_t k = 0; k < owed; k++) queue[at + k] = 0xFFFFFFFFu;
      have = false;
    }
    if (have) {
      if (only < 0 || u == (uint32_t)only) {
        e.end();
        const uint32_t my_row = only < 0 ? base + u : base;
        publish = my_row + 1;
        uint8_t* row = rows + (size_t)my_row * ROW_STATE_BYTES;
        store_column<BLOCK>(col, row);
        int m, n;
        const uint64_t l1 = e.choices(0, (result >> 4) & 3, &m);
        const uint64_t l2 = e.choices(1, (result >> 6) & 3, &n);
        uint32_t cw[5] = {0, 0, 0, 0, 0};
        for (int i = 0; i < 9; i++) {
          if (i < m) cw[i >> 2] |= E::choice_at(l1, i) << (8 * (i & 3));
          if (i < n) cw[(9 + i) >> 2] |= E::choice_at(l2, i) << (8 * ((9 + i) & 3));
        }
        uint4* tail = reinterpret_cast<uint4*>(row + 384);
        tail[0] = make_uint4(e.dur[0], e.dur[1], result | (stored << 8) | ((uint32_t)m << 16) | ((uint32_t)n << 24), (uint32_t)p);
        tail[1] = make_uint4(cw[0], cw[1], cw[2], cw[3]);
        tail[2] = make_uint4(cw[4], 0, 0, 0);
      }
    }
    // ---- publish: one queue reservation per warp, entries written after the staged states are visible device-wide
    {
      const uint32_t mask = __ballot_sync(0xFFFFFFFFu, publish != 0);
      if (mask) {
        __threadfence();
        uint32_t at = 0;
        if (lane == (__ffs(mask) - 1)) at = atomicAdd(q_tail, (uint32_t)__popc(mask));
        at = __shfl_sync(0xFFFFFFFFu, at, __ffs(mask) - 1);
        if (publish != 0) queue[at + __popc(mask & ((1u << lane) - 1u))] = publish;
      }
    }
    if (have) {
      const uint32_t hdr = stream[p], c1 = stream[p + 1], c2 = stream[p + 2];
      e.act[0] = e.act[1] = 0;
      result = e.update(c1, c2);
      p += 11 + 4 * ((hdr & 15u) + (hdr >> 4) + 2);
      u++;
      if (u == n_updates) {
        if (only < 0 && (result != stored || p != end)) atomicMin(first_bad, g);
        have = false;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// MCTS::randomize_hidden_variables (reference cpp/include/search/durations.h:25-97). Touches a few
// bytes per battle, so it works on the AoS records in place. Every draw uses the same battle.rng.
__device__ __constant__ uint64_t HIDDEN_BINDING_CUM[4] = {
    // thresholds (cumulative counts out of 40) of the 4x40 lookup rows, 8 bits each: n(<=1), n(<=2), n(<=3)
    15ull | (30ull << 8) | (36ull << 16), 24ull | (32ull << 8) | (40ull << 16), 20ull | (40ull << 8) | (40ull << 16),
    40ull | (40ull << 8) | (40ull << 16)};

__global__ void randomize_hidden_kernel(uint8_t* __restrict__ battles, const uint8_t* __restrict__ durations, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint8_t* b = battles + (size_t)i * 384;
  const uint64_t rng = *reinterpret_cast<const uint64_t*>(b + B_RNG);
  const uint2 dd = *reinterpret_cast<const uint2*>(durations + (size_t)i * 8);
  for (int s = 0; s < 2; s++) {
    const uint32_t d = s == 0 ? dd.x : dd.y;
    if (d == 0) continue;
    uint8_t* side = b + s * SIDE_SIZE;
    uint64_t v = *reinterpret_cast<uint64_t*>(side + A_VOL);
    const uint32_t confusion = bf_get(d, DF_CONFUSION);
    if (confusion) {
      const uint32_t one = confusion == 1;
      const uint32_t max = 6 - (confusion + one);
      v = vf_set(v, VF_CONFUSION, (uint32_t)(rng % max) + 1 + one);
    }
    const uint32_t disable = bf_get(d, DF_DISABLE);
    if (disable) v = vf_set(v, VF_DISABLE_DURATION, (uint32_t)(rng % (9 - disable)) + 1);
    const uint32_t attacking = bf_get(d, DF_ATTACKING);
    if (attacking && (v & (V_BIDE | V_THRASHING))) {
      v = vf_set(v, VF_ATTACKS, attacking == 3 ? 1u : 4u - (attacking + (uint32_t)(rng % 2)));
    }
    const uint32_t binding = bf_get(d, DF_BINDING);
    if (binding) {
      const uint32_t idx = (uint32_t)(rng % 40);
      const uint64_t cum = HIDDEN_BINDING_CUM[binding - 1];
      const uint32_t a = 1 + (idx >= (cum & 255)) + (idx >= ((cum >> 8) & 255)) + (idx >= ((cum >> 16) & 255));
      v = vf_set(v, VF_ATTACKS, a);
    }
    *reinterpret_cast<uint64_t*>(side + A_VOL) = v;
    for (int k = 0; k < 6; k++) {
      const uint32_t sleep = (d >> (3 * k)) & 7u;
      if (!sleep) continue;
      uint8_t* pk = side + ((int)side[S_ORDER + k] - 1) * PK_SIZE;
      const uint32_t st = pk[PK_STATUS];
      if ((st & 7u) && !(st & SB_EXT)) pk[PK_STATUS] = (uint8_t)((st & 0xF8u) | ((uint32_t)(rng % (8 - sleep)) + 1));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// build_kernel: PKMN::battle(p1, p2, seed) + update(0, 0) (reference cpp/include/libpkmn/pkmn.h:50-57,
// init.h:35-40,90-154, data/moves.h:1794-1796, cpp/src/generate.cc:238-240).
// teams[n][2][6][5] = {species, move x4}; level 100, DV 15, max stat experience.
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK) build_kernel(const uint8_t* __restrict__ teams, const uint64_t* __restrict__ seeds,
                                                       uint8_t* __restrict__ battles, uint8_t* __restrict__ durations,
                                                       uint8_t* __restrict__ results, uint32_t n, uint32_t only_finished) {
  extern __shared__ uint32_t tile[];
  const uint32_t i = blockIdx.x * BLOCK + threadIdx.x;
  if (i >= n) return;
  if (only_finished && (results[i] & 15u) == 0) return;  // respawn: only slots whose battle has ended are rebuilt
  uint32_t* col = tile + threadIdx.x;
  for (int w = 0; w < BATTLE_WORDS; w++) col[w * BLOCK] = 0;
  Mem<BLOCK> m;
  m.set(reinterpret_cast<uint8_t*>(col));
  const uint8_t* t = teams + (size_t)i * 60;
  for (int s = 0; s < 2; s++) {
    for (int k = 0; k < 6; k++) {
      const uint8_t* set = t + (s * 6 + k) * 5;
      const uint32_t species = set[0];
      const int pk = s * SIDE_SIZE + k * PK_SIZE;
      if (species == 0) {
        if (k == 0) m.st8(s * SIDE_SIZE + S_ORDER, 1);
        continue;
      }
      for (int q = 0; q < 5; q++) {
        const uint32_t base = GEN1_SPECIES_STATS[5 * species + q];
        const uint32_t core = 2 * (base + 15) + 63;
        m.st16(pk + PK_STATS + 2 * q, core * 100 / 100 + (q == 0 ? 110u : 5u));
      }
      for (int q = 0; q < 4; q++) {
        const uint32_t mv = set[1 + q];
        const uint32_t pp = GEN1_MOVE_PP[mv] / 5 * 8;
        m.st8(pk + PK_MOVES + 2 * q, mv);
        m.st8(pk + PK_MOVES + 2 * q + 1, mv ? (pp < 61 ? pp : 61) : 0);
      }
      m.st16(pk + PK_HP, m.ld16(pk + PK_STATS));
      m.st8(pk + PK_SPECIES, species);
      m.st8(pk + PK_TYPES, GEN1_SPECIES_TYPES[species]);
      m.st8(pk + PK_LEVEL, 100);
      m.st8(s * SIDE_SIZE + S_ORDER + k, k + 1);
    }
  }
  m.st64(B_RNG, seeds[i]);
  Engine<BLOCK, false, true> e;
  e.begin(reinterpret_cast<uint8_t*>(col), 0, 0);
  const uint32_t r = e.update(0, 0);
  e.end();
  store_column<BLOCK>(col, battles + (size_t)i * 384);
  *reinterpret_cast<uint2*>(durations + (size_t)i * 8) = make_uint2(e.dur[0], e.dur[1]);
  results[i] = (uint8_t)r;
}

}  // namespace oakb200
