// Tensor-core leaf evaluation for sm_100a (tcgen05 + Tensor Memory), included by value_net.cu.
//
// Same function as eval_kernel (value_net.cu) and as the reference's
// NN::Battle::NetworkImpl::value_policy_inference (cpp/include/nn/battle/network.h:102-175), organised as
// batched GEMMs so that every dense contraction runs on the 5th-generation tensor cores:
//
//   CTA tile = 128 evaluations; thread r of the four worker warps owns evaluation r = TMEM lane r.
//   For each of the 12 party slots (2 sides x {active, 5 benched}):
//     L1  hidden[128 x 128]  = onehot-ish features[128 x K] * W0^T   K = 208 (Pokemon) or 240 + 208 (Active)
//     L2  embed [128 x 64|96] = act(hidden + b0) * W1^T
//     T   trunk [128 x 64]  += [embed | hp] * fc0[:, slot columns]^T  (the 768-wide battle embedding never exists)
//   then fc1, and one N = 160 GEMM for the three head hidden layers (value 32 | policy p1 64 | policy p2 64).
//   The last layers (value 32 -> 1, policy 64 -> the <= 9 legal rows of 315) are per-thread fp32 dot products.
//
// Precision: operands are split bf16 pairs x = hi + lo (umma::split_bf16); each GEMM issues
// A_hi*B_hi + A_hi*B_lo + A_lo*B_hi into one fp32 accumulator, i.e. ~16 mantissa bits per operand
// (error ~1e-5 relative, well inside the 1e-3 of the north star) at the smem footprint of one tf32 plane.
// One-hot feature columns are exact in bf16, so the first layer needs its A_lo plane only for the 14 real-valued
// columns, which the host-side packing permutes to the front of K.
//
// Where the operands live: only the first-layer features of one party slot are in shared memory (scattered one-hot
// stores); every layer OUTPUT is rewritten in place in Tensor Memory as the split-bf16 A operand of the next layer
// (tcgen05.st after the epilogue, tcgen05.mma with A in TMEM; see tmem_operand), so the feature buffer is free as soon as a
// first layer has run and the twelve slots of a tile form one software pipeline (worker loop below).
//
// Roles: warps 0-7 workers, two threads per row (thread r of group g = warps 4g..4g+3 serves evaluation r = TMEM lane r;
// the groups split every epilogue by accumulator columns and the feature block by core columns), warp 8 lane 0 streams
// weight chunks L2 -> shared memory with cp.async.bulk through a 3-slot mbarrier ring (and owns the TMEM allocation),
// warp 9's elected lane issues tcgen05.mma, warps 10-13 run the last-layer dot products of the previous tile.
// The per-tile schedule is a static chunk program built on the host at oakb200_net_load (value_net.cu tc_pack).
// What bounds it (profiles/r2n_eval_tc_pipeline.txt): the workers' serial instruction latency per slot and the MMA count
// (an M = 128, K = 16 step costs 50-64 cycles whatever its N), not tensor throughput, shared-memory or L2 bandwidth.
#pragma once

#include "gen1_mem_ops.cuh"
#include "umma.cuh"

// hand-overs between the workers and the MMA issuer: OAKB200_TC_SPIN polls instead of the suspending try_wait (A/B switch)
#ifdef OAKB200_TC_SPIN
#define OAKB200_TC_WAIT(bar, parity) umma::mbar_wait_spin(bar, parity)
#else
#define OAKB200_TC_WAIT(bar, parity) umma::mbar_wait(bar, parity)
#endif

namespace tc {

using namespace oakb200;

struct Chunk {        // 16 bytes; one weight chunk = `ksteps` K=16 steps of a [N x K] matrix, hi plane then lo plane
  uint32_t gofs16;    // offset in the packed weight buffer, 16-byte units
  uint16_t bytes16;   // chunk bytes / 16
  uint8_t ksteps;
  uint8_t n8;         // N / 8
  uint16_t a_off16;   // shared-memory offset / 16 of the A hi plane at this chunk's first K step
  uint16_t alo_off16; // same for the A lo plane
  uint8_t lo_steps;   // leading K steps of this chunk for which an A lo plane exists
  uint8_t flags;
  uint16_t dcol;      // TMEM column of D
};
enum : uint8_t { F_STAGE_BEGIN = 1, F_OVERWRITE = 2, F_END_D = 4, F_END_T = 8, F_A_TMEM = 16, F_WAIT_A2 = 32, F_WAIT_E = 64, F_END_E = 128 };
// The benched slots are software-pipelined, so two hand-overs of one direction can be outstanding at once; each kind has its
// own mbarrier (a parity wait must never fall two phases behind): a stage begins on bar_a (features / everything else),
// bar_a2 (F_WAIT_A2: hidden layer in Tensor Memory) or bar_ae (F_WAIT_E: benched embedding operand); its end is committed to
// bar_d, bar_e (F_END_E: second layer of a benched slot) or bar_t.
// F_A_TMEM: the A operand is one the workers left in Tensor Memory in place of the previous layer's accumulator; alo_off16 =
// tmem_operand(...) below, a_off16 = the index of the chunk's first K step, both planes exist for every step

const int ROWS = 128;
#ifndef OAKB200_TC_GROUPS
#define OAKB200_TC_GROUPS 2
#endif
const int GROUPS = OAKB200_TC_GROUPS;  // threads per row (2 or 4): they split the epilogue columns and the zero-fill
static_assert(GROUPS == 2 || GROUPS == 4, "two or four worker threads per row");
const int WORKERS = GROUPS * ROWS;
const int THREADS = WORKERS + 64 + ROWS;
const int W_STREAM = WORKERS / 32, W_MMA = WORKERS / 32 + 1, W_HEADS = WORKERS / 32 + 2;  // W_HEADS .. +3: the head warps
static_assert(W_HEADS % 4 == 2, "head warp w serves TMEM lane quarter w % 4");
const uint32_t OFF_UNION = 0;        // A1: the first-layer features of one party slot (the only operand in shared memory)
const uint32_t OFF_TILE = 65536;     // 128 battle records as word-transposed columns (gen1_state.cuh)
const uint32_t OFF_RING = 114688;    // weight chunk ring
#ifndef OAKB200_TC_RING_SLOTS
#define OAKB200_TC_RING_SLOTS 3
#endif
// 3 x 32 KB: a bulk copy costs the copy engine ~400 cycles whatever its size (tools/probes/bulk_rate_probe.cu),
// so big chunks; three slots let the streamer run two chunks ahead of the MMA issuer
const uint32_t RING_SLOTS = OAKB200_TC_RING_SLOTS, RING_SLOT = 98304 / RING_SLOTS;
static_assert(RING_SLOTS >= 2 && RING_SLOTS <= 4, "four full / empty barrier pairs are laid out");
const uint32_t OFF_PROG = OFF_RING + RING_SLOTS * RING_SLOT;
const uint32_t PROG_MAX = 192;
const uint32_t OFF_CONST = OFF_PROG + PROG_MAX * 16;
const uint32_t CONST_FLOATS = 740;
const uint32_t OFF_BAR = OFF_CONST + CONST_FLOATS * 4;  // full[4], empty[4], a_ready, d_ready, t_done, heads_ready, heads_free, a2_ready, ae_ready, e_done
const uint32_t OFF_TMEM_SLOT = OFF_BAR + 16 * 8;
const uint32_t OFF_ROWIDX = OFF_TMEM_SLOT + 8;          // [2 players][128 rows][12] u16: last-layer rows of the legal choices
const uint32_t SMEM_BYTES = OFF_ROWIDX + 2 * ROWS * 24;
static_assert(OFF_ROWIDX % 8 == 0, "row index records are written as 32-bit words");
static_assert(OFF_BAR % 8 == 0, "mbarriers are 8-byte aligned");
static_assert(SMEM_BYTES <= 232448, "fits the 227 KB of one CTA");

// operand planes (byte offsets from the start of dynamic shared memory); one core column = 128 rows x 16 B
const uint32_t CORE = ROWS * 16;
const uint32_t A1_HI = OFF_UNION, A1_LO = OFF_UNION + 30 * CORE;   // K <= 240 hi, 16 lo
// TMEM columns
const uint32_t TMEM_COLS = 512, D_WORK = 0, D_TRUNK = 192, D_HEADS = 256, D_ACT = 416;  // one CTA per SM: all of Tensor Memory
const uint32_t D_EMB = 128;  // second-layer accumulator of the benched slots (64 columns): beside the first layer's, not over it
// Operands that stay in Tensor Memory. An accumulator block of NC fp32 columns owned by one worker group (NC = 64: half of
// a hidden layer, 32: half of a benched embedding) is rewritten IN PLACE by that group as the split-bf16 A operand of the
// next layer (two K elements per 32-bit column, a K = 16 step = 8 columns): hi plane in the block's first NC / 2 columns,
// lo plane in the other NC / 2, so a thread only ever overwrites columns it has itself read (no barrier between groups).
// K step k of such an operand, kpb = NC / 16 steps per block: hi at base + 16 kpb (k / kpb) + 8 (k % kpb), lo 8 kpb on.
// Layout of an A operand in Tensor Memory: tools/probes/ts_mma_probe*.cu.
const int HID_NC = 128 / GROUPS, EMB_NC = 64 / GROUPS;  // columns per worker group of a hidden layer / a benched embedding
const int ACT_NC = 96 / GROUPS;                          // ... of the active Pokemon's embedding (K = 96 of the trunk)
static_assert(GROUPS == 2, "the in-place operand blocks (64 / 48 / 32 columns per group) are laid out for two worker groups");
__host__ __device__ inline uint32_t tmem_operand(uint32_t base, uint32_t kpb) { return base | (kpb << 12); }  // Chunk::alo_off16
// constants block (floats)
const int C_PK0 = 0, C_PK1 = 128, C_AC0 = 192, C_AC1 = 320, C_FC0 = 416, C_FC1 = 480, C_HEADS = 544, C_V3W = 704, C_V3B = 736;
const int K_POK = 208, K_ACT = 240;

// Feature columns after the host-side permutation. Every block of columns starts on a core column (8 columns), the
// real-valued columns come first (they alone need an A lo plane) and the move block comes last, so that two worker
// groups can each zero-fill and write their own core columns without meeting: group 0 everything but the moves, group 1
// the moves.
__host__ __device__ inline int act_col(int idx) {   // Active block (battle.h:480-489), 229 -> K_ACT = 240
  if (idx < 5) return idx;           // stats                          0..4   ->   0..4
  if (idx < 20) return idx + 11;     // types                          5..19  ->  16..30
  if (idx < 26) return idx - 15;     // boosts                        20..25  ->   5..10
  if (idx < 42) return idx + 6;      // flags                         26..41  ->  32..47
  if (idx < 45) return idx - 31;     // state / substitute / toxic    42..44  ->  11..13
  if (idx < 209) return idx + 27;    // moves                         45..208 ->  72..235
  return idx - 161;                  // durations                    209..228 ->  48..67
}
__host__ __device__ inline int pok_col(int idx) {   // Pokemon block (battle.h:208-214), 198 -> K_POK = 208
  if (idx < 5) return idx;           // stats                          0..4   ->   0..4
  if (idx < 169) return idx + 3;     // moves                          5..168 ->   8..171
  if (idx < 183) return idx + 7;     // status                       169..182 -> 176..189
  return idx + 9;                    // types                        183..197 -> 192..206
}
const int POK_MOVE_CORE0 = 1, POK_MOVE_CORE1 = 22, ACT_MOVE_CORE0 = 9, ACT_MOVE_CORE1 = 30;  // core columns of the move blocks

struct Params {
  const uint8_t* wpack;
  const Chunk* prog;
  const float* consts;
  const float *p1bW, *p1bb, *p2bW, *p2bb;
  int nchunks, act, pod, aod, vhd, pohd;
};

// AoS record (384 B, 16-byte aligned) -> column of the word-transposed tile; the four threads of a row take a quarter each
__device__ __forceinline__ void load_record_part(uint32_t* col, const uint8_t* record, int part) {
  if ((reinterpret_cast<uintptr_t>(record) & 31u) == 0) {  // whole 32-byte sectors with 256-bit loads (battle_kernels.cuh: load_column)
    const int per8 = 12 / GROUPS;
#pragma unroll
    for (int k = per8 * part; k < per8 * part + per8; k++) {
      uint32_t v[8];
      asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                   : "l"(record + 32 * k));
#pragma unroll
      for (int j = 0; j < 8; j++) col[(8 * k + j) * ROWS] = v[j];
    }
    return;
  }
  const uint4* src = reinterpret_cast<const uint4*>(record);
  const int per = 24 / GROUPS;
#pragma unroll
  for (int k = per * part; k < per * part + per; k++) {
    const uint4 v = __ldg(src + k);
    col[(4 * k + 0) * ROWS] = v.x;
    col[(4 * k + 1) * ROWS] = v.y;
    col[(4 * k + 2) * ROWS] = v.z;
    col[(4 * k + 3) * ROWS] = v.w;
  }
}

__device__ __forceinline__ void st16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((uint16_t)v) : "memory"); }
__device__ __forceinline__ void stv4(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
__device__ __forceinline__ uint32_t elem(uint32_t plane, int r, int k) { return plane + (uint32_t)((((k >> 3) * ROWS + r) << 4) + ((k & 7) << 1)); }
__device__ __forceinline__ float actf(float x, int act) {  // 1 relu, 2 clamp, 4 the int8 net's ClippedReLU (value_net.cu: activate)
  if (act == 4) return fminf(fmaxf(floorf(x * 0.015625f), 0.0f), 127.0f);
  x = fmaxf(x, 0.0f);
  return act == 2 ? fminf(x, 1.0f) : x;
}
__device__ __forceinline__ uint32_t bf16_count(int c) {  // 1.0, 2.0, 3.0, 4.0
  return c == 1 ? 0x3F80u : c == 2 ? 0x4000u : c == 3 ? 0x4040u : 0x4080u;
}

// eight consecutive K elements of one row -> one 16-byte store per plane
struct Row8 {
  uint32_t h[4], l[4];
  __device__ __forceinline__ void set(int i, float x) {
    uint32_t hi, lo;
    umma::split_bf16(x, hi, lo);
    if (i & 1) { h[i >> 1] |= hi << 16; l[i >> 1] |= lo << 16; }
    else { h[i >> 1] = hi; l[i >> 1] = lo; }
  }
  __device__ __forceinline__ void store(uint32_t hi_plane, uint32_t lo_plane, int core, int r) const {
    const uint32_t o = (uint32_t)((core * ROWS + r) << 4);
    stv4(hi_plane + o, h[0], h[1], h[2], h[3]);
    stv4(lo_plane + o, l[0], l[1], l[2], l[3]);
  }
};

// one-hot entries for up to four move slots, summing duplicates as the reference's sparse list does
__device__ __forceinline__ void put_moves(uint32_t plane, int r, int col0, const uint32_t (&id)[4], const uint32_t (&pp)[4]) {
  bool ok[4];
#pragma unroll
  for (int k = 0; k < 4; k++) ok[k] = id[k] != MOVE_None && id[k] != MOVE_Struggle && pp[k] != 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    if (!ok[k]) continue;
    int cnt = 0;
    bool first = true;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const bool same = ok[j] && id[j] == id[k];
      cnt += same ? 1 : 0;
      if (j < k && same) first = false;
    }
    if (first) st16(elem(plane, r, col0 + (int)id[k] - 1), bf16_count(cnt));
  }
}

__device__ __forceinline__ void zero_cores(uint32_t hi, int r, int c0, int c1) {
  for (int c = c0; c < c1; c++) stv4(hi + (uint32_t)((c * ROWS + r) << 4), 0, 0, 0, 0);
}
// Pokemon block, pok_col() column order: group 0 writes stats / status / types, group 1 the moves. Two steps so that the
// loads, divisions and conversions of the NEXT slot run while this slot's first-layer MMAs do (the workers only wait
// then); what is left on the critical path is the stores.
struct PokemonFeatures {
  Row8 q;            // group 0: core column 0 (the five stats)
  uint32_t at[4];    // shared-memory addresses of up to four one-hot entries (0 = none) ...
  uint32_t val[4];   // ... and their bf16 values (1.0, or the count of a duplicated move)
};
__device__ __forceinline__ void prepare_pokemon(PokemonFeatures& f, uint32_t smem, int r, const Mem<ROWS>& m, int s, int slot,
                                                uint32_t dur, int grp) {
  const uint32_t hi = smem + A1_HI;
  const uint32_t id0 = m.ld8(Mem<ROWS>::sb(s), S_ORDER + slot - 1);
  const PH pk = Mem<ROWS>::pokemon(s, id0 ? (int)id0 - 1 : 0);
  const bool alive = id0 != 0 && m.ld16(pk, PK_HP) != 0;
#pragma unroll
  for (int k = 0; k < 4; k++) { f.at[k] = 0; f.val[k] = 0x3F80u; }
#pragma unroll
  for (int k = 0; k < 4; k++) f.q.h[k] = f.q.l[k] = 0;
  if (!alive) return;
  if (grp == 1) {
    uint32_t id[4], pp[4];
    bool ok[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      id[k] = m.ld8(pk, PK_MOVES + 2 * k);
      pp[k] = m.ld8(pk, PK_MOVES + 2 * k + 1);
      ok[k] = id[k] != MOVE_None && id[k] != MOVE_Struggle && pp[k] != 0;
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {  // as put_moves: duplicates add up on their first occurrence
      int cnt = 0;
      bool first = ok[k];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const bool same = ok[j] && id[j] == id[k];
        cnt += same ? 1 : 0;
        if (j < k && same) first = false;
      }
      if (first) {
        f.at[k] = elem(hi, r, 8 + (int)id[k] - 1);
        f.val[k] = bf16_count(cnt);
      }
    }
  } else if (grp == 0) {
#pragma unroll
    for (int c = 0; c < 5; c++) f.q.set(c, (float)m.ld16(pk, PK_STATS + 2 * c) / (c == 0 ? 703.0f : 999.0f));
    const uint32_t sleep = (dur >> (3 * (slot - 1))) & 7u;
    const uint32_t st = m.ld8(pk, PK_STATUS);
    if (st) f.at[0] = elem(hi, r, 176 + status_index(st, sleep));
    const uint32_t ty = m.ld8(pk, PK_TYPES), t1 = ty & 15u, t2 = ty >> 4;
    f.at[1] = elem(hi, r, 192 + (int)t1);
    if (t2 != t1) f.at[2] = elem(hi, r, 192 + (int)t2);
  }
}
__device__ __forceinline__ void emit_pokemon(const PokemonFeatures& f, uint32_t smem, int r, int grp) {
  const uint32_t hi = smem + A1_HI, lo = smem + A1_LO;
  if (grp == 1) {
    zero_cores(hi, r, POK_MOVE_CORE0, POK_MOVE_CORE1);
  } else if (grp == 0) {
    zero_cores(hi, r, POK_MOVE_CORE1, K_POK / 8);
    stv4(lo + (uint32_t)((ROWS + r) << 4), 0, 0, 0, 0);
    f.q.store(hi, lo, 0, r);
  } else {
    return;
  }
#pragma unroll
  for (int k = 0; k < 4; k++)
    if (f.at[k]) st16(f.at[k], f.val[k]);
}

// Active block, act_col() column order: group 0 writes everything but the moves, group 1 the moves
__device__ __forceinline__ void build_active(uint32_t smem, int r, const Mem<ROWS>& m, int s, uint32_t dur, bool alive, int grp) {
  const uint32_t hi = smem + A1_HI, lo = smem + A1_LO;
  const H sb = Mem<ROWS>::sb(s);
  if (grp == 1) {
    zero_cores(hi, r, ACT_MOVE_CORE0, ACT_MOVE_CORE1);
    if (!alive) return;
    uint32_t id[4], pp[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      id[k] = m.ld8(sb, A_MOVES + 2 * k);
      pp[k] = m.ld8(sb, A_MOVES + 2 * k + 1);
    }
    put_moves(hi, r, 72, id, pp);
    return;
  }
  if (grp != 0) return;
  zero_cores(hi, r, 2, ACT_MOVE_CORE0);
  Row8 q0, q1;
  if (!alive) {
#pragma unroll
    for (int i = 0; i < 4; i++) q0.h[i] = q0.l[i] = q1.h[i] = q1.l[i] = 0;
    q0.store(hi, lo, 0, r);
    q1.store(hi, lo, 1, r);
    return;
  }
#pragma unroll
  for (int c = 0; c < 5; c++) q0.set(c, (float)m.ld16(sb, A_STATS + 2 * c) / (c == 0 ? 703.0f : 999.0f));
  const uint32_t b = m.ld32(sb, A_BOOSTS) & 0xFFFFFFu;
  float bv[6];
#pragma unroll
  for (int k = 0; k < 6; k++) {
    const int stage = boost_get(b, k);
    bv[k] = ((float)boost_num(stage) / (float)boost_den(stage)) * (k < 4 ? 0.25f : (1.0f / 3.0f));
  }
  q0.set(5, bv[0]);
  q0.set(6, bv[1]);
  q0.set(7, bv[2]);
  q1.set(0, bv[3]);
  q1.set(1, bv[4]);
  q1.set(2, bv[5]);
  const uint32_t vlo = m.vlo(s), vhi = m.vhi(s);
  q1.set(3, (float)((vlo >> 24) | ((vhi & 255u) << 8)) / 65535.0f);
  q1.set(4, (float)bf_get(vhi, VHI_SUBSTITUTE) / 177.0f);
  q1.set(5, (float)bf_get(vhi, VHI_TOXIC) / 16.0f);
  q1.h[3] = q1.l[3] = 0;
  q0.store(hi, lo, 0, r);
  q1.store(hi, lo, 1, r);
  const uint32_t ty = m.ld8(sb, A_TYPES), t1 = ty & 15u, t2 = ty >> 4;
  st16(elem(hi, r, 16 + (int)t1), 0x3F80u);
  if (t2 != t1) st16(elem(hi, r, 16 + (int)t2), 0x3F80u);
  // bide thrashing | charging binding invulnerable confusion mist focus_energy substitute recharging rage leech_seed
  // toxic light_screen reflect transform (MultiHit = bit 2 and Flinch = bit 3 are not encoded)
  uint32_t fl = (vlo & 3u) | ((vlo >> 2) & 0xFFFCu);
  while (fl) {
    const int j = __ffs(fl) - 1;
    fl &= fl - 1;
    st16(elem(hi, r, 32 + j), 0x3F80u);
  }
  uint32_t n;
  if ((n = bf_get(dur, DF_CONFUSION)) != 0) st16(elem(hi, r, 48 + (int)n - 1), 0x3F80u);
  if ((n = bf_get(dur, DF_DISABLE)) != 0) st16(elem(hi, r, 53 + (int)n - 1), 0x3F80u);
  if ((n = bf_get(dur, DF_ATTACKING)) != 0) st16(elem(hi, r, 61 + (int)n - 1), 0x3F80u);
  if ((n = bf_get(dur, DF_BINDING)) != 0) st16(elem(hi, r, 64 + (int)n - 1), 0x3F80u);
}

// two fp32 -> packed split-bf16: hi = {bf16(x1), bf16(x0)}, lo = the rounded residuals (x0 in the low half)
__device__ __forceinline__ void split2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(x1), "f"(x0));
  const float r0 = x0 - __uint_as_float(hi << 16), r1 = x1 - __uint_as_float(hi & 0xFFFF0000u);
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(r1), "f"(r0));
}

// One accumulator element -> activation: min(max(D + bias, 0), bound); bound = 1 for the clamp activation, +inf for ReLU,
// 0 for an empty slot's row. MODE (discrete networks): 1 = then floor(127 x), the u8 an embedding becomes (cache.h:106,234); 2 = the int8 net's
// ClippedReLU clamp((D + bias) >> 6, 0, 127) instead. Both leave small integers: the hi plane is exact, the lo plane zero.
template <int MODE>
__device__ __forceinline__ float epi_act(float v, float b, float bound) {
  if (MODE == 2) return fminf(fmaxf(floorf((v + b) * 0.015625f), 0.0f), 127.0f);
  const float x = fminf(fmaxf(v + b, 0.0f), bound);
  return MODE == 1 ? floorf(127.0f * x) : x;
}
// In-place epilogue (see tmem_operand): NC = 16 j fp32 accumulator columns at t0 -> + bias, activation (MODE as epi_act)
// -> hi plane in columns [t0, t0 + NC / 2), lo plane in [t0 + NC / 2, t0 + NC). hp_col in [0, NC): that K element is the
// hp / max-hp input of the next layer instead (network.h:147,165); pass -1 for none.
template <int MODE, int NC>
__device__ __forceinline__ void epilogue_inplace_m(uint32_t t0, const float* bias, float bound, int hp_col, float hp) {
  static_assert(NC == 16 || NC == 32 || NC == 48 || NC == 64, "blocks of 16, 32, 48 or 64 accumulator columns");
  uint32_t v[64];
  if (NC == 16) {
    umma::tmem_ld16(t0, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
  } else {
    uint32_t(&a)[32] = *reinterpret_cast<uint32_t(*)[32]>(&v[0]);
    umma::tmem_ld32(t0, a);
    if (NC == 64) {
      uint32_t(&b)[32] = *reinterpret_cast<uint32_t(*)[32]>(&v[32]);
      umma::tmem_ld32(t0 + 32u, b);
    } else if (NC == 48) {
      uint32_t(&b)[16] = *reinterpret_cast<uint32_t(*)[16]>(&v[32]);
      umma::tmem_ld16(t0 + 32u, b);
    }
  }
  umma::tmem_ld_wait();
  uint32_t h[32], l[32];
#pragma unroll
  for (int g = 0; g < NC / 8; g++) {
    const float4 ba = *reinterpret_cast<const float4*>(bias + 8 * g), bb = *reinterpret_cast<const float4*>(bias + 8 * g + 4);
    float x[8];
    x[0] = epi_act<MODE>(__uint_as_float(v[8 * g + 0]), ba.x, bound);
    x[1] = epi_act<MODE>(__uint_as_float(v[8 * g + 1]), ba.y, bound);
    x[2] = epi_act<MODE>(__uint_as_float(v[8 * g + 2]), ba.z, bound);
    x[3] = epi_act<MODE>(__uint_as_float(v[8 * g + 3]), ba.w, bound);
    x[4] = epi_act<MODE>(__uint_as_float(v[8 * g + 4]), bb.x, bound);
    x[5] = epi_act<MODE>(__uint_as_float(v[8 * g + 5]), bb.y, bound);
    x[6] = epi_act<MODE>(__uint_as_float(v[8 * g + 6]), bb.z, bound);
    x[7] = epi_act<MODE>(__uint_as_float(v[8 * g + 7]), bb.w, bound);
#pragma unroll
    for (int j = 0; j < 8; j++) x[j] = (8 * g + j == hp_col) ? hp : x[j];
#pragma unroll
    for (int j = 0; j < 4; j++) split2(x[2 * j], x[2 * j + 1], h[4 * g + j], l[4 * g + j]);
  }
  if (NC == 16) {
    umma::tmem_st8(t0, *reinterpret_cast<uint32_t(*)[8]>(&h[0]));
    umma::tmem_st8(t0 + 8u, *reinterpret_cast<uint32_t(*)[8]>(&l[0]));
  } else {
    umma::tmem_st16(t0, *reinterpret_cast<uint32_t(*)[16]>(&h[0]));
    if (NC == 64) umma::tmem_st16(t0 + 16u, *reinterpret_cast<uint32_t(*)[16]>(&h[16]));
    if (NC == 48) umma::tmem_st8(t0 + 16u, *reinterpret_cast<uint32_t(*)[8]>(&h[16]));
    umma::tmem_st16(t0 + NC / 2, *reinterpret_cast<uint32_t(*)[16]>(&l[0]));
    if (NC == 64) umma::tmem_st16(t0 + NC / 2 + 16u, *reinterpret_cast<uint32_t(*)[16]>(&l[16]));
    if (NC == 48) umma::tmem_st8(t0 + NC / 2 + 16u, *reinterpret_cast<uint32_t(*)[8]>(&l[16]));
  }
  umma::tmem_st_wait();
}
template <int NC>
__device__ __forceinline__ void epilogue_inplace(uint32_t t0, const float* bias, float bound, int mode, int hp_col = -1, float hp = 0.0f) {
  if (mode == 0) epilogue_inplace_m<0, NC>(t0, bias, bound, hp_col, hp);
  else if (mode == 1) epilogue_inplace_m<1, NC>(t0, bias, bound, hp_col, hp);
  else epilogue_inplace_m<2, NC>(t0, bias, bound, hp_col, hp);
}
// CL = CTAs per cluster. CL = 2: the two CTAs of a cluster walk the chunk program in step on two different tiles, and
// each weight chunk is read from L2 ONCE for both — each CTA's streamer fetches one half and multicasts it into both
// rings — because the L2 -> SM weight stream (2.2 MB per tile, every SM asking for the same lines) is what bounds the
// single-CTA form (profiles/r1f_eval_tc_ncu_summary.txt). A ring slot is reused when BOTH issuers have released it.
template <int CL>
__global__ void __launch_bounds__(THREADS, 1)
eval_tc_kernel(Params P, const uint8_t* __restrict__ battles, const uint8_t* __restrict__ durations, const uint8_t* __restrict__ results,
               float* __restrict__ values, float* __restrict__ logits, uint8_t* __restrict__ counts, uint32_t n) {
  extern __shared__ __align__(1024) uint8_t dsm[];
  const uint32_t smem = umma::smem_u32(dsm);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar = smem + OFF_BAR;
  const uint32_t bar_full = bar, bar_empty = bar + 32, bar_a = bar + 64, bar_d = bar + 72, bar_t = bar + 80;
  const uint32_t bar_h = bar + 88, bar_hfree = bar + 96, bar_a2 = bar + 104, bar_ae = bar + 112, bar_e = bar + 120;
  const uint32_t ntiles = (n + ROWS - 1) / ROWS;
  // CTA `crank` of cluster `blockIdx.x / CL` takes tile base + crank of every group of CL tiles the cluster visits; a
  // tile index >= ntiles is a dummy (rows clamp to the last evaluation, nothing is written) so that the CTAs of a
  // cluster always run the same number of chunk programs
  const uint32_t crank = CL > 1 ? umma::cluster_ctarank() : 0u;
  const uint32_t tile0 = (blockIdx.x / CL) * CL, tile_step = gridDim.x;
  const uint16_t cmask = (uint16_t)((1u << CL) - 1u);

  if (threadIdx.x == 0) {
    for (int i = 0; i < (int)RING_SLOTS; i++) {
      umma::mbar_init(bar_full + 8 * i, 1);
      umma::mbar_init(bar_empty + 8 * i, CL);
    }
    umma::mbar_init(bar_a, WORKERS / 32);
    umma::mbar_init(bar_a2, WORKERS / 32);
    umma::mbar_init(bar_ae, WORKERS / 32);
    umma::mbar_init(bar_e, 1);
    umma::mbar_init(bar_d, 1);
    umma::mbar_init(bar_t, 1);
    umma::mbar_init(bar_h, 2 * ROWS / 32);  // one arrival per warp of worker groups 0 and 1
    umma::mbar_init(bar_hfree, ROWS / 32);  // one per head warp
    umma::mbar_fence_init();
  }
  if (warp == W_STREAM) umma::tmem_alloc(smem + OFF_TMEM_SLOT, TMEM_COLS);
  {
    const uint4* src = reinterpret_cast<const uint4*>(P.prog);
    uint4* dst = reinterpret_cast<uint4*>(dsm + OFF_PROG);
    for (int i = threadIdx.x; i < P.nchunks; i += THREADS) dst[i] = src[i];
    float* cdst = reinterpret_cast<float*>(dsm + OFF_CONST);
    for (int i = threadIdx.x; i < (int)CONST_FLOATS; i += THREADS) cdst[i] = P.consts[i];
  }
  umma::tc_fence_before();
  __syncthreads();
  if (CL > 1) umma::cluster_sync();  // every CTA's barriers exist before any peer multicasts into them
  umma::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(dsm + OFF_TMEM_SLOT);
  const Chunk* prog = reinterpret_cast<const Chunk*>(dsm + OFF_PROG);

  if (warp == W_STREAM) {
    // ===== weight streamer: packed chunks L2 -> ring =====
    if (lane == 0) {
      uint32_t g = 0;
#ifdef OAKB200_TC_TRACE  // diagnostic build only: clock64 stamps of the issuer over the stages of one tile of CTA 0
      long long tr_i[400];
      int tr_n = 0, tr_tile = 0;
#define TR_I(tag) do { if (blockIdx.x == 0 && tr_tile == 2 && tr_n < 400) tr_i[tr_n++] = (clock64() << 4) | (tag); } while (0)
#else
#define TR_I(tag) do { } while (0)
#endif
      for (uint32_t t = tile0; t < ntiles; t += tile_step) {
#ifdef OAKB200_TC_TRACE
        tr_tile++;
#endif
        for (int c = 0; c < P.nchunks; c++, g++) {
          const uint32_t slot = g % RING_SLOTS;
          TR_I(6);
          if (g >= RING_SLOTS) umma::mbar_wait(bar_empty + 8 * slot, ((g / RING_SLOTS) - 1) & 1);
          TR_I(7);
          const uint32_t bytes = (uint32_t)prog[c].bytes16 << 4;
#ifdef OAKB200_TC_NO_STREAM  // timing experiment only: the ring is never filled, results are garbage
          umma::mbar_arrive(bar_full + 8 * slot);
          continue;
#endif
          umma::mbar_expect_tx(bar_full + 8 * slot, bytes);
          const uint8_t* src = P.wpack + ((size_t)prog[c].gofs16 << 4);
          const uint32_t dst = smem + OFF_RING + slot * RING_SLOT;
          if (CL > 1) {  // this CTA's half of the chunk, into both rings; the peer sends the other half
            const uint32_t h0 = ((uint32_t)prog[c].bytes16 >> 1) << 4;
            const uint32_t ofs = crank ? h0 : 0u, len = crank ? bytes - h0 : h0;
            umma::bulk_g2s_multicast(dst + ofs, src + ofs, len, bar_full + 8 * slot, cmask);
          } else {
            umma::bulk_g2s(dst, src, bytes, bar_full + 8 * slot);
          }
        }
      }
#ifdef OAKB200_TC_TRACE
      if (blockIdx.x == 0)
        for (int i = 0; i < tr_n; i++) printf("[trace S] %d %lld\n", (int)(tr_i[i] & 15), tr_i[i] >> 4);
#endif
    }
  } else if (warp == W_MMA) {
    // ===== MMA issuer =====
    if (umma::elect_one()) {
      uint32_t g = 0, stage = 0, stage_a2 = 0, stage_e = 0;
#ifdef OAKB200_TC_TIMING  // diagnostic build only: where the issuer's time goes, by N of the stage (printed by CTA 0)
      long long tm0 = 0, tm1 = 0, tm2 = 0, tm3 = 0, tm4 = 0;  // wait workers, wait weights, issue, drain, commit
      uint32_t tm_d = 0, tm_t = 0;
      long long tm_c = clock64();
#define TM_LAP(acc) do { const long long now_ = clock64(); acc += now_ - tm_c; tm_c = now_; } while (0)
#endif
#ifdef OAKB200_TC_TRACE  // diagnostic build only: clock64 stamps of the issuer over the stages of one tile of CTA 0
      long long tr_i[400];
      int tr_n = 0, tr_tile = 0;
#define TR_I(tag) do { if (blockIdx.x == 0 && tr_tile == 2 && tr_n < 400) tr_i[tr_n++] = (clock64() << 4) | (tag); } while (0)
#else
#define TR_I(tag) do { } while (0)
#endif
      for (uint32_t t = tile0; t < ntiles; t += tile_step) {
#ifdef OAKB200_TC_TRACE
        tr_tile++;
#endif
        for (int c = 0; c < P.nchunks; c++, g++) {
          // The chunk record is read with one explicit 16-byte load and decoded by shifts: with the struct copy
          // `const Chunk ch = prog[c]` ptxas 12.9 was seen to take dcol from a register the load never wrote
          // (profiles/r2l_eval_tc_tmem.txt), which sent every accumulator but column 0 astray.
          struct { uint32_t ksteps, n8, a_off16, alo_off16, lo_steps, flags, dcol; } ch;
          {
            uint32_t w0, w1, w2, w3;
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3) : "r"(smem + OFF_PROG + 16u * (uint32_t)c));
            (void)w0;
            ch.ksteps = (w1 >> 16) & 0xffu;
            ch.n8 = w1 >> 24;
            ch.a_off16 = w2 & 0xffffu;
            ch.alo_off16 = w2 >> 16;
            ch.lo_steps = w3 & 0xffu;
            ch.flags = (w3 >> 8) & 0xffu;
            ch.dcol = w3 >> 16;
          }
#ifdef OAKB200_TC_TIMING
          TM_LAP(tm2);
#endif
          if (ch.flags & F_STAGE_BEGIN) TR_I(1);
          if (ch.flags & F_STAGE_BEGIN) {
            if (ch.flags & F_WAIT_A2) {
              OAKB200_TC_WAIT(bar_a2, stage_a2 & 1);
              stage_a2++;
            } else if (ch.flags & F_WAIT_E) {
              OAKB200_TC_WAIT(bar_ae, stage_e & 1);
              stage_e++;
            } else {
              OAKB200_TC_WAIT(bar_a, stage & 1);
              stage++;
            }
          }
#ifdef OAKB200_TC_TIMING
          TM_LAP(tm0);
#endif
          if (ch.flags & F_STAGE_BEGIN) TR_I(2);
          const uint32_t slot = g % RING_SLOTS;
          umma::mbar_wait(bar_full + 8 * slot, (g / RING_SLOTS) & 1);
          umma::tc_fence_after();
          TR_I(4);
#ifdef OAKB200_TC_TIMING
          TM_LAP(tm1);
#endif
          const uint32_t N = (uint32_t)ch.n8 << 3;
          const uint32_t idesc = umma::idesc_bf16_f32(N);
          const uint32_t kstep_b = N * 32;  // bytes of one K = 16 step of B (two core columns of N rows)
          const uint32_t b_hi = smem + OFF_RING + slot * RING_SLOT, b_lo = b_hi + ch.ksteps * kstep_b;
          const uint32_t a_hi = smem + ((uint32_t)ch.a_off16 << 4), a_lo = smem + ((uint32_t)ch.alo_off16 << 4);
          const uint32_t d = tmem + ch.dcol;
          // One thread issues every MMA of the CTA, so its instruction stream is the tile's critical path: descriptors
          // are built once per chunk and advanced by adding to their 14-bit address field (operands live below 256 KB,
          // so the field never carries into its neighbours).
          uint64_t da = umma::smem_desc(a_hi, CORE, 128), dal = umma::smem_desc(a_lo, CORE, 128);
          uint64_t dbh = umma::smem_desc(b_hi, N * 16, 128), dbl = umma::smem_desc(b_lo, N * 16, 128);
#ifdef OAKB200_TC_SWZ_TIMING  // timing experiment only: 128-byte-swizzled descriptors over the un-swizzled data, results are garbage
          da = umma::smem_desc(a_hi, 16, 1024) | (2ull << 61); dal = umma::smem_desc(a_lo, 16, 1024) | (2ull << 61);
          dbh = dbl = umma::smem_desc(smem + OFF_RING, 16, 1024) | (2ull << 61);  // stays inside the ring whatever N
          const uint64_t a_step = 2, b_step = 2;
#else
          const uint64_t a_step = (2 * CORE) >> 4, b_step = kstep_b >> 4;
#endif
          uint32_t acc = (ch.flags & F_OVERWRITE) ? 0u : 1u;
          const uint32_t lo_steps = ch.lo_steps, ksteps = ch.ksteps;
          if (ch.flags & F_A_TMEM) {  // A = an operand the workers left in Tensor Memory (tmem_operand)
            const uint32_t kpb = ch.alo_off16 >> 12, lo_ofs = 8u * kpb;
            uint32_t j = ch.a_off16 % kpb, a_t = tmem + (ch.alo_off16 & 0x3FFu) + 16u * kpb * (ch.a_off16 / kpb) + 8u * j;
            for (uint32_t i = 0; i < ksteps; i++) {
#ifdef OAKB200_TC_NO_COLLECTOR  // A/B switch: every MMA fetches its own A operand
              umma::mma_bf16_ts(d, a_t, dbh, idesc, acc);
              umma::mma_bf16_ts(d, a_t, dbl, idesc, 1u);
#else
              umma::mma_bf16_ts_pair(d, a_t, dbh, dbl, idesc, acc);
#endif
              umma::mma_bf16_ts(d, a_t + lo_ofs, dbh, idesc, 1u);
              acc = 1u;
              dbh += b_step;
              dbl += b_step;
              a_t += 8u;
              if (++j == kpb) {  // next block: skip this block's lo plane
                j = 0;
                a_t += lo_ofs;
              }
            }
          }
          const uint32_t ss_steps = (ch.flags & F_A_TMEM) ? 0u : ksteps;  // the K steps whose A operand is in shared memory
#ifdef OAKB200_TC_NO_MMA  // timing experiment only: no tensor-core work is issued, results are garbage
          if (ksteps) { da = dal = dbh = dbl = 0; }
          for (uint32_t i = 0; i < 0; i++) {
#elif defined(OAKB200_TC_HALF_MMA)  // timing experiment only: the hi x hi product alone
          for (uint32_t i = 0; i < ksteps; i++) {
            umma::mma_bf16(d, da, dbh, idesc, acc);
            acc = 1u;
            da += a_step;
            dbh += b_step;
            continue;
#else
          for (uint32_t i = 0; i < ss_steps; i++) {
#endif
#ifdef OAKB200_TC_WIDE_TIMING  // timing experiment only: A_hi x [B_hi | B_lo] as ONE instruction of twice the N, results are garbage
            if (N <= 128) umma::mma_bf16(d, da, dbh, umma::idesc_bf16_f32(2 * N), acc);
            else { umma::mma_bf16(d, da, dbh, idesc, acc); umma::mma_bf16(d, da, dbl, idesc, 1u); }
#elif defined(OAKB200_TC_NO_COLLECTOR)
            umma::mma_bf16(d, da, dbh, idesc, acc);
            umma::mma_bf16(d, da, dbl, idesc, 1u);
#else
            umma::mma_bf16_pair(d, da, dbh, dbl, idesc, acc);  // A_hi x B_hi, A_hi x B_lo: A is read once
#endif
            if (i < lo_steps) umma::mma_bf16(d, dal, dbh, idesc, 1u);
            acc = 1u;
            da += a_step;
            dal += a_step;
            dbh += b_step;
            dbl += b_step;
          }
#ifdef OAKB200_TC_TIMING
          TM_LAP(tm2);
#endif
          TR_I(5);
          if (CL > 1) umma::mma_commit_multicast(bar_empty + 8 * slot, cmask);
          else umma::mma_commit(bar_empty + 8 * slot);
          if (ch.flags & F_END_D) umma::mma_commit(bar_d);
          if (ch.flags & F_END_E) umma::mma_commit(bar_e);
          if (ch.flags & (F_END_D | F_END_E | F_END_T)) TR_I(3);
          if (ch.flags & F_END_T) umma::mma_commit(bar_t);
#ifdef OAKB200_TC_TIMING
          TM_LAP(tm4);
          if (ch.flags & F_END_D) { umma::mbar_wait(bar_d, tm_d & 1); tm_d++; TM_LAP(tm3); }
          if (ch.flags & F_END_T) { umma::mbar_wait(bar_t, tm_t & 1); tm_t++; TM_LAP(tm3); }
#endif
        }
      }
#ifdef OAKB200_TC_TRACE
      if (blockIdx.x == 0)
        for (int i = 0; i < tr_n; i++) printf("[trace I] %d %lld\n", (int)(tr_i[i] & 15), tr_i[i] >> 4);
#endif
#ifdef OAKB200_TC_TIMING
      if (blockIdx.x == 0)
        printf("[eval_tc timing] issuer: wait workers %lld, wait weights %lld, issue %lld, drain %lld, commit %lld cycles\n", tm0, tm1, tm2,
               tm3, tm4);
#endif
    }
  } else if (warp >= W_HEADS) {
    // ===== head warps: thread r serves row r = TMEM lane r of the tile the workers have just finished: value head
    // (32 -> 1) and the <= 9 legal rows of each policy head (64 -> 315), per-thread fp32 dot products, while the workers
    // and the tensor pipe are already on the next tile.
    const int r = 32 * (warp & 3) + lane;
    const float* cst = reinterpret_cast<const float*>(dsm + OFF_CONST);
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const bool disc = P.act == 3;
    const int act = disc ? 4 : P.act;
    const float out_div = disc ? 8128.0f : 1.0f;
    uint32_t iter = 0;
    for (uint32_t tb = tile0; tb < ntiles; tb += tile_step, iter++) {
      const uint32_t e = (tb + crank) * ROWS + (uint32_t)r;
      umma::mbar_wait(bar_h, iter & 1);
      umma::tc_fence_after();
      {
        uint32_t v[32];
        umma::tmem_ld32(tl + D_HEADS, v);
        umma::tmem_ld_wait();
        float acc = cst[C_V3B];
#pragma unroll
        for (int j = 0; j < 32; j++)
          if (j < P.vhd) acc = fmaf(actf(__uint_as_float(v[j]) + cst[C_HEADS + j], act), cst[C_V3W + j], acc);
        if (e < n) values[e] = 1.0f / (1.0f + expf(-(acc / out_div)));
      }
#pragma unroll 1
      for (int s = 0; s < 2; s++) {
        float h[64];
        {
          uint32_t v[32];
          umma::tmem_ld32(tl + D_HEADS + 32 + 64 * s, v);
          umma::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j++) h[j] = actf(__uint_as_float(v[j]) + cst[C_HEADS + 32 + 64 * s + j], act);
          umma::tmem_ld32(tl + D_HEADS + 64 + 64 * s, v);
          umma::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j++) h[32 + j] = actf(__uint_as_float(v[j]) + cst[C_HEADS + 64 + 64 * s + j], act);
        }
        const float* W = s == 0 ? P.p1bW : P.p2bW;
        const float* B = s == 0 ? P.p1bb : P.p2bb;
        const uint32_t ia = smem + OFF_ROWIDX + (uint32_t)((s * ROWS + r) * 24);
        for (int k = 0; k < 9; k++) {
          uint32_t idx;
          asm volatile("ld.shared.u16 %0, [%1];" : "=r"(idx) : "r"(ia + 2 * k) : "memory");
          float lg = 0.0f;
          if (idx != 0xFFFFu) {
            const float4* row = reinterpret_cast<const float4*>(W + (size_t)idx * P.pohd);
            float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
            for (int j = 0; j < 16; j++) {
              if (4 * j < P.pohd) {
                const float4 w = __ldg(row + j);
                a0 = fmaf(h[4 * j + 0], w.x, a0);
                a1 = fmaf(h[4 * j + 1], w.y, a1);
                a0 = fmaf(h[4 * j + 2], w.z, a0);
                a1 = fmaf(h[4 * j + 3], w.w, a1);
              }
            }
            lg = (a0 + a1 + __ldg(B + idx)) / out_div;
          }
          if (e < n) logits[(size_t)e * 18 + s * 9 + k] = lg;
        }
      }
      umma::tc_fence_before();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(bar_hfree);
    }
  } else {
    // ===== workers: threads r, r + 128, r + 256, r + 384 share evaluation r of the tile = TMEM lane r; group 0 writes the
    // features, and the groups split every epilogue by accumulator columns (a warp may touch TMEM lanes
    // 32 * (warp % 4) .. + 31)
    const int r = threadIdx.x & (ROWS - 1), grp = threadIdx.x >> 7;
    uint32_t* col = reinterpret_cast<uint32_t*>(dsm + OFF_TILE) + r;
    Mem<ROWS> m;
    m.set(reinterpret_cast<uint8_t*>(col));
    const float* cst = reinterpret_cast<const float*>(dsm + OFF_CONST);
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t dphase = 0, tphase = 0, iter = 0;
    bool t_pending = false;
    // discrete networks (P.act == 3): clamp in the embedding nets, embeddings / hp quantised to 0..127 (mode 1), the int8
    // main net's ClippedReLU on exact integer accumulators (mode 2, act 4), outputs divided by 127 * 64
    const bool disc = P.act == 3;
    const int m_emb = disc ? 1 : 0, m_main = disc ? 2 : 0;
    const float act_bound = (P.act == 2 || disc) ? 1.0f : __int_as_float(0x7F800000);
#ifdef OAKB200_TC_TIMING  // diagnostic build only: where a worker's time goes (thread 0 of CTA 0 prints)
    long long wt[7] = {};  // waiting on MMAs, epilogues, publish, feature build, tile load, heads
    long long wt_c = clock64();
#define WT_LAP(k) do { const long long now_ = clock64(); wt[k] += now_ - wt_c; wt_c = now_; } while (0)
#else
#define WT_LAP(k) do { } while (0)
#endif
    // hand-over to the MMA issuer, one arrival per warp. An operand in Tensor Memory needs the tcgen05 fence only (the stores
    // were waited for in the epilogue); the feature block was written through the generic proxy and is read by the async
    // proxy, hence the proxy fence there.
    auto publish_on = [&](uint32_t which) {
      umma::tc_fence_before();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(which);
      WT_LAP(2);
    };
    auto publish = [&]() {
      umma::tc_fence_before();
      umma::fence_proxy_async();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(bar_a);
      WT_LAP(2);
    };
    uint32_t ephase = 0;
    auto wait_e = [&]() {
      OAKB200_TC_WAIT(bar_e, ephase & 1);
      ephase++;
      umma::tc_fence_after();
      WT_LAP(0);
    };
    auto wait_d = [&]() {
      OAKB200_TC_WAIT(bar_d, dphase & 1);
      dphase++;
      umma::tc_fence_after();
      WT_LAP(0);
    };
    auto wait_t = [&]() {
      if (t_pending) {
        OAKB200_TC_WAIT(bar_t, tphase & 1);
        tphase++;
        t_pending = false;
        umma::tc_fence_after();
        WT_LAP(0);
      }
    };
    auto workers_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(WORKERS) : "memory"); };
#ifdef OAKB200_TC_TRACE
    long long tr_w[96];
    int tr_wn = 0, tr_wtile = 0;
#define TR_W(tag) do { if (blockIdx.x == 0 && threadIdx.x == 0 && tr_wtile == 2 && tr_wn < 96) tr_w[tr_wn++] = (clock64() << 4) | (tag); } while (0)
#else
#define TR_W(tag) do { } while (0)
#endif
    for (uint32_t tb = tile0; tb < ntiles; tb += tile_step) {
#ifdef OAKB200_TC_TRACE
      tr_wtile++;
#endif
      const uint32_t t = tb + crank;
      const uint32_t e = t * ROWS + (uint32_t)r;
      const uint32_t ec = e < n ? e : n - 1;
      workers_sync();  // the other groups have finished with the previous tile
      load_record_part(col, battles + (size_t)ec * 384, grp);
      const uint2 durw = *reinterpret_cast<const uint2*>(durations + (size_t)ec * 8);
      workers_sync();
      WT_LAP(4);
      PokemonFeatures feat;
      // ---- the twelve party slots of the tile as ONE software pipeline: q = 6 s + slot - 1, slot 1 = the active Pokémon of
      // side s. Every layer output stays in Tensor Memory as the next layer's A operand (tmem_operand), so the only shared
      // buffer is the feature block A1, free again as soon as a first layer has run. Per slot q the workers do
      //   wait L1(q) -> E1(q) -> [L2(q)] ; features of q + 1 -> [L1(q + 1)] ; wait L2(q) -> E2(q) -> [T(q)] ; prepare q + 2
      // and the MMA order (= the chunk program, value_net.cu tc_pack) is
      //   L1(0)a L1(0)b | L2(q) L1(q + 1) T(q) ... ; an active slot's first layer comes in two K parts (a: the Active block,
      //   b: its Pokémon block, accumulated), the second part goes out after T(q).
      auto slot_state = [&](int s, int slot, float& hp, float& bound) {
        // hp fraction and activation bound of party slot `slot` (an empty slot contributes a zero block, network.h:131-175)
        const H sb = Mem<ROWS>::sb(s);
        const uint32_t id = m.ld8(sb, S_ORDER + slot - 1);
        PH pk = Mem<ROWS>::pokemon(s, id ? (int)id - 1 : 0);
        const uint32_t hpv = id ? m.ld16(pk, PK_HP) : 0;
        const bool alive = hpv != 0;
        hp = alive ? (float)hpv / (float)m.ld16(pk, PK_STATS) : 0.0f;
        if (disc) hp = floorf(hp * 127.0f);  // network.h:147,165
        bound = alive ? act_bound : 0.0f;
        return alive;
      };
      auto active_features = [&](int s) {  // part a of an active slot's first layer, then its Pokémon block is prepared
        float hp, bound;
        const bool alive = slot_state(s, 1, hp, bound);
        const uint32_t dur = s ? durw.y : durw.x;
        build_active(smem, r, m, s, dur, alive, grp);
        WT_LAP(3);
        publish();
        prepare_pokemon(feat, smem, r, m, s, 1, dur, grp);  // under the Active block's MMAs
        WT_LAP(3);
      };
      auto pokemon_features = [&]() {
#ifndef OAKB200_TC_NO_BUILD  // timing experiment only
        emit_pokemon(feat, smem, r, grp);
#endif
        WT_LAP(3);
        publish();
      };
      active_features(0);
      wait_d();  // L1(0)a
      pokemon_features();  // -> L1(0)b
      prepare_pokemon(feat, smem, r, m, 0, 2, durw.x, grp);
      WT_LAP(3);
      for (int q = 0; q < 12; q++) {
        const int s = q >= 6 ? 1 : 0, slot = q - 6 * s + 1;
        const bool act = slot == 1, last = q == 11, next_act = q == 5;
        TR_W(1);
        wait_d();  // L1(q), for an active slot its second part
        TR_W(2);
        epilogue_inplace<HID_NC>(tl + D_WORK + (uint32_t)(HID_NC * grp), cst + (act ? C_AC0 : C_PK0) + HID_NC * grp, act_bound, 0);
        WT_LAP(1);
        TR_W(3);
        publish_on(bar_a2);  // -> L2(q), A = the hidden layer in Tensor Memory
        TR_W(4);
        if (next_act) active_features(1);       // -> L1(6)a
        else if (!last) pokemon_features();     // -> L1(q + 1), queued behind L2(q)
        TR_W(5);
        float hp, bound;
        slot_state(s, slot, hp, bound);
        TR_W(6);
        wait_e();  // L2(q)
        TR_W(7);
        if (act) epilogue_inplace<ACT_NC>(tl + D_ACT + (uint32_t)(ACT_NC * grp), cst + C_AC1 + ACT_NC * grp, bound, m_emb, P.aod - ACT_NC * grp, hp);
        else epilogue_inplace<EMB_NC>(tl + D_EMB + (uint32_t)(EMB_NC * grp), cst + C_PK1 + EMB_NC * grp, bound, m_emb, P.pod - EMB_NC * grp, hp);
        WT_LAP(1);
        TR_W(8);
        publish_on(bar_ae);  // -> T(q), A = the embedding in Tensor Memory; T(11) commits to bar_t
        TR_W(9);
        if (next_act) {
          wait_d();  // L1(6)a
          pokemon_features();  // -> L1(6)b
          prepare_pokemon(feat, smem, r, m, 1, 2, durw.y, grp);
        } else if (q + 2 < 12 && q + 2 != 6) {  // registers only, under L1(q + 1)
          const int s2 = q + 2 >= 6 ? 1 : 0;
          prepare_pokemon(feat, smem, r, m, s2, q + 2 - 6 * s2 + 1, s2 ? durw.y : durw.x, grp);
        }
        WT_LAP(3);
      }
      t_pending = true;
      // The head warps read the previous tile's head accumulators and row indices while this tile runs; both are rewritten
      // from here on (a fresh barrier passes the parity-1 wait of the first tile).
      umma::mbar_wait(bar_hfree, (iter + 1) & 1);
      WT_LAP(6);
      // Policy heads (policy.h:29-58), worker part, done under the last trunk product: legal choices -> row indices of the
      // last layer. The dot products themselves (one 256-byte weight row per legal choice, gathered from L2) are
      // latency-bound, 15 % of the tile time when the workers did them (profiles/r1o_eval_tc_timing.txt): the head warps
      // run them off the critical path.
      if (grp < 2) {
        const int s = grp;
        const uint32_t res = results[ec];
        const uint64_t list = mem_choices(m, s, (res >> (4 + 2 * s)) & 3);
        const int cnt = choice_count(list);
        const H sb = Mem<ROWS>::sb(s);
        uint32_t idx[12];
#pragma unroll
        for (int k = 0; k < 12; k++) {
          idx[k] = 0xFFFFu;
          if (k < 9 && k < cnt) {
            const uint32_t c = choice_at(list, k);
            const uint32_t type = c & 3, data = c >> 2;
            idx[k] = 0;
            if (type == CH_MOVE) {
              // data == 0 (forced move): the reference reads moves[-1] = the low byte of stats.spc (policy.h:34-36), kept
              const uint32_t mid = m.ld8(m.stored(s), data >= 1 ? PK_MOVES + 2 * ((int)data - 1) : PK_STATS + ST_SPC);
              idx[k] = mid == 0 ? 0 : mid - 1;
            } else if (type == CH_SWITCH) {
              const uint32_t pid = m.ld8(sb, S_ORDER + (int)data - 1);
              idx[k] = 164 + m.ld8(Mem<ROWS>::pokemon(s, (int)pid - 1), PK_SPECIES) - 1;
            }
          }
        }
        const uint32_t ia = smem + OFF_ROWIDX + (uint32_t)((s * ROWS + r) * 24);
#pragma unroll
        for (int k = 0; k < 6; k++) asm volatile("st.shared.u32 [%0], %1;" ::"r"(ia + 4 * k), "r"(idx[2 * k] | (idx[2 * k + 1] << 16)) : "memory");
        if (e < n) counts[(size_t)e * 2 + s] = (uint8_t)cnt;
      }
      WT_LAP(5);
      // ---- trunk (main-net.h:66-107)
      wait_t();
      // both trunk layers' activations become the next layer's A operand where their accumulator was (tmem_operand)
      epilogue_inplace<EMB_NC>(tl + D_TRUNK + (uint32_t)(EMB_NC * grp), cst + C_FC0 + EMB_NC * grp, act_bound, m_main);
      WT_LAP(1);
      publish_on(bar_a);
      wait_d();
      epilogue_inplace<EMB_NC>(tl + D_WORK + (uint32_t)(EMB_NC * grp), cst + C_FC1 + EMB_NC * grp, act_bound, m_main);
      WT_LAP(1);
      publish_on(bar_a);
      wait_d();
      if (grp < 2) {  // head accumulators and row indices of this tile are ready
        umma::tc_fence_before();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bar_h);
      }
      iter++;
      WT_LAP(5);
    }
#ifdef OAKB200_TC_TRACE
    if (blockIdx.x == 0 && threadIdx.x == 0)
      for (int i = 0; i < tr_wn; i++) printf("[trace W] %d %lld\n", (int)(tr_w[i] & 15), tr_w[i] >> 4);
#endif
#ifdef OAKB200_TC_TIMING
    if (blockIdx.x == 0 && threadIdx.x == 0)
      printf("[eval_tc timing] worker 0: waiting on MMAs %lld, epilogues %lld, publish %lld, feature build %lld, tile load %lld, wait for the head warps %lld, row indices %lld cycles\n",
             wt[0], wt[1], wt[2], wt[3], wt[4], wt[6], wt[5]);
#endif
  }
  umma::tc_fence_before();
  __syncthreads();
  if (CL > 1) umma::cluster_sync();  // no CTA leaves while its peer's last commits are still on their way to its barriers
  if (warp == W_STREAM) {
    umma::tc_fence_after();
    umma::tmem_dealloc(tmem, TMEM_COLS);
  }
}

}  // namespace tc
