// One training row from one staged state (see train_rows.cu for the reference citations and the design). Shared by the
// stand-alone rows_encode_kernel (train_rows.cu) and the fused replay_encode_kernel (battle_kernels.cuh).
#pragma once

#include <stdint.h>

#include "../../include/oakb200.h"
#include "gen1_state.cuh"

namespace oakb200 {
namespace rows {

const int POKEMON_DIM = 198, ACTIVE_DIM = 229, POLICY_DIM = 315;
const int ROW_POKEMON = 2 * 6 * POKEMON_DIM;  // 2376 floats = 9504 B: every row of `pokemon` is 16-byte aligned
const int ROW_ACTIVE = 2 * ACTIVE_DIM;        // 458 floats = 1832 B: rows of `active` are 8-byte aligned only
const int STATE_BYTES = 432;  // staged state (battle_kernels.cuh: ROW_STATE_BYTES)

__device__ __forceinline__ int dense_status_index(uint32_t status, uint32_t sleeps) {  // battle.h:106-123
  if ((status & 7u) == 0) return __ffs(status) - 1 - 3;
  if (!(status & 0x80u)) return 3 + (int)sleeps;
  return 14 - (int)(status & 7u);
}
__device__ __forceinline__ uint32_t ld16(const uint8_t* p) { return p[0] | (p[1] << 8); }

__device__ __forceinline__ uint32_t row_policy_index(const uint8_t* side, uint32_t choice) {  // policy.h:29-58
  const uint32_t type = choice & 3u, data = choice >> 2;
  const uint8_t* stored = side + (side[S_ORDER] - 1) * PK_SIZE;
  if (type == 1) {
    // data == 0 (the forced move of recharge / locked-in turns): the reference indexes moves[-1], the two bytes before
    // Pokemon::moves, i.e. the low byte of stats.spc — its compiled behaviour, kept
    const uint32_t id = data >= 1 ? stored[PK_MOVES + 2 * (data - 1)] : stored[PK_STATS + ST_SPC];
    return id == 0 ? 0 : id - 1;
  }
  if (type == 2) return 164 + side[(side[S_ORDER + data - 1] - 1) * PK_SIZE + PK_SPECIES] - 1;
  return 0;
}

// Entities of a row: e = 0..11 the party slots (side e / 6, slot e % 6), e = 12, 13 the two active Pokemon. An entity
// that is absent or fainted leaves its block zero (encoded-frames.h:73-99).
// Whole warp; `st` = this row's staged state in shared memory, already visible to every lane.
__device__ __forceinline__ void encode_row(const uint8_t* st, const uint8_t* __restrict__ stream, const oakb200_training_rows& out,
                                           uint32_t row, int lane) {
  float* pk_g = out.pokemon ? out.pokemon + (size_t)row * ROW_POKEMON : nullptr;
  float* ac_g = out.active ? out.active + (size_t)row * ROW_ACTIVE : nullptr;
  // ---- clear the row
  if (pk_g) {
    uint4* z = reinterpret_cast<uint4*>(pk_g);
#pragma unroll 4
    for (int i = lane; i < ROW_POKEMON / 4; i += 32) z[i] = make_uint4(0, 0, 0, 0);
  }
  if (ac_g) {
    uint2* z = reinterpret_cast<uint2*>(ac_g);
#pragma unroll 4
    for (int i = lane; i < ROW_ACTIVE / 2; i += 32) z[i] = make_uint2(0, 0);
  }
  __syncwarp();  // the state is in shared memory; the zeros are ordered before the stores below
  const uint32_t d0 = *reinterpret_cast<const uint32_t*>(st + 384), d1 = *reinterpret_cast<const uint32_t*>(st + 388);

  // ---- entity table: lane e holds {valid, byte offset of the Pokemon / ActivePokemon in the state}
  uint32_t ent = 0;
  if (lane < 12) {
    const uint8_t* side = st + (lane / 6) * SIDE_SIZE;
    const uint32_t id = side[S_ORDER + lane % 6];
    if (id != 0) {
      const uint32_t off = (lane / 6) * SIDE_SIZE + (id - 1) * PK_SIZE;
      ent = off | (ld16(st + off + PK_HP) != 0 ? 0x80000000u : 0u);
    }
  } else if (lane < 14) {
    const uint8_t* side = st + (lane - 12) * SIDE_SIZE;
    const uint32_t hp = ld16(side + (side[S_ORDER] - 1) * PK_SIZE + PK_HP);
    ent = ((lane - 12) * SIDE_SIZE + S_ACTIVE) | (hp != 0 ? 0x80000000u : 0u);
  }
  const uint32_t valid = __ballot_sync(0xFFFFFFFFu, ent >> 31);
  // hp (encoded-frames.h:77,94)
  if (lane < 12 && out.hp) {
    const uint32_t off = ent & 0xFFFFu;
    out.hp[(size_t)row * 12 + lane] = (ent >> 31) ? (float)ld16(st + off + PK_HP) / (float)ld16(st + off) : 0.0f;
  }
  // ---- stats: 14 entities x 5 (battle.h:22-29)
#pragma unroll
  for (int r = 0; r < 3; r++) {
    const int t = lane + 32 * r, e = min(t / 5, 13), i = t - 5 * e;
    const uint32_t off = __shfl_sync(0xFFFFFFFFu, ent, e) & 0xFFFFu;
    if (t < 70 && ((valid >> e) & 1u)) {
      const float v = (float)ld16(st + off + 2 * i) / (i == 0 ? 703.0f : 999.0f);
      float* dst = e < 12 ? pk_g + e * POKEMON_DIM : ac_g + (e - 12) * ACTIVE_DIM;
      if (e < 12 ? pk_g != nullptr : ac_g != nullptr) dst[i] = v;
    }
  }
  // ---- move slots: 14 entities x 4 (battle.h:64-71); slots are written in order there, so among equal ids (Mimic
  // can duplicate a move) the last slot decides — and a slot out of pp leaves the zero
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const int t = lane + 32 * r, e = min(t >> 2, 13), i = t & 3;
    const uint32_t off = (__shfl_sync(0xFFFFFFFFu, ent, e) & 0xFFFFu) + (e < 12 ? PK_MOVES : 24);
    if (t < 56 && ((valid >> e) & 1u)) {
      const uint32_t id = st[off + 2 * i], pp = st[off + 2 * i + 1];
      bool later = false;
      for (int j = i + 1; j < 4; j++) later |= st[off + 2 * j] == id;
      if (id != MOVE_None && id != MOVE_Struggle && pp != 0 && !later) {
        float* dst = e < 12 ? pk_g + e * POKEMON_DIM + 5 : ac_g + (e - 12) * ACTIVE_DIM + 45;
        if (e < 12 ? pk_g != nullptr : ac_g != nullptr) dst[id - 1] = 1.0f;
      }
    }
  }
  // ---- types (battle.h:171-178) and, for party slots, status (battle.h:146-151)
  if (lane < 14 && (ent >> 31)) {
    const uint32_t off = ent & 0xFFFFu;
    const uint32_t types = st[off + (lane < 12 ? PK_TYPES : 11)];
    float* dst = lane < 12 ? pk_g + lane * POKEMON_DIM + 183 : ac_g + (lane - 12) * ACTIVE_DIM + 5;
    if (lane < 12 ? pk_g != nullptr : ac_g != nullptr) {
      dst[types & 15u] = 1.0f;
      dst[types >> 4] = 1.0f;
      if (lane < 12) {
        const uint32_t status = st[off + PK_STATUS];
        const uint32_t sleep = ((lane >= 6 ? d1 : d0) >> (3 * (lane % 6))) & 7u;
        if (status) pk_g[lane * POKEMON_DIM + 169 + dense_status_index(status, sleep)] = 1.0f;
      }
    }
  }
  // ---- active: the 16 volatile flags of both sides (battle.h:297-316; multi-hit and flinch are not encoded)
  if (ac_g) {
    const int s = lane >> 4, j = lane & 15;
    const uint8_t* a = st + s * SIDE_SIZE + S_ACTIVE;
    const uint32_t vlo = a[16] | (a[17] << 8) | (a[18] << 16);
    if (((valid >> (12 + s)) & 1u) && ((vlo >> (j < 2 ? j : j + 2)) & 1u)) ac_g[s * ACTIVE_DIM + 26 + j] = 1.0f;
    // ---- active: boosts (lanes 0..11), volatile fields (12..17), durations (18..25)
    if (lane < 12) {
      const int sb = lane / 6, k = lane % 6;
      const uint8_t* b = st + sb * SIDE_SIZE + S_ACTIVE + 12;
      const int stage = boost_get(b[0] | (b[1] << 8) | (b[2] << 16), k);
      if ((valid >> (12 + sb)) & 1u)  // battle.h:251-264
        ac_g[sb * ACTIVE_DIM + 20 + k] = ((float)boost_num(stage) / (float)boost_den(stage)) * (k < 4 ? 0.25f : (float)(1 / 3.0));
    } else if (lane < 18) {
      const int sv = (lane - 12) / 3, w = (lane - 12) % 3;
      const uint8_t* v = st + sv * SIDE_SIZE + S_ACTIVE + 16;
      float val;  // battle.h:319-325: state / 65535, substitute hp / (706 / 4 + 1), toxic counter / 16
      if (w == 0) val = (float)(v[3] | (v[4] << 8)) / 65535.0f;
      else if (w == 1) val = (float)v[5] / 177.0f;
      else val = (float)((double)(v[7] >> 3) / 16.0);
      if (((valid >> (12 + sv)) & 1u) && val != 0.0f) ac_g[sv * ACTIVE_DIM + 42 + w] = val;
    } else if (lane < 26) {
      const int sd = (lane - 18) >> 2, w = (lane - 18) & 3;
      const uint32_t dur = sd ? d1 : d0;  // battle.h:375-400: confusion 5, disable 8, attacking 3, binding 4
      const uint32_t n = w == 0 ? bf_get(dur, DF_CONFUSION) : w == 1 ? bf_get(dur, DF_DISABLE) : w == 2 ? bf_get(dur, DF_ATTACKING)
                                                                                                    : bf_get(dur, DF_BINDING);
      const int base = w == 0 ? 209 : w == 1 ? 214 : w == 2 ? 222 : 225;
      if (((valid >> (12 + sd)) & 1u) && n != 0) ac_g[sd * ACTIVE_DIM + base + n - 1] = 1.0f;
      // battle.h:472-475 (the disabled move's entry is cleared AFTER t moved past the move block): that store lands
      // at or beyond this Duration block on a zero that later writes fill in — no effect on a cleared buffer, none here
    }
  }

  // ---- the targets, straight from the stored update (Update::write_to_tensor) and the staged choices
  const uint8_t* upd = stream + *reinterpret_cast<const uint32_t*>(st + 396);
  const uint32_t hdr = upd[0], m = (hdr & 15u) + 1, n = (hdr >> 4) + 1;
  if (lane < 18) {
    const int s = lane / 9, i = lane % 9;
    const uint32_t cnt = s ? n : m;
    const bool on = (uint32_t)i < cnt;
    if (out.choice_indices) {
      uint32_t idx = POLICY_DIM;
      if (on && (uint32_t)i < st[394 + s]) idx = row_policy_index(st + s * SIDE_SIZE, st[400 + 9 * s + i]);
      out.choice_indices[(size_t)row * 18 + lane] = (int64_t)idx;
    }
    const uint8_t* pe = upd + 11 + (s ? 4 * m : 0) + 2 * i;  // p1 empirical, p1 nash, p2 empirical, p2 nash
    const uint8_t* pn = pe + 2 * cnt;
    if (out.empirical_policies) out.empirical_policies[(size_t)row * 18 + lane] = on ? (float)ld16(pe) / 65535.0f : 0.0f;
    if (out.nash_policies) out.nash_policies[(size_t)row * 18 + lane] = on ? (float)ld16(pn) / 65535.0f : 0.0f;
  } else if (lane == 18) {
    if (out.k) {
      out.k[(size_t)row * 2] = (uint8_t)m;
      out.k[(size_t)row * 2 + 1] = (uint8_t)n;
    }
    if (out.choice) {
      out.choice[(size_t)row * 2] = upd[1];
      out.choice[(size_t)row * 2 + 1] = upd[2];
    }
  } else if (lane == 19) {
    if (out.iterations) out.iterations[row] = upd[3] | (upd[4] << 8) | (upd[5] << 16) | ((uint32_t)upd[6] << 24);
    if (out.empirical_value) out.empirical_value[row] = (float)ld16(upd + 7) / 65535.0f;
    if (out.nash_value) out.nash_value[row] = (float)ld16(upd + 9) / 65535.0f;
    const uint32_t t = st[393] & 15u;
    if (out.score) out.score[row] = t == RES_WIN ? 1.0f : (t == RES_LOSE ? 0.0f : 0.5f);
  }
  // ---- the uncompressed frame (Py::Battle::Frames::write, py/battle/frames.h:39-51): the state the row was built from
  if (out.battle) {
    uint8_t* dst = out.battle + (size_t)row * 384;
    if ((reinterpret_cast<uintptr_t>(out.battle) & 15u) == 0) {
      if (lane < 24) reinterpret_cast<uint4*>(dst)[lane] = reinterpret_cast<const uint4*>(st)[lane];
    } else {
      for (int i = lane; i < 384; i += 32) dst[i] = st[i];
    }
  }
  if (out.durations && lane < 8) out.durations[(size_t)row * 8 + lane] = st[384 + lane];
  if (out.result && lane == 8) out.result[row] = st[392];
  if (out.legal_choices && lane >= 9 && lane < 27) {
    const int j = lane - 9, s = j / 9, i = j % 9;
    out.legal_choices[(size_t)row * 18 + j] = (uint32_t)i < st[394 + s] ? st[400 + j] : (uint8_t)0;
  }
}

}  // namespace rows
}  // namespace oakb200
