// Battle-state access for the B200 engine.
//
// HBM layout: the reference's own records, untouched — battle[n][384] (libpkmn Battle, reference
// engine/src/lib/gen1/data.zig:54-68), durations[n][8] (gen1/chance.zig:255-318), result[n] (u8).
// On-chip layout: one battle per thread, its 96 32-bit words stored as a COLUMN of a word-transposed
// shared-memory tile, word w of the battle owned by thread t at tile[w * STRIDE + t]. Thread t
// therefore always hits bank (t mod 32) whatever field it touches, so the data-dependent field
// addressing of the simulator (party slot, move slot, player) is bank-conflict-free by construction.
//
// Everything here is `OAK_HD`: under nvcc it is __device__ (device only), under a plain host compiler it is
// ordinary inline C++ with STRIDE = 1 so that tests can differential-test the very same engine source
// against the oracle without a GPU (tests only; the product has no CPU path).
#pragma once

#include <stdint.h>

#ifdef __CUDACC__
#define OAK_HD __device__ __forceinline__
#define OAK_HD_NOINLINE __device__ __noinline__
#else
#define OAK_HD inline
#define OAK_HD_NOINLINE inline
#endif

#include "gen1_tables.cuh"

namespace oakb200 {

// ---- byte offsets inside the 384-byte battle (reference cpp/include/libpkmn/layout.h:17-96)
enum : int {
  SIDE_SIZE = 184,
  PK_SIZE = 24,
  PK_STATS = 0,  // hp, atk, def, spe, spc (u16 each)
  PK_MOVES = 10,  // 4 x {id, pp}
  PK_HP = 18,
  PK_STATUS = 20,
  PK_SPECIES = 21,
  PK_TYPES = 22,
  PK_LEVEL = 23,
  S_ACTIVE = 144,
  A_STATS = S_ACTIVE + 0,
  A_SPECIES = S_ACTIVE + 10,
  A_TYPES = S_ACTIVE + 11,
  A_BOOSTS = S_ACTIVE + 12,
  A_VOL = S_ACTIVE + 16,
  A_MOVES = S_ACTIVE + 24,
  S_ORDER = 176,
  S_LAST_SELECTED = 182,
  S_LAST_USED = 183,
  B_TURN = 368,
  B_LAST_DAMAGE = 370,
  B_LAST_MOVES = 372,  // per player {index u8, counterable u8}
  B_RNG = 376,
  ST_HP = 0, ST_ATK = 2, ST_DEF = 4, ST_SPE = 6, ST_SPC = 8,
};

enum : uint8_t { CH_PASS = 0, CH_MOVE = 1, CH_SWITCH = 2 };
enum : uint8_t { RES_NONE = 0, RES_WIN = 1, RES_LOSE = 2, RES_TIE = 3, RES_ERROR = 4, RESULT_DEFAULT = 0x50 };

// Status byte (data.zig:235-301)
enum : uint8_t { SB_PSN = 1u << 3, SB_BRN = 1u << 4, SB_FRZ = 1u << 5, SB_PAR = 1u << 6, SB_EXT = 1u << 7, SB_TOX = 0x88 };

// Volatiles (data.zig:325-383), as masks on the packed u64
#define OAK_VBIT(n) (1ull << (n))
static const uint64_t V_BIDE = OAK_VBIT(0), V_THRASHING = OAK_VBIT(1), V_MULTIHIT = OAK_VBIT(2), V_FLINCH = OAK_VBIT(3),
                      V_CHARGING = OAK_VBIT(4), V_BINDING = OAK_VBIT(5), V_INVULNERABLE = OAK_VBIT(6),
                      V_CONFUSION = OAK_VBIT(7), V_MIST = OAK_VBIT(8), V_FOCUS_ENERGY = OAK_VBIT(9),
                      V_SUBSTITUTE = OAK_VBIT(10), V_RECHARGING = OAK_VBIT(11), V_RAGE = OAK_VBIT(12),
                      V_LEECH_SEED = OAK_VBIT(13), V_TOXIC = OAK_VBIT(14), V_LIGHT_SCREEN = OAK_VBIT(15),
                      V_REFLECT = OAK_VBIT(16), V_TRANSFORM = OAK_VBIT(17);
// multi-bit fields: shift, width
#define VF_CONFUSION 18, 3
#define VF_ATTACKS 21, 3
#define VF_STATE 24, 16
#define VF_SUBSTITUTE 40, 8
#define VF_TRANSFORM 48, 4
#define VF_DISABLE_DURATION 52, 4
#define VF_DISABLE_MOVE 56, 3
#define VF_TOXIC 59, 5

OAK_HD uint32_t vf_get(uint64_t v, int shift, int bits) { return (uint32_t)((v >> shift) & ((1ull << bits) - 1)); }
OAK_HD uint64_t vf_set(uint64_t v, int shift, int bits, uint32_t x) {
  const uint64_t m = ((1ull << bits) - 1) << shift;
  return (v & ~m) | (((uint64_t)x << shift) & m);
}

// Boost nibbles (data.zig:448-463)
enum : int { BO_ATK = 0, BO_DEF = 1, BO_SPE = 2, BO_SPC = 3, BO_ACC = 4, BO_EVA = 5 };
OAK_HD int boost_get(uint32_t b, int which) {
  const int v = (int)((b >> (4 * which)) & 15u);
  return v >= 8 ? v - 16 : v;
}
OAK_HD uint32_t boost_set(uint32_t b, int which, int v) {
  return (b & ~(15u << (4 * which))) | (((uint32_t)v & 15u) << (4 * which));
}

// chance Action fields (chance.zig:160-206): shift, width
#define AF_DAMAGE 0, 8
#define AF_HIT 8, 2
#define AF_CRIT 10, 2
#define AF_SECONDARY 12, 2
#define AF_SPEED_TIE 14, 2
#define AF_CONFUSED 16, 2
#define AF_PARALYZED 18, 2
#define AF_DURATION 20, 4
#define AF_SLEEP 24, 3
#define AF_CONFUSION 27, 4
#define AF_DISABLE 31, 3
#define AF_ATTACKING 34, 2
#define AF_BINDING 36, 2
#define AF_MOVE_SLOT 38, 3
#define AF_MULTI_HIT 41, 3
#define AF_PP 44, 4
#define AF_PSYWAVE 48, 8
#define AF_METRONOME 56, 8
enum : uint32_t { OPT_NONE = 0, OPT_FALSE = 1, OPT_TRUE = 2 };
enum : uint32_t { OBS_NONE = 0, OBS_STARTED = 1, OBS_CONTINUING = 2, OBS_ENDED = 3, OBS_OVERWRITTEN = 4 };

// chance Duration fields (chance.zig:279-287): sleeps 6x3 bits, then these
#define DF_CONFUSION 18, 3
#define DF_DISABLE 21, 4
#define DF_ATTACKING 25, 3
#define DF_BINDING 28, 3
OAK_HD uint32_t bf_get(uint32_t d, int shift, int bits) { return (d >> shift) & ((1u << bits) - 1); }
OAK_HD uint32_t bf_set(uint32_t d, int shift, int bits, uint32_t x) {
  const uint32_t m = ((1u << bits) - 1) << shift;
  return (d & ~m) | ((x << shift) & m);
}

// ---- constant tables (generated; reference moves.zig / species.zig / types.zig)
OAK_HD uint32_t move_word(uint32_t id) { return GEN1_MOVE_DATA[id]; }
OAK_HD uint32_t mv_effect(uint32_t w) { return w & 255u; }
OAK_HD uint32_t mv_bp(uint32_t w) { return (w >> 8) & 255u; }
OAK_HD uint32_t mv_accuracy(uint32_t w) { return (w >> 16) & 255u; }
OAK_HD uint32_t mv_type(uint32_t w) { return (w >> 24) & 15u; }
OAK_HD bool mv_self(uint32_t w) { return ((w >> 28) & 1u) != 0; }
OAK_HD uint32_t type_eff(uint32_t atk, uint32_t def) { return GEN1_TYPE_CHART[atk * 15 + def]; }
OAK_HD bool types_include(uint32_t types, uint32_t t) { return (types & 15u) == t || (types >> 4) == t; }

// Effect ordinal ranges (moves.zig:1600-1640)
OAK_HD bool eff_onBegin(uint32_t e) { return e - 1u < 16u; }
OAK_HD bool eff_isStatDown(uint32_t e) { return e - 17u < 5u; }
OAK_HD bool eff_onEnd(uint32_t e) { return e - 17u < 15u; }
OAK_HD bool eff_alwaysHappens(uint32_t e) { return e - 32u < 7u; }
OAK_HD bool eff_isSpecial(uint32_t e) { return e - 32u < 15u; }
OAK_HD bool eff_isMulti(uint32_t e) { return e - 45u < 3u; }
OAK_HD bool eff_isStatDownChance(uint32_t e) { return e - 48u < 4u; }

// Stat-stage multipliers num/den (mechanics.zig:34-48), index = stage + 6
OAK_HD uint32_t boost_num(int stage) {
  // 25 28 33 40 50 66 1 15 2 25 3 35 4 packed 8 bits each would not fit one word; two words
  const uint64_t lo = 0x0F01423228211C19ull;  // stages -6..+1: 25,28,33,40,50,66,1,15
  const uint64_t hi = 0x0000000423031902ull;  // stages +2..+6: 2,25,3,35,4
  const int i = stage + 6;
  return (uint32_t)(((i < 8 ? lo >> (8 * i) : hi >> (8 * (i - 8)))) & 255u);
}
OAK_HD uint32_t boost_den(int stage) {
  // 100 x6, 1, 10, 1, 10, 1, 10, 1
  return stage < 0 ? 100u : ((stage & 1) ? 10u : 1u);
}
// x * num / den of a stage with the three possible denominators as CONSTANT divisions (a multiply-high each) and a select,
// instead of a division by a run-time value (~30 instructions, 5 % of the playout kernel's instructions when it was one)
OAK_HD uint32_t boost_scale(uint32_t x, int stage) {
  const uint32_t n = x * boost_num(stage);
  const uint32_t q100 = n / 100u, q10 = n / 10u;
  return stage < 0 ? q100 : ((stage & 1) ? q10 : n);
}

// volatile flag masks on the LOW word of the packed u64 (bits 0..17), fields confusion/attacks
enum : uint32_t {
  VL_BIDE = 1u << 0, VL_THRASHING = 1u << 1, VL_MULTIHIT = 1u << 2, VL_FLINCH = 1u << 3, VL_CHARGING = 1u << 4,
  VL_BINDING = 1u << 5, VL_INVULNERABLE = 1u << 6, VL_CONFUSION = 1u << 7, VL_MIST = 1u << 8,
  VL_FOCUS_ENERGY = 1u << 9, VL_SUBSTITUTE = 1u << 10, VL_RECHARGING = 1u << 11, VL_RAGE = 1u << 12,
  VL_LEECH_SEED = 1u << 13, VL_TOXIC = 1u << 14, VL_LIGHT_SCREEN = 1u << 15, VL_REFLECT = 1u << 16,
  VL_TRANSFORM = 1u << 17,
};
// (shift, width) of fields inside the low / high volatile words
#define VLO_CONFUSION 18, 3
#define VLO_ATTACKS 21, 3
#define VHI_SUBSTITUTE 8, 8
#define VHI_TRANSFORM 16, 4
#define VHI_DISABLE_DURATION 20, 4
#define VHI_DISABLE_MOVE 24, 3
#define VHI_TOXIC 27, 5

// ---- the per-battle memory window
// H is a WORD-ALIGNED position inside the battle record, already scaled to column addressing
// (word index * STRIDE * 4). Fields are then addressed as (H, constant byte offset): the offset folds
// into the load/store immediate, so a field access costs one LDS/STS and no address arithmetic.
// (The first B200 build addressed everything by runtime byte offset and spent ~40% of its
// instructions on shift/mask/multiply address math.)
struct H {
  int v;
};
// Handle of one party Pokémon: v = its position in the on-chip column (scaled like H), c = the byte offset of its 24-byte
// record inside the 384-byte battle (only the SPLIT layout reads it).
struct PH {
  int v, c;
};

// SPLIT layout: what an update can never write stays where it is — in the battle's own record in global memory — and only
// the rest is staged on chip. Per party Pokémon (24 B: data.zig:185-203) the stored stats, move ids, species, types and
// level are fixed for the whole battle (Transform, Mimic, Conversion and stat changes write the ACTIVE block only,
// mechanics.zig:2156-2179,2438-2465,2692-2696); PP, HP and status (7 B) change. On chip a battle is then 48 words instead of 96:
//   per side 22 words: 6 x {word 0: pp[4], word 1: hp lo, hp hi, status, -} | active 8 words | order[6], last_selected, last_used
//   tail 4 words: turn, last_damage | last_moves | rng
// which doubles the battles (warps) an SM holds. The constant fields are read through the read-only path (ld.global.nc):
// they hit L1/L2, the record was fetched when the column was filled.
template <bool SPLIT>
struct ConstRecord {};
template <>
struct ConstRecord<true> {
  const uint8_t* crec;  // the battle's 384-byte record (global memory on the device)
};

template <int STRIDE, bool SPLIT = false>
struct Mem : ConstRecord<SPLIT> {
  static const int W = STRIDE * 4;
  static const bool IS_SPLIT = SPLIT;
  static const int WORDS = SPLIT ? 48 : 96;           // words per battle on chip
  static const int SIDE_WORDS = SPLIT ? 22 : 46;      // per side
  static const int PK_WORDS = SPLIT ? 2 : 6;          // per party Pokémon
  static const int ACTIVE_WORD = SPLIT ? 12 : 36;     // word of `active` inside a side (S_ACTIVE / 4 in the record)
  static const int TAIL_WORD = SPLIT ? 44 : 92;       // word of `turn` (B_TURN / 4 in the record)
  static OAK_HD int enc(int off) { return (off >> 2) * W + (off & 3); }

#ifdef __CUDA_ARCH__
  // Device: the column lives in shared memory; keep its 32-bit shared-window address and issue
  // explicit ld.shared / st.shared so that non-inlined routines (which cannot see where the pointer
  // came from) still get LDS/STS with the field offset folded in, never generic LD/ST.
  uint32_t sbase;
  OAK_HD void set(uint8_t* column) { sbase = (uint32_t)__cvta_generic_to_shared(column); }
  OAK_HD uint32_t ld8a(uint32_t a) const { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
  OAK_HD uint32_t ld16a(uint32_t a) const { uint32_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
  OAK_HD uint32_t ld32a(uint32_t a) const { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
  OAK_HD void st8a(uint32_t a, uint32_t v) const { asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
  OAK_HD void st16a(uint32_t a, uint32_t v) const { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
  OAK_HD void st32a(uint32_t a, uint32_t v) const { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
  // raw battle offsets: battle-level fields (B_*) in either layout; anything below B_TURN only in the full layout
  OAK_HD uint32_t at(int off) const { return sbase + (uint32_t)(enc(off) + (TAIL_WORD - B_TURN / 4) * W); }
  OAK_HD uint32_t at(H h, int off) const { return sbase + (uint32_t)(h.v + enc(off)); }
  OAK_HD uint32_t atv(int v) const { return sbase + (uint32_t)v; }
  OAK_HD uint32_t cld8(int off) const { return SPLIT ? (uint32_t)__ldg(crec_ptr() + off) : 0u; }
  OAK_HD uint32_t cld16(int off) const { return SPLIT ? (uint32_t)__ldg(reinterpret_cast<const uint16_t*>(crec_ptr() + off)) : 0u; }
  OAK_HD uint2 cld64(int off) const { return SPLIT ? __ldg(reinterpret_cast<const uint2*>(crec_ptr() + off)) : make_uint2(0u, 0u); }  // off % 8 == 0
#else
  uint8_t* base;  // host (tests only): address of byte 0 of word 0 of this battle's column
  OAK_HD void set(uint8_t* column) { base = column; }
  OAK_HD uint32_t ld8a(uint8_t* a) const { return *a; }
  OAK_HD uint32_t ld16a(uint8_t* a) const { return *reinterpret_cast<const uint16_t*>(a); }
  OAK_HD uint32_t ld32a(uint8_t* a) const { return *reinterpret_cast<const uint32_t*>(a); }
  OAK_HD void st8a(uint8_t* a, uint32_t v) const { *a = (uint8_t)v; }
  OAK_HD void st16a(uint8_t* a, uint32_t v) const { *reinterpret_cast<uint16_t*>(a) = (uint16_t)v; }
  OAK_HD void st32a(uint8_t* a, uint32_t v) const { *reinterpret_cast<uint32_t*>(a) = v; }
  OAK_HD uint8_t* at(int off) const { return base + (enc(off) + (TAIL_WORD - B_TURN / 4) * W); }
  OAK_HD uint8_t* at(H h, int off) const { return base + (h.v + enc(off)); }
  OAK_HD uint8_t* atv(int v) const { return base + v; }
  OAK_HD uint32_t cld8(int off) const { return SPLIT ? (uint32_t)crec_ptr()[off] : 0u; }
  OAK_HD uint32_t cld16(int off) const { return SPLIT ? (uint32_t)crec_ptr()[off] | ((uint32_t)crec_ptr()[off + 1] << 8) : 0u; }
#endif
  OAK_HD const uint8_t* crec_of(const ConstRecord<true>& c) const { return c.crec; }
  OAK_HD const uint8_t* crec_of(const ConstRecord<false>&) const { return nullptr; }
  OAK_HD const uint8_t* crec_ptr() const { return crec_of(*this); }
  OAK_HD void set_record(ConstRecord<true>& c, const uint8_t* record) { c.crec = record; }
  OAK_HD void set_record(ConstRecord<false>&, const uint8_t*) {}
  OAK_HD void set_const(const uint8_t* record) { set_record(*this, record); }
  // raw byte offset from the start of the battle (battle-level fields, cold paths)
  OAK_HD uint32_t ld8(int off) const { return ld8a(at(off)); }
  OAK_HD void st8(int off, uint32_t v) const { st8a(at(off), v); }
  OAK_HD uint32_t ld16(int off) const { return ld16a(at(off)); }
  OAK_HD void st16(int off, uint32_t v) const { st16a(at(off), v); }
  OAK_HD uint32_t ld32(int off) const { return ld32a(at(off)); }
  OAK_HD void st32(int off, uint32_t v) const { st32a(at(off), v); }
  OAK_HD uint64_t ld64(int off) const { return (uint64_t)ld32(off) | ((uint64_t)ld32(off + 4) << 32); }
  OAK_HD void st64(int off, uint64_t v) const {
    st32(off, (uint32_t)v);
    st32(off + 4, (uint32_t)(v >> 32));
  }
  // (aligned handle, byte offset) addressing
  OAK_HD uint32_t ld8(H h, int off) const { return ld8a(at(h, off)); }
  OAK_HD void st8(H h, int off, uint32_t v) const { st8a(at(h, off), v); }
  OAK_HD uint32_t ld16(H h, int off) const { return ld16a(at(h, off)); }
  OAK_HD void st16(H h, int off, uint32_t v) const { st16a(at(h, off), v); }
  OAK_HD uint32_t ld32(H h, int off) const { return ld32a(at(h, off)); }
  OAK_HD void st32(H h, int off, uint32_t v) const { st32a(at(h, off), v); }

  // ---- handles
  // side p: a handle such that (sb(p), A_* / S_* offset) lands on the side's active / order / last-move words in either layout
  static OAK_HD H sb(int p) { return H{(p * SIDE_WORDS + ACTIVE_WORD - S_ACTIVE / 4) * W}; }
  static OAK_HD PH pokemon(int p, int idx) { return PH{(p * SIDE_WORDS + idx * PK_WORDS) * W, p * SIDE_SIZE + idx * PK_SIZE}; }  // party index 0..5
  static OAK_HD H hadd(H h, int aligned_off) { return H{h.v + (aligned_off >> 2) * W}; }   // + word-aligned byte offset
  OAK_HD PH stored(int p) const { return pokemon(p, (int)ld8(sb(p), S_ORDER) - 1); }

  // ---- party Pokémon fields: (handle, PK_* byte offset). With a compile-time offset the layout test folds away.
  static OAK_HD bool pk_is_const(int off) { return off < PK_MOVES || off > PK_STATUS || (off < PK_HP && !(off & 1)); }
  static OAK_HD int pk_dyn(int off) {  // position of a dynamic byte inside the two on-chip words of the SPLIT layout
    return off < PK_HP ? ((off - PK_MOVES) >> 1) : W + (off - PK_HP);
  }
  OAK_HD uint32_t ld8(PH h, int off) const {
    if (!SPLIT) return ld8a(at(H{h.v}, off));
    return pk_is_const(off) ? cld8(h.c + off) : ld8a(atv(h.v + pk_dyn(off)));
  }
  OAK_HD void st8(PH h, int off, uint32_t v) const {  // PP and status are the only bytes ever written
    if (!SPLIT) st8a(at(H{h.v}, off), v);
    else st8a(atv(h.v + pk_dyn(off)), v);
  }
  OAK_HD uint32_t ld16(PH h, int off) const {  // stats (constant), hp, or a {move id, pp} pair
    if (!SPLIT) return ld16a(at(H{h.v}, off));
    if (off == PK_HP) return ld16a(atv(h.v + W));
    if (off < PK_MOVES) return cld16(h.c + off);
    return cld8(h.c + off) | (ld8a(atv(h.v + pk_dyn(off + 1))) << 8);
  }
  // run-time indexed fields: PP of move slot k (0..3), stat at byte offset st (ST_*) of the stored stats
  OAK_HD uint32_t pp_at(PH h, int k) const { return SPLIT ? ld8a(atv(h.v + k)) : ld8a(at(H{h.v}, PK_MOVES + 2 * k + 1)); }
  OAK_HD void set_pp_at(PH h, int k, uint32_t v) const {
    if (SPLIT) st8a(atv(h.v + k), v);
    else st8a(at(H{h.v}, PK_MOVES + 2 * k + 1), v);
  }
  OAK_HD uint32_t stat_at(PH h, int st) const { return SPLIT ? cld16(h.c + PK_STATS + st) : ld16a(at(H{h.v}, PK_STATS + st)); }
  OAK_HD void st16(PH h, int off, uint32_t v) const {  // hp is the only 16-bit field ever written
    if (!SPLIT) st16a(at(H{h.v}, off), v);
    else st16a(atv(h.v + W), v);
  }

  // ---- field helpers
  OAK_HD uint32_t vlo(int p) const { return ld32(sb(p), A_VOL); }
  OAK_HD void set_vlo(int p, uint32_t v) const { st32(sb(p), A_VOL, v); }
  OAK_HD uint32_t vhi(int p) const { return ld32(sb(p), A_VOL + 4); }
  OAK_HD void set_vhi(int p, uint32_t v) const { st32(sb(p), A_VOL + 4, v); }
  OAK_HD bool vflag(int p, uint32_t mask) const { return (vlo(p) & mask) != 0; }
  OAK_HD void vclear(int p, uint32_t mask) const { set_vlo(p, vlo(p) & ~mask); }
  OAK_HD void vraise(int p, uint32_t mask) const { set_vlo(p, vlo(p) | mask); }
  OAK_HD uint32_t vlo_field(int p, int shift, int bits) const { return bf_get(vlo(p), shift, bits); }
  OAK_HD void set_vlo_field(int p, int shift, int bits, uint32_t x) const { set_vlo(p, bf_set(vlo(p), shift, bits, x)); }
  OAK_HD uint32_t vhi_field(int p, int shift, int bits) const { return bf_get(vhi(p), shift, bits); }
  OAK_HD void set_vhi_field(int p, int shift, int bits, uint32_t x) const { set_vhi(p, bf_set(vhi(p), shift, bits, x)); }
  OAK_HD uint32_t vstate(int p) const { return (vlo(p) >> 24) | ((vhi(p) & 255u) << 8); }
  OAK_HD void set_vstate(int p, uint32_t s) const {
    set_vlo(p, (vlo(p) & 0x00FFFFFFu) | (s << 24));
    set_vhi(p, (vhi(p) & ~255u) | ((s >> 8) & 255u));
  }
  OAK_HD uint32_t status_of(PH pk) const { return ld8(pk, PK_STATUS); }
  OAK_HD uint32_t hp_of(PH pk) const { return ld16(pk, PK_HP); }
  OAK_HD uint32_t last_selected(int p) const { return ld8(sb(p), S_LAST_SELECTED); }
  OAK_HD uint32_t active_types(int p) const { return ld8(sb(p), A_TYPES); }
  OAK_HD uint32_t active_stat(int p, int st) const { return ld16(sb(p), A_STATS + st); }
  OAK_HD void set_active_stat(int p, int st, uint32_t v) const { st16(sb(p), A_STATS + st, v); }
  OAK_HD uint32_t ident(int p) const { return ((uint32_t)p << 3) | ld8(sb(p), S_ORDER); }

  // ---- protocol log: nothing in this window type (MemL below writes); calls compile away
  static const bool LOGGING = false;
  OAK_HD void lg1(uint32_t) const {}
  OAK_HD void lg2(uint32_t, uint32_t) const {}
  OAK_HD void lg3(uint32_t, uint32_t, uint32_t) const {}
  OAK_HD void lg4(uint32_t, uint32_t, uint32_t, uint32_t) const {}
  OAK_HD void lg_hp(uint32_t, uint32_t, PH, uint32_t) const {}
  OAK_HD void lg_hp_of(uint32_t, uint32_t, PH, uint32_t, uint32_t) const {}
  OAK_HD void lg_switched(uint32_t, PH) const {}
  OAK_HD void lg_boost(uint32_t, uint32_t, int) const {}
};

// ---- SPLIT layout: record <-> column. LD::ld(const uint32_t*) reads one record word (the kernels pass a cache policy).
struct PlainLoad {
  static OAK_HD uint32_t ld(const uint32_t* p) { return *p; }
};
// the dynamic fields of the record (96 words at `rec`) -> the 48-word column
template <int STRIDE, class LD>
OAK_HD void split_fill(uint32_t* col, const uint32_t* rec) {
  for (int p = 0; p < 2; p++) {
    for (int idx = 0; idx < 6; idx++) {
      const uint32_t* r = rec + p * (SIDE_SIZE / 4) + idx * (PK_SIZE / 4);
      // record bytes 8..23: spc | id0 pp0 | id1 pp1 id2 pp2 | id3 pp3 hp | status species types level
      const uint32_t w2 = LD::ld(r + 2), w3 = LD::ld(r + 3), w4 = LD::ld(r + 4), w5 = LD::ld(r + 5);
      uint32_t* c = col + (p * 22 + idx * 2) * STRIDE;
      c[0] = (w2 >> 24) | ((w3 >> 8) & 0xFFu) << 8 | (w3 >> 24) << 16 | ((w4 >> 8) & 0xFFu) << 24;
      c[STRIDE] = (w4 >> 16) | (w5 & 0xFFu) << 16;
    }
    for (int k = 0; k < 10; k++) col[(p * 22 + 12 + k) * STRIDE] = LD::ld(rec + p * (SIDE_SIZE / 4) + S_ACTIVE / 4 + k);
  }
  for (int k = 0; k < 4; k++) col[(44 + k) * STRIDE] = LD::ld(rec + B_TURN / 4 + k);
}
// the column's fields written back into a record that already holds the constant bytes (in place: `rec` is read and written)
template <int STRIDE, class LD>
OAK_HD void split_merge(const uint32_t* col, uint32_t* rec) {
  for (int p = 0; p < 2; p++) {
    for (int idx = 0; idx < 6; idx++) {
      uint32_t* r = rec + p * (SIDE_SIZE / 4) + idx * (PK_SIZE / 4);
      const uint32_t* c = col + (p * 22 + idx * 2) * STRIDE;
      const uint32_t pp = c[0], hs = c[STRIDE];
      r[2] = (LD::ld(r + 2) & 0x00FFFFFFu) | (pp << 24);
      r[3] = (LD::ld(r + 3) & 0x00FF00FFu) | ((pp >> 8) & 0xFFu) << 8 | ((pp >> 16) & 0xFFu) << 24;
      r[4] = (LD::ld(r + 4) & 0x000000FFu) | (pp >> 24) << 8 | (hs & 0xFFFFu) << 16;
      r[5] = (LD::ld(r + 5) & 0xFFFFFF00u) | ((hs >> 16) & 0xFFu);
    }
    for (int k = 0; k < 10; k++) rec[p * (SIDE_SIZE / 4) + S_ACTIVE / 4 + k] = col[(p * 22 + 12 + k) * STRIDE];
  }
  for (int k = 0; k < 4; k++) rec[B_TURN / 4 + k] = col[(44 + k) * STRIDE];
}

// Binary protocol log of one update (reference engine/src/lib/common/protocol.zig:313-863). The sink is the caller's
// buffer (`Writer`, :831-863): a write that does not fit is refused whole and raises `overflow`, after which nothing more
// is written, so an undersized buffer ends with the same bytes as the reference's; the update then returns 0b1100
// (bindings/c.zig:6,181). The byte groups passed to lgN are the reference's writeAll / writeInt / writeByte groups.
struct LogSink {
  uint8_t* buf;
  uint32_t cap, pos, overflow;
  OAK_HD bool room(uint32_t n) {
    if (overflow || pos + n > cap) {
      overflow = 1;
      return false;
    }
    return true;
  }
};
static const uint32_t RESULT_LOG_OVERFLOW = 0x0C;

// The memory window of a LOG build: Mem plus the sink, so that the non-inlined state routines (gen1_mem_ops.cuh), which
// receive the window by value, can log. The non-LOG kernels never see this type: their code is unchanged.
template <int STRIDE>
struct MemL : Mem<STRIDE> {
  LogSink* sink;
  static const bool LOGGING = true;
  OAK_HD void lg1(uint32_t a) const {
    if (!sink->room(1)) return;
    sink->buf[sink->pos++] = (uint8_t)a;
  }
  OAK_HD void lg2(uint32_t a, uint32_t b) const {
    if (!sink->room(2)) return;
    uint8_t* o = sink->buf + sink->pos;
    o[0] = (uint8_t)a;
    o[1] = (uint8_t)b;
    sink->pos += 2;
  }
  OAK_HD void lg3(uint32_t a, uint32_t b, uint32_t c) const {
    if (!sink->room(3)) return;
    uint8_t* o = sink->buf + sink->pos;
    o[0] = (uint8_t)a;
    o[1] = (uint8_t)b;
    o[2] = (uint8_t)c;
    sink->pos += 3;
  }
  OAK_HD void lg4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) const {
    if (!sink->room(4)) return;
    uint8_t* o = sink->buf + sink->pos;
    o[0] = (uint8_t)a;
    o[1] = (uint8_t)b;
    o[2] = (uint8_t)c;
    o[3] = (uint8_t)d;
    sink->pos += 4;
  }
  OAK_HD void lg16(uint32_t v) const { lg2(v & 255u, v >> 8); }
  // damage / heal (protocol.zig:432-458): op, ident | hp | max hp | status, reason [| of]
  OAK_HD void lg_hp(uint32_t op, uint32_t id, PH pk, uint32_t reason) const {
    lg2(op, id);
    lg16(this->ld16(pk, PK_HP));
    lg16(this->ld16(pk, PK_STATS + ST_HP));
    lg2(this->ld8(pk, PK_STATUS), reason);
  }
  OAK_HD void lg_hp_of(uint32_t op, uint32_t id, PH pk, uint32_t reason, uint32_t of) const {
    lg_hp(op, id, pk, reason);
    lg1(of);
  }
  // switched (protocol.zig:355-378): op, ident | species, level | hp | max hp | status
  OAK_HD void lg_switched(uint32_t id, PH pk) const {
    lg2(LG_Switch, id);
    lg2(this->ld8(pk, PK_SPECIES), this->ld8(pk, PK_LEVEL));
    lg16(this->ld16(pk, PK_HP));
    lg16(this->ld16(pk, PK_STATS + ST_HP));
    lg1(this->ld8(pk, PK_STATUS));
  }
  OAK_HD void lg_boost(uint32_t id, uint32_t reason, int num) const { lg4(LG_Boost, id, reason, (uint32_t)(num + 6)); }
};

template <int STRIDE, bool LOG, bool SPLIT = false>
struct MemFor {
  typedef Mem<STRIDE, SPLIT> type;
};
template <int STRIDE>
struct MemFor<STRIDE, true, false> {
  typedef MemL<STRIDE> type;
};

}  // namespace oakb200
