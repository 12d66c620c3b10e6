// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement of the reference's Generation I battle engine (libpkmn, Zig) in plain C++17.
// Nothing in the product path (oak_b200/) may include, link or call this file: it exists so
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs can check
// and time the algorithm on the host.
//
// Source followed (paths relative to /root/reference/cpp/extern/engine/src/lib):
//   gen1/mechanics.zig  (update :53, start :81, selectMove :104, saveMove :205, switchIn :225,
//                        turnOrder :265, doTurn :318, executeMove :383, beforeMove :484,
//                        canMove :746, decrementPP :808, incrementPP :828, doMove :844,
//                        checkCriticalHit :1104, calcDamage :1118, adjustDamage :1197,
//                        randomizeDamage :1228, specialDamage :1236, counterDamage :1266,
//                        applyDamage :1314, mirrorMove :1351, metronome :1378, checkHit :1398,
//                        moveHit :1422, checkFaint :1507, faint :1545, handleResidual :1585,
//                        endTurn :1639, checkEBC :1662, disabled :1716, buildRage :1734,
//                        Effects.* :1821-2690, clearVolatiles :2711, Rolls.* :2774-3115,
//                        choices :3139-3251)
//   gen1/data.zig       (layouts :54-68,123-138,157-173,185-203,215-232,235-301,325-383,448-463)
//   gen1/chance.zig     (Action :160-206, Duration :279-287, recorders :367-756)
//   gen1/calc.zig       (Calc.overridden :94-108, Calc.confusion :110-116)
//   common/rng.zig      (PSRNG :34-59, Gen56 :149-164, FixedRNG :181-222)
//
// The engine is a template over the reference's build options. Pokémon Showdown compatibility
// mode (`-Dshowdown`, hence mod=true) is the only mode restated; the remaining options are
// template parameters so the same code runs (a) with libpkmn's defaults {miss, advance, ebc} and
// a scripted roll list, which is how the reference's own known-answer tests (gen1/test.zig) are
// written, and (b) with Oak's build flags {miss=false, advance=false, ebc=false, key=true}
// (reference dev/libpkmn:9), which is the parity target for the CUDA engine.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cstring>

#include "gen1_tables.h"

namespace oracle {

#pragma pack(push, 1)
struct MoveSlot {
  uint8_t id, pp;
};
struct Stats {
  uint16_t hp, atk, def, spe, spc;
};
struct Pokemon {  // data.zig:185-203
  Stats stats;
  MoveSlot moves[4];
  uint16_t hp;
  uint8_t status, species, types, level;
};
struct Active {  // data.zig:157-173
  Stats stats;
  uint8_t species, types;
  uint32_t boosts;
  uint64_t volatiles;
  MoveSlot moves[4];
};
struct Side {  // data.zig:123-138
  Pokemon pokemon[6];
  Active active;
  uint8_t order[6];
  uint8_t last_selected_move, last_used_move;
};
struct MoveDetails {  // data.zig:227-232 (showdown layout)
  uint8_t index, counterable;
};
struct Battle {  // data.zig:54-68
  Side sides[2];
  uint16_t turn, last_damage;
  MoveDetails last_moves[2];
  uint64_t rng;
};
#pragma pack(pop)
static_assert(sizeof(Pokemon) == 24 && sizeof(Active) == 32 && sizeof(Side) == 184 && sizeof(Battle) == 384, "layout");

// Caller-visible option block (our own layout; the libpkmn-compatible 128-byte opaque lives in
// the C API shim). actions/durations/overrides follow chance.zig bit-for-bit.
struct Options {
  uint64_t actions[2];    // chance.zig:160-206, written per update
  uint32_t durations[2];  // chance.zig:279-287, carried between updates
  uint64_t overrides[2];  // calc.zig:88-92, read (and `confusion` cleared) per update
};

// Scripted roll source (rng.zig:181-222) used only when FIXED.
struct Rolls32 {
  const uint32_t* rolls;
  size_t len, index;
  bool overrun;
};

enum : uint8_t { P1 = 0, P2 = 1 };
enum : uint8_t { CH_PASS = 0, CH_MOVE = 1, CH_SWITCH = 2 };
enum : uint8_t { RES_NONE = 0, RES_WIN = 1, RES_LOSE = 2, RES_TIE = 3, RES_ERROR = 4 };
static const uint8_t RESULT_DEFAULT = 0x50;  // p1 = Move, p2 = Move (common/data.zig:109)

// Status (data.zig:235-301)
enum : uint8_t { ST_SLP = 2, ST_PSN = 3, ST_BRN = 4, ST_FRZ = 5, ST_PAR = 6, ST_EXT = 7 };
static const uint8_t STATUS_TOX = 0x88;
static inline bool status_is(uint8_t s, int which) {
  if (which == ST_SLP) return (s & 7) > 0;
  return ((s >> which) & 1) != 0;
}
static inline uint8_t status_init(int which) { return (uint8_t)(1u << which); }

// Volatiles bit positions (data.zig:325-383)
enum {
  V_Bide = 0, V_Thrashing, V_MultiHit, V_Flinch, V_Charging, V_Binding, V_Invulnerable, V_Confusion,
  V_Mist, V_FocusEnergy, V_Substitute, V_Recharging, V_Rage, V_LeechSeed, V_Toxic, V_LightScreen,
  V_Reflect, V_Transform,
};
struct VField { int shift, bits; };
static const VField VF_confusion{18, 3}, VF_attacks{21, 3}, VF_state{24, 16}, VF_substitute{40, 8},
    VF_transform{48, 4}, VF_disable_duration{52, 4}, VF_disable_move{56, 3}, VF_toxic{59, 5};

static inline bool vget(const Active& a, int bit) { return (a.volatiles >> bit) & 1; }
static inline void vset(Active& a, int bit, bool v) {
  a.volatiles = (a.volatiles & ~(1ull << bit)) | ((uint64_t)v << bit);
}
static inline uint32_t vfget(const Active& a, VField f) {
  return (uint32_t)((a.volatiles >> f.shift) & ((1ull << f.bits) - 1));
}
static inline void vfset(Active& a, VField f, uint32_t v) {
  const uint64_t m = ((1ull << f.bits) - 1) << f.shift;
  a.volatiles = (a.volatiles & ~m) | (((uint64_t)v << f.shift) & m);
}

// Boosts: six i4 nibbles (data.zig:448-463)
enum { B_atk = 0, B_def, B_spe, B_spc, B_accuracy, B_evasion };
static inline int bget(const Active& a, int which) {
  int v = (a.boosts >> (4 * which)) & 15;
  return v >= 8 ? v - 16 : v;
}
static inline void bset(Active& a, int which, int v) {
  a.boosts = (a.boosts & ~(15u << (4 * which))) | (((uint32_t)v & 15u) << (4 * which));
}

// Action fields (chance.zig:160-206); Optional(T) puts None at 0 (optional.zig:8-27)
struct AField { int shift, bits; };
static const AField A_damage{0, 8}, A_hit{8, 2}, A_critical_hit{10, 2}, A_secondary_chance{12, 2},
    A_speed_tie{14, 2}, A_confused{16, 2}, A_paralyzed{18, 2}, A_duration{20, 4}, A_sleep{24, 3},
    A_confusion{27, 4}, A_disable{31, 3}, A_attacking{34, 2}, A_binding{36, 2}, A_move_slot{38, 3},
    A_multi_hit{41, 3}, A_pp{44, 4}, A_psywave{48, 8}, A_metronome{56, 8};
static inline uint32_t aget(uint64_t a, AField f) { return (uint32_t)((a >> f.shift) & ((1ull << f.bits) - 1)); }
static inline void aset(uint64_t& a, AField f, uint32_t v) {
  const uint64_t m = ((1ull << f.bits) - 1) << f.shift;
  a = (a & ~m) | (((uint64_t)v << f.shift) & m);
}
enum : uint32_t { OPT_NONE = 0, OPT_FALSE = 1, OPT_TRUE = 2 };
enum : uint32_t { OBS_NONE = 0, OBS_STARTED = 1, OBS_CONTINUING = 2, OBS_ENDED = 3, OBS_OVERWRITTEN = 4 };

// Duration fields (chance.zig:279-287)
static const AField D_confusion{18, 3}, D_disable{21, 4}, D_attacking{25, 3}, D_binding{28, 3};
static inline uint32_t dget(uint32_t d, AField f) { return (d >> f.shift) & ((1u << f.bits) - 1); }
static inline void dset(uint32_t& d, AField f, uint32_t v) {
  const uint32_t m = ((1u << f.bits) - 1) << f.shift;
  d = (d & ~m) | ((v << f.shift) & m);
}
static inline uint32_t sleeps_get(uint32_t d, int i) { return (d >> (3 * i)) & 7; }
static inline void sleeps_set(uint32_t& d, int i, uint32_t v) {
  d = (d & ~(7u << (3 * i))) | ((v & 7u) << (3 * i));
}

// Move data accessors over the generated table (moves.zig:184-194)
struct MoveData {
  uint8_t effect, bp, accuracy, type;
  bool target_self;
};
static inline MoveData move_get(uint8_t id) {
  const uint32_t w = GEN1_MOVE_DATA[id];
  return MoveData{(uint8_t)(w & 255), (uint8_t)((w >> 8) & 255), (uint8_t)((w >> 16) & 255),
                  (uint8_t)((w >> 24) & 15), ((w >> 28) & 1) != 0};
}
// Effect ordinal ranges (moves.zig:1600-1640)
static inline bool eff_onBegin(uint8_t e) { return e > 0 && e <= 16; }
static inline bool eff_isStatDown(uint8_t e) { return e > 16 && e <= 21; }
static inline bool eff_onEnd(uint8_t e) { return e > 16 && e <= 31; }
static inline bool eff_alwaysHappens(uint8_t e) { return e > 31 && e <= 38; }
static inline bool eff_isSpecial(uint8_t e) { return e > 31 && e <= 46; }
static inline bool eff_isMulti(uint8_t e) { return e > 44 && e <= 47; }
static inline bool eff_isStatDownChance(uint8_t e) { return e > 47 && e <= 51; }

static inline uint8_t type_eff(uint8_t atk_type, uint8_t def_type) { return GEN1_TYPE_CHART[atk_type * 15 + def_type]; }
static inline bool types_includes(uint8_t types, uint8_t t) { return (types & 15) == t || (types >> 4) == t; }
static inline bool type_special(uint8_t t) { return t >= TYPE_Fire; }

static const uint8_t BOOSTS[13][2] = {  // mechanics.zig:34-48
    {25, 100}, {28, 100}, {33, 100}, {40, 100}, {50, 100}, {66, 100}, {1, 1},
    {15, 10},  {2, 1},    {25, 10},  {3, 1},    {35, 10},  {4, 1}};
static const uint8_t DISTRIBUTION[8] = {2, 2, 2, 3, 3, 3, 4, 5};  // mechanics.zig:3038
static const int NULLR = -1;  // "no result yet" (Zig `null` of ?Result)

#ifndef ORACLE_ASSERT
#define ORACLE_ASSERT(x) ((void)0)
#endif

// Protocol log sink (common/protocol.zig:831-863 `Writer`): fixed caller buffer; a write that does not fit
// aborts the update (Zig `try` -> error.NoSpaceLeft -> result byte 0b1100, bindings/c.zig:6,181).
struct LogBuf {
  uint8_t* buf;
  size_t cap, pos;
};
struct LogOverflow {};
static const uint8_t RESULT_LOG_OVERFLOW = 0x0C;

// ArgType / reason ordinals (common/protocol.zig:16-311), Generation I subset
enum : uint8_t {
  LG_None = 0, LG_LastStill, LG_LastMiss, LG_Move, LG_Switch, LG_Cant, LG_Faint, LG_Turn, LG_Win, LG_Tie, LG_Damage,
  LG_Heal, LG_Status, LG_CureStatus, LG_Boost, LG_ClearAllBoost, LG_Fail, LG_Miss, LG_HitCount, LG_Prepare,
  LG_MustRecharge, LG_Activate, LG_FieldActivate, LG_Start, LG_End, LG_OHKO, LG_Crit, LG_SuperEffective, LG_Resisted,
  LG_Immune, LG_Transform,
};
enum : uint8_t { CANT_Sleep = 0, CANT_Freeze, CANT_Paralysis, CANT_Bound, CANT_Flinch, CANT_Disable, CANT_Recharge };
enum : uint8_t { HEAL_None = 0, HEAL_Silent, HEAL_Drain };
enum : uint8_t { DMG_None = 0, DMG_Poison, DMG_Burn, DMG_Confusion, DMG_LeechSeed, DMG_RecoilOf };
enum : uint8_t { STS_None = 0, STS_Silent, STS_From };
enum : uint8_t { CURE_Message = 0, CURE_Silent };
enum : uint8_t { BST_Rage = 0, BST_Attack, BST_Defense, BST_Speed, BST_SpecialAttack, BST_SpecialDefense, BST_Accuracy, BST_Evasion };
enum : uint8_t { FAIL_None = 0, FAIL_Sleep, FAIL_Poison, FAIL_Burn, FAIL_Freeze, FAIL_Paralysis, FAIL_Toxic, FAIL_Substitute, FAIL_Weak };
enum : uint8_t { ACT_Bide = 0, ACT_Confusion, ACT_Haze, ACT_Mist, ACT_Struggle, ACT_Substitute, ACT_Splash };
enum : uint8_t {
  START_Bide = 0, START_Confusion, START_ConfusionSilent, START_FocusEnergy, START_LeechSeed, START_LightScreen, START_Mist,
  START_Reflect, START_Substitute, START_TypeChange, START_Disable, START_Mimic,
};
enum : uint8_t {
  END_Disable = 0, END_Confusion, END_Bide, END_Substitute, END_DisableSilent = 15, END_ConfusionSilent, END_MistSilent,
  END_FocusEnergySilent, END_LeechSeedSilent, END_ToxicSilent, END_LightScreenSilent, END_ReflectSilent,
};
enum : uint8_t { IMM_None = 0, IMM_OHKO };

template <bool MISS, bool ADVANCE, bool EBC, bool KEY, bool FIXED, bool LOG = false>
struct Engine {
  Battle& b;
  Options& o;
  Rolls32* fx;
  LogBuf* lg;
  // flag cross-check instrumentation (tests/test_oracle_flags.py): how often this update reached a site where the build
  // flags change behaviour — an `advance` call (ADVANCE), a hit roll that landed exactly on the accuracy (MISS: the 1/256 miss)
  uint32_t n_advance = 0, n_hit_edge = 0;

  Engine(Battle& battle, Options& options, Rolls32* fixed = nullptr, LogBuf* log = nullptr)
      : b(battle), o(options), fx(fixed), lg(log) {}

  // ------------------------------------------------------------------ protocol log (common/protocol.zig:313-824)
  // Every writer below is one `Log.*` method; the grouping of bytes into `lw` calls is the reference's grouping into
  // writeAll / writeInt / writeByte calls, so that an overflowing buffer is left with exactly the same prefix.
  void lw(const uint8_t* bytes, size_t n) {
    if (lg->pos + n > lg->cap) throw LogOverflow{};
    memcpy(lg->buf + lg->pos, bytes, n);
    lg->pos += n;
  }
  template <class... A>
  void lwb(A... a) {
    const uint8_t t[] = {(uint8_t)a...};
    lw(t, sizeof...(A));
  }
  void lw16(uint16_t v) { lwb(v & 255, v >> 8); }
  void log_move(uint8_t src, uint8_t move, uint8_t target, uint8_t from = 0) {  // :337-352
    if (!LOG) return;
    lwb(LG_Move, src, move, target);
    if (from == 0) lwb(0); else lwb(1, from);
  }
  void log_switched(uint8_t id, const Pokemon& pk) {  // :355-378
    if (!LOG) return;
    lwb(LG_Switch, id);
    lwb(pk.species, pk.level);
    lw16(pk.hp);
    lw16(pk.stats.hp);
    lwb(pk.status);
  }
  void log_cant(uint8_t id, uint8_t reason) {  // :382-393
    if (!LOG) return;
    lwb(LG_Cant, id, reason);
  }
  void log_cant_disable(uint8_t id, uint8_t move) {
    if (!LOG) return;
    lwb(LG_Cant, id, CANT_Disable);
    lwb(move);
  }
  void log_faint(uint8_t id, bool done) {  // :396-401
    if (!LOG) return;
    lwb(LG_Faint, id);
    if (done) lwb(LG_None);
  }
  void log_turn(uint16_t n) {  // :404-410
    if (!LOG) return;
    lwb(LG_Turn);
    lw16(n);
    lwb(LG_None);
  }
  void log_win(int player) {  // :413-421
    if (!LOG) return;
    lwb(LG_Win, player, LG_None);
  }
  void log_tie() {  // :424-428
    if (!LOG) return;
    lwb(LG_Tie, LG_None);
  }
  void log_hp(uint8_t op, uint8_t id, const Pokemon& pk, uint8_t reason, int of) {  // damage :432-444, heal :448-458
    if (!LOG) return;
    lwb(op, id);
    lw16(pk.hp);
    lw16(pk.stats.hp);
    lwb(pk.status, reason);
    if (of >= 0) lwb(of);
  }
  void log_damage(uint8_t id, const Pokemon& pk, uint8_t reason, int of = -1) { log_hp(LG_Damage, id, pk, reason, of); }
  void log_heal(uint8_t id, const Pokemon& pk, uint8_t reason, int of = -1) { log_hp(LG_Heal, id, pk, reason, of); }
  void log_status(uint8_t id, uint8_t value, uint8_t reason, int from = -1) {  // :462-473
    if (!LOG) return;
    lwb(LG_Status, id, value, reason);
    if (from >= 0) lwb(from);
  }
  void log_curestatus(uint8_t id, uint8_t value, uint8_t reason) {  // :476-485
    if (!LOG) return;
    lwb(LG_CureStatus, id, value, reason);
  }
  void log_boost(uint8_t id, uint8_t reason, int num) {  // :488-498
    if (!LOG) return;
    lwb(LG_Boost, id, reason, num + 6);
  }
  void log_op(uint8_t op) {  // clearallboost :501, fieldactivate :599, ohko :644, laststill :810, lastmiss :817
    if (!LOG) return;
    lwb(op);
  }
  void log_id(uint8_t op, uint8_t id) {  // miss :519, mustrecharge :549, crit :651, supereffective :658, resisted :668
    if (!LOG) return;
    lwb(op, id);
  }
  void log_id_arg(uint8_t op, uint8_t id, uint8_t arg) {  // fail :508, hitcount :526, prepare :537, activate :563, start :606, end :633, immune :678, transform :689
    if (!LOG) return;
    lwb(op, id, arg);
  }
  void log_fail(uint8_t id, uint8_t reason) { log_id_arg(LG_Fail, id, reason); }
  void log_lastmiss_miss(int p) {
    log_op(LG_LastMiss);
    log_id(LG_Miss, ident(p));
  }
  void log_activate(uint8_t id, uint8_t reason) { log_id_arg(LG_Activate, id, reason); }
  void log_start(uint8_t id, uint8_t reason) { log_id_arg(LG_Start, id, reason); }
  void log_start_move(uint8_t id, uint8_t reason, uint8_t move) {  // .Disable / .Mimic
    if (!LOG) return;
    lwb(LG_Start, id, reason);
    lwb(move);
  }
  void log_start_typechange(uint8_t id, uint8_t types, uint8_t target) {
    if (!LOG) return;
    lwb(LG_Start, id, START_TypeChange);
    lwb(types, target);
  }
  void log_end(uint8_t id, uint8_t reason) { log_id_arg(LG_End, id, reason); }

  // ------------------------------------------------------------------ accessors
  Side& side(int p) { return b.sides[p]; }
  Side& foe(int p) { return b.sides[p ^ 1]; }
  static Pokemon& get(Side& s, int slot) { return s.pokemon[s.order[slot - 1] - 1]; }
  static Pokemon& stored(Side& s) { return get(s, 1); }
  uint8_t ident(int p) { return (uint8_t)((p << 3) | b.sides[p].order[0]); }  // common/data.zig ID

  // ------------------------------------------------------------------ PRNG (rng.zig:34-59,149-164)
  uint32_t next() {
    if (FIXED) {
      if (fx->index >= fx->len) {
        fx->overrun = true;
        return 0;
      }
      return fx->rolls[fx->index++];
    }
    b.rng = 0x5D588B656C078965ull * b.rng + 0x269EC3ull;
    return (uint32_t)(b.rng >> 32);
  }
  void advance(int n) {
    n_advance += (uint32_t)n;
    if (!ADVANCE) return;
    for (int i = 0; i < n; i++) (void)next();
  }
  uint32_t range(uint32_t from, uint32_t to) { return (uint32_t)(((uint64_t)next() * (to - from)) >> 32) + from; }
  bool chance(uint32_t num, uint32_t den) { return range(0, den) < num; }

  // ------------------------------------------------------------------ calc overrides (calc.zig:94-116)
  uint32_t overridden(int p, AField f) { return aget(o.overrides[p], f); }
  void calc_confusion(int p) {
    if (aget(o.overrides[p], A_confusion) == OBS_STARTED) aset(o.overrides[p], A_confusion, OBS_NONE);
  }

  // ------------------------------------------------------------------ chance recorders (chance.zig:435-756)
  void ch_switched(int p, int slot) {
    uint32_t& d = o.durations[p];
    const uint32_t out = sleeps_get(d, 0);
    sleeps_set(d, 0, sleeps_get(d, slot - 1));
    sleeps_set(d, slot - 1, out);
    dset(d, D_confusion, 0);
    dset(d, D_disable, 0);
    dset(d, D_attacking, 0);
    dset(d, D_binding, 0);
    dset(o.durations[p ^ 1], D_binding, 0);
  }
  void ch_bool(int p, AField f, bool v) { aset(o.actions[p], f, v ? OPT_TRUE : OPT_FALSE); }
  void ch_speedTie(bool p1) {
    aset(o.actions[0], A_speed_tie, p1 ? 1 : 2);
    aset(o.actions[1], A_speed_tie, p1 ? 1 : 2);
  }
  void ch_duration(int p, uint32_t turns) { aset(o.actions[p], A_duration, KEY ? 2 : turns); }
  enum Obs { F_sleep, F_confusion, F_disable, F_attacking, F_binding };
  static AField obs_afield(Obs f) {
    switch (f) {
      case F_sleep: return A_sleep;
      case F_confusion: return A_confusion;
      case F_disable: return A_disable;
      case F_attacking: return A_attacking;
      default: return A_binding;
    }
  }
  void dur_set(int p, Obs f, uint32_t v) {
    uint32_t& d = o.durations[p];
    switch (f) {
      case F_sleep: sleeps_set(d, 0, v); break;
      case F_confusion: dset(d, D_confusion, v); break;
      case F_disable: dset(d, D_disable, v); break;
      case F_attacking: dset(d, D_attacking, v); break;
      case F_binding: dset(d, D_binding, v); break;
    }
  }
  void ch_observe(Obs f, int p, uint32_t obs) {  // chance.zig:556-600
    uint64_t& a = o.actions[p];
    switch (obs) {
      case OBS_OVERWRITTEN:
        ORACLE_ASSERT(f == F_confusion);
        aset(a, A_confusion, obs);
        dset(o.durations[p], D_confusion, 1);
        break;
      case OBS_NONE:
        aset(a, obs_afield(f), OBS_NONE);
        dur_set(p, f, 0);
        break;
      case OBS_STARTED:
        ORACLE_ASSERT(aget(a, obs_afield(f)) == OBS_NONE || aget(a, obs_afield(f)) == OBS_ENDED);
        aset(a, obs_afield(f), OBS_STARTED);
        dur_set(p, f, 1);
        break;
      default:
        // chance.zig:598 `unreachable`, but mechanics.zig:714-717 passes .continuing when
        // overwrite=false; overwrite is always true here.
        ORACLE_ASSERT(false);
    }
  }
  void ch_sleep(int p, uint32_t obs) {  // chance.zig:602-624
    aset(o.actions[p], A_sleep, obs);
    uint32_t& d = o.durations[p];
    const uint32_t n = sleeps_get(d, 0);
    if (obs == OBS_NONE || obs == OBS_ENDED) {
      sleeps_set(d, 0, 0);
    } else {
      ORACLE_ASSERT(n >= 1 && n < 7);
      sleeps_set(d, 0, n + 1);
    }
  }
  void ch_tick(int p, Obs f, AField df, uint32_t obs) {  // chance.zig:626-733 (probability dropped)
    aset(o.actions[p], obs_afield(f), obs);
    uint32_t& d = o.durations[p];
    if (obs == OBS_ENDED) {
      dset(d, df, 0);
    } else {
      dset(d, df, dget(d, df) + 1);
    }
  }
  void ch_confusion(int p, uint32_t obs) { ch_tick(p, F_confusion, D_confusion, obs); }
  void ch_disable(int p, uint32_t obs) { ch_tick(p, F_disable, D_disable, obs); }
  void ch_attacking(int p, uint32_t obs) { ch_tick(p, F_attacking, D_attacking, obs); }
  void ch_binding(int p, uint32_t obs) { ch_tick(p, F_binding, D_binding, obs); }

  // ------------------------------------------------------------------ Rolls (mechanics.zig:2774-3115)
  bool roll_speedTie() {
    const uint32_t ov = overridden(P1, A_speed_tie);
    const bool p1 = ov ? (ov == 1) : (range(0, 2) == 0);
    ch_speedTie(p1);
    return p1;
  }
  bool roll_criticalHit(int p, uint8_t rate) {
    const uint32_t ov = overridden(p, A_critical_hit);
    const bool crit = ov ? (ov == OPT_TRUE) : chance(rate, 256);
    ch_bool(p, A_critical_hit, crit);
    return crit;
  }
  uint8_t roll_damage(int p) {
    const uint32_t ov = overridden(p, A_damage);
    const uint8_t roll = ov ? (uint8_t)ov : (uint8_t)range(217, 256);
    ORACLE_ASSERT(roll >= 217);
    aset(o.actions[p], A_damage, roll);
    return roll;
  }
  bool roll_hit(int p, uint8_t accuracy) {
    const uint32_t acc = MISS ? accuracy : (uint32_t)accuracy + 1;
    const uint32_t ov = overridden(p, A_hit);
    bool ok;
    if (ov) {
      ok = ov == OPT_TRUE;
    } else {
      const uint32_t r = range(0, 256);  // chance(acc, 256) = range(0, 256) < acc
      ok = r < acc;
      if (r == accuracy) n_hit_edge++;
    }
    ch_bool(p, A_hit, ok);
    return ok;
  }
  bool roll_confused(int p) {
    const uint32_t ov = overridden(p, A_confused);
    const bool cfz = ov ? (ov == OPT_TRUE) : !chance(128, 256);
    ch_bool(p, A_confused, cfz);
    return cfz;
  }
  bool roll_paralyzed(int p) {
    const uint32_t ov = overridden(p, A_paralyzed);
    const bool par = ov ? (ov == OPT_TRUE) : chance(63, 256);
    ch_bool(p, A_paralyzed, par);
    return par;
  }
  bool roll_secondary(int p, uint8_t rate) {  // confusionChance/secondaryChance/poisonChance/unboost
    const uint32_t ov = overridden(p, A_secondary_chance);
    const bool proc = ov ? (ov == OPT_TRUE) : chance(rate, 256);
    ch_bool(p, A_secondary_chance, proc);
    return proc;
  }
  uint8_t roll_metronome(int p) {
    const uint32_t ov = overridden(p, A_metronome);
    uint8_t m;
    if (ov) {
      m = (uint8_t)ov;
    } else {
      const uint32_t r = range(0, 163);  // Move.METRONOME (moves.zig:1840-1851): 1..164 minus Metronome
      m = (uint8_t)(r + 1 < MOVE_Metronome ? r + 1 : r + 2);
    }
    aset(o.actions[p], A_metronome, m);
    return m;
  }
  uint8_t roll_psywave(int p, uint8_t max) {
    const uint32_t ov = overridden(p, A_psywave);
    const uint8_t power = ov ? (uint8_t)(ov - 1) : (uint8_t)range(0, max);
    aset(o.actions[p], A_psywave, (uint32_t)power + 1);
    return power;
  }
  uint32_t roll_duration(int p, uint32_t lo, uint32_t hi_excl) {  // sleep/disable/confusion/attacking
    const uint32_t ov = overridden(p, A_duration);
    const uint32_t d = (ov >= lo && ov < hi_excl) ? ov : range(lo, hi_excl);
    ch_duration(p, d);
    return d;
  }
  uint32_t roll_distribution(int p, bool binding) {
    const uint32_t ov = overridden(p, binding ? A_duration : A_multi_hit);
    const uint32_t n = (ov >= 2 && ov <= 5) ? ov : DISTRIBUTION[range(0, 8)];
    if (binding) {
      ch_duration(p, n);
    } else {
      aset(o.actions[p], A_multi_hit, n);
    }
    return n;
  }
  uint32_t roll_moveSlot(int p, MoveSlot* moves, uint32_t check_pp) {
    const uint32_t ov = overridden(p, A_move_slot);
    uint32_t slot = 0;
    if (ov && moves[ov - 1].id != MOVE_None && (check_pp == 0 || moves[ov - 1].pp > 0)) {
      slot = ov;
    } else if (check_pp == 0) {
      for (int i = 4; i > 0;) {
        i--;
        if (moves[i].id != MOVE_None) {
          slot = range(0, (uint32_t)i + 1) + 1;
          break;
        }
      }
    } else {
      uint32_t r = range(0, check_pp) + 1;
      for (int i = 0; i < 4 && r > 0; i++) {
        if (moves[i].pp > 0) {
          r--;
          if (r == 0) slot = (uint32_t)i + 1;
        }
      }
    }
    aset(o.actions[p], A_move_slot, slot);
    return slot;
  }
  template <class T>
  T decrement(AField f, int p, T n) {  // mechanics.zig:2761-2772
    const uint32_t obs = overridden(p, f);
    if (obs == OBS_NONE) return (T)(n - 1);
    if (obs == OBS_STARTED || obs == OBS_ENDED) return 0;
    return n > 1 ? (T)(n - 1) : n;
  }

  // ------------------------------------------------------------------ update (mechanics.zig:53-79)
  uint8_t update(uint8_t c1, uint8_t c2) {
    const int t1 = c1 & 3, t2 = c2 & 3;
    if (b.turn == 0) return start();

    bool s1 = false, s2 = false;
    selectMove(P1, c1, s1);
    selectMove(P2, c2, s2);

    Side& p1 = side(P1);
    Side& p2 = side(P2);
    const bool r1 = vget(p1.active, V_Binding) && t2 == CH_SWITCH;
    const bool r2 = vget(p2.active, V_Binding) && t1 == CH_SWITCH;

    int r;
    if (turnOrder(c1, c2) == P1) {
      r = doTurn(P1, c1, r1, s1, P2, c2, r2, s2);
    } else {
      r = doTurn(P2, c2, r2, s2, P1, c1, r1, s1);
    }
    if (r != NULLR) return (uint8_t)r;

    if (vfget(p1.active, VF_attacks) == 0) vset(p1.active, V_Binding, false);
    if (vfget(p2.active, VF_attacks) == 0) vset(p2.active, V_Binding, false);
    return endTurn();
  }

  static int findFirstAlive(const Side& s) {  // :99-102
    for (int i = 0; i < 6; i++)
      if (s.pokemon[i].hp > 0) return s.order[i];
    return 0;
  }

  uint8_t start() {  // :81-97
    const int s1 = findFirstAlive(side(P1));
    if (s1 == 0) return findFirstAlive(side(P2)) == 0 ? RES_TIE : RES_LOSE;
    const int s2 = findFirstAlive(side(P2));
    if (s2 == 0) return RES_WIN;
    switchIn(P1, s1, true);
    switchIn(P2, s2, true);
    return endTurn();
  }

  void selectMove(int p, uint8_t choice, bool& skip_turn) {  // :104-203 (showdown branches)
    const int type = choice & 3;
    if (type == CH_PASS) return;
    Side& s = side(p);
    Active& a = s.active;
    const Pokemon& st = stored(s);
    const bool foe_binding = vget(foe(p).active, V_Binding);

    if (vget(a, V_Recharging)) {
      if (foe_binding) skip_turn = true;
      return;
    }
    if (vget(a, V_Rage)) {
      if (foe_binding) skip_turn = true;
      return;  // saveMove(null) is a no-op
    }
    if (vget(a, V_Thrashing) || vget(a, V_Charging)) {
      if (foe_binding) skip_turn = true;
      return;
    }
    if (type == CH_SWITCH) return;
    if (status_is(st.status, ST_FRZ) || status_is(st.status, ST_SLP) || vget(a, V_Bide)) {
      if (vget(a, V_Bide) && foe_binding) skip_turn = true;
      saveMove(p, choice);
      return;
    }
    if (vget(a, V_Binding)) {
      saveMove(p, choice);
      return;
    }
    if (foe_binding) {
      skip_turn = true;
      saveMove(p, choice);
      return;
    }
    vfset(a, VF_state, 0);
    saveMove(p, choice);
  }

  void saveMove(int p, uint8_t choice) {  // :205-223
    Side& s = side(p);
    const int data = choice >> 2;
    if (data == 0) {
      s.last_selected_move = MOVE_Struggle;
    } else {
      s.last_selected_move = s.active.moves[data - 1].id;
      b.last_moves[p].index = (uint8_t)data;
    }
  }

  static void statusModify(uint8_t status, Stats& stats) {  // :2698-2704
    if (status_is(status, ST_PAR)) {
      stats.spe = (uint16_t)(stats.spe / 4 > 1 ? stats.spe / 4 : 1);
    } else if (status_is(status, ST_BRN)) {
      stats.atk = (uint16_t)(stats.atk / 2 > 1 ? stats.atk / 2 : 1);
    }
  }

  void switchIn(int p, int slot, bool /*initial*/) {  // :225-263
    Side& s = side(p);
    Side& f = foe(p);
    Active& a = s.active;
    Pokemon& incoming = get(s, slot);

    const uint8_t out = s.order[0];
    s.order[0] = s.order[slot - 1];
    s.order[slot - 1] = out;

    b.last_moves[p].index = 1;
    s.last_used_move = MOVE_None;
    f.last_used_move = MOVE_None;

    a.stats = incoming.stats;
    a.species = incoming.species;
    a.types = incoming.types;
    a.boosts = 0;
    a.volatiles = 0;
    memcpy(a.moves, incoming.moves, sizeof(a.moves));

    statusModify(incoming.status, a.stats);
    ch_switched(p, slot);
    vset(f.active, V_Binding, false);
    log_switched(ident(p), incoming);
    if (incoming.status == STATUS_TOX) {
      incoming.status = status_init(ST_PSN);
      log_status(ident(p), incoming.status, STS_Silent);
    }
  }

  int turnOrder(uint8_t c1, uint8_t c2) {  // :265-316
    const int t1 = c1 & 3, t2 = c2 & 3;
    if (t1 == CH_PASS) return P2;
    if (t2 == CH_PASS) return P1;
    if ((t1 == CH_SWITCH) != (t2 == CH_SWITCH)) return t1 == CH_SWITCH ? P1 : P2;
    const bool double_switch = t1 == CH_SWITCH && t2 == CH_SWITCH;
    const uint8_t m1 = side(P1).last_selected_move, m2 = side(P2).last_selected_move;
    if (!double_switch) {
      if ((m1 == MOVE_QuickAttack) != (m2 == MOVE_QuickAttack)) {
        return m1 == MOVE_QuickAttack ? P1 : P2;
      } else if ((m1 == MOVE_Counter) != (m2 == MOVE_Counter)) {
        return m1 == MOVE_Counter ? P2 : P1;
      }
    }
    const uint16_t spe1 = side(P1).active.stats.spe, spe2 = side(P2).active.stats.spe;
    if (spe1 == spe2) {
      if (!double_switch && m1 == MOVE_Counter && m2 == MOVE_Counter) advance(1);
      const bool p1 = roll_speedTie();
      const bool th1 = vget(side(P1).active, V_Thrashing), th2 = vget(side(P2).active, V_Thrashing);
      if (th1 && th2) return P1;
      return p1 ? ((th1 && !th2) ? P2 : P1) : ((th2 && !th1) ? P1 : P2);
    }
    return spe1 > spe2 ? P1 : P2;
  }

  int doTurn(int p, uint8_t pc, bool p_rewrap, bool p_skip, int fp, uint8_t fc, bool f_rewrap, bool f_skip) {  // :318-381
    bool residual = true;
    bool replace = stored(side(p)).hp == 0;
    int r = executeMove(p, pc, p_rewrap, p_skip, residual);
    if (r != NULLR) return r;
    if (!replace) {
      if ((pc & 3) != CH_SWITCH) {
        r = checkFaint(fp);
        if (r != NULLR) return r;
      }
      if (residual) handleResidual(p);
      r = checkFaint(p);
      if (r != NULLR) return r;
    } else if ((fc & 3) == CH_PASS) {
      return NULLR;
    }

    residual = true;
    replace = stored(side(fp)).hp == 0;
    const bool calc = (fc & 3) == CH_PASS;  // pkmn.options.calc is on in Oak's build (dev/libpkmn:9)
    if (!calc) {
      r = executeMove(fp, fc, f_rewrap, f_skip, residual);
      if (r != NULLR) return r;
    }
    if (!replace) {
      if (!calc && (fc & 3) != CH_SWITCH) {
        r = checkFaint(p);
        if (r != NULLR) return r;
      }
      if (residual) handleResidual(fp);
      r = checkFaint(fp);
      if (r != NULLR) return r;
    }
    vset(side(P1).active, V_Flinch, false);
    vset(side(P2).active, V_Flinch, false);
    return NULLR;
  }

  int executeMove(int p, uint8_t choice, bool rewrap, bool skip, bool& residual) {  // :383-480
    Side& s = side(p);
    if ((choice & 3) == CH_SWITCH) {
      switchIn(p, choice >> 2, false);
      return NULLR;
    }
    ORACLE_ASSERT((choice & 3) == CH_MOVE);
    int mslot = choice >> 2;
    const bool autosel = s.last_selected_move != MOVE_None;
    if (vget(s.active, V_Charging)) mslot = b.last_moves[p].index;  // :440-444

    bool skip_can = false, skip_pp = false;
    switch (beforeMove(p, skip, residual)) {
      case BM_done: return NULLR;
      case BM_skip_can: skip_can = true; break;
      case BM_skip_pp: skip_pp = true; break;
      case BM_ok: break;
      case BM_err: return RES_ERROR;
    }

    // :459-473 Hyper Beam automatic-selection PP quirk
    if (!skip_pp && s.last_selected_move == MOVE_HyperBeam && s.last_selected_move != s.active.moves[mslot - 1].id) {
      bool has_beam = false;
      for (int i = 0; i < 4; i++) {
        if (s.active.moves[i].id == MOVE_HyperBeam) {
          mslot = i + 1;
          has_beam = true;
          break;
        }
      }
      if (!has_beam) skip_pp = true;
    }

    const bool can = skip_can || canMove(p, mslot, autosel, skip_pp, residual);
    if (!can) return NULLR;
    return doMove(p, mslot, rewrap, autosel, residual);
  }

  enum BeforeMove { BM_done, BM_skip_can, BM_skip_pp, BM_ok, BM_err };

  BeforeMove beforeMove(int p, bool skip, bool& residual) {  // :484-744
    Side& s = side(p);
    Side& f = foe(p);
    Active& a = s.active;
    Pokemon& st = stored(s);

    const uint8_t id = ident(p);

    if (status_is(st.status, ST_SLP)) {
      const uint8_t before = st.status;
      const bool slf = status_is(st.status, ST_EXT);
      const uint32_t ov = overridden(p, A_sleep);
      if (ov) {
        if (ov == OBS_STARTED || ov == OBS_ENDED) {
          st.status = 0;
        } else if ((st.status & 7) > 1) {
          st.status -= 1;
        }
      } else {
        st.status -= 1;
      }
      const uint32_t duration = st.status & 7;
      ch_sleep(p, slf ? OBS_NONE : (duration == 0 ? OBS_ENDED : OBS_CONTINUING));
      if (duration == 0) {
        log_curestatus(id, before, CURE_Message);
        st.status = 0;
      } else {
        log_cant(id, CANT_Sleep);
      }
      s.last_used_move = MOVE_None;
      return BM_done;
    }
    if (status_is(st.status, ST_FRZ)) {
      log_cant(id, CANT_Freeze);
      s.last_used_move = MOVE_None;
      return BM_done;
    }
    if (skip || vget(f.active, V_Binding)) {
      log_cant(id, CANT_Bound);
      return BM_done;
    }
    if (vget(a, V_Flinch)) {
      log_cant(id, CANT_Flinch);
      return BM_done;
    }
    if (vget(a, V_Recharging)) {
      vset(a, V_Recharging, false);
      log_cant(id, CANT_Recharge);
      return BM_done;
    }
    if (vfget(a, VF_disable_duration) > 0) {
      vfset(a, VF_disable_duration, decrement<uint32_t>(A_disable, p, vfget(a, VF_disable_duration)));
      ch_disable(p, vfget(a, VF_disable_duration) == 0 ? OBS_ENDED : OBS_CONTINUING);
      if (vfget(a, VF_disable_duration) == 0) {
        vfset(a, VF_disable_move, 0);
        log_end(id, END_Disable);
      }
    }
    if (disabled(s, id)) return BM_done;
    if (s.last_selected_move == MOVE_None) return BM_err;  // unreachable with mod=true (assert :572)

    if (vget(a, V_Confusion)) {
      ORACLE_ASSERT(vfget(a, VF_confusion) > 0);
      const uint32_t ov = overridden(p, A_confusion);
      if (ov) {
        if (ov == OBS_STARTED || ov == OBS_ENDED) {
          vfset(a, VF_confusion, 0);
        } else if (ov == OBS_CONTINUING) {
          if (vfget(a, VF_confusion) > 1) vfset(a, VF_confusion, vfget(a, VF_confusion) - 1);
        }
      } else {
        vfset(a, VF_confusion, vfget(a, VF_confusion) - 1);
      }
      ch_confusion(p, vfget(a, VF_confusion) == 0 ? OBS_ENDED : OBS_CONTINUING);
      if (vfget(a, VF_confusion) == 0) {
        vset(a, V_Confusion, false);
        log_end(id, END_Confusion);
      } else if (log_activate(id, ACT_Confusion), roll_confused(p)) {
        if (!vget(a, V_Rage)) vfset(a, VF_state, 0);
        vset(a, V_Bide, false);
        vset(a, V_Thrashing, false);
        vset(a, V_MultiHit, false);
        vset(a, V_Flinch, false);
        vset(a, V_Charging, false);
        vset(a, V_Binding, false);
        vset(a, V_Invulnerable, false);
        ch_observe(F_attacking, p, OBS_NONE);
        ch_observe(F_binding, p, OBS_NONE);
        {
          const uint16_t def = f.active.stats.def;
          f.active.stats.def = a.stats.def;
          const bool ok = calcDamage(p, p ^ 1, nullptr, false);
          f.active.stats.def = def;
          if (!ok) return BM_err;
        }
        const uint16_t uncapped = b.last_damage;
        applyDamage(p, p ^ 1, DMG_Confusion);
        b.last_damage = uncapped;
        return BM_done;
      }
    }

    if (status_is(st.status, ST_PAR) && roll_paralyzed(p)) {
      if (!vget(a, V_Rage)) vfset(a, VF_state, 0);
      vset(a, V_Bide, false);
      vset(a, V_Thrashing, false);
      vset(a, V_Charging, false);
      vset(a, V_Binding, false);
      ch_observe(F_attacking, p, OBS_NONE);
      ch_observe(F_binding, p, OBS_NONE);
      log_cant(id, CANT_Paralysis);
      return BM_done;
    }

    if (vget(a, V_Bide)) {
      uint32_t state = vfget(a, VF_state) + b.last_damage;  // saturating add (showdown)
      if (state > 65535) state = 65535;
      vfset(a, VF_state, state);
      vfset(a, VF_attacks, decrement<uint32_t>(A_attacking, p, vfget(a, VF_attacks)));
      ch_attacking(p, vfget(a, VF_attacks) == 0 ? OBS_ENDED : OBS_CONTINUING);
      if (vfget(a, VF_attacks) != 0) {
        log_activate(id, ACT_Bide);
        return BM_done;
      }

      vset(a, V_Bide, false);
      log_end(id, END_Bide);
      b.last_damage = (uint16_t)(vfget(a, VF_state) * 2);  // *% 2 wraps to u16
      vfset(a, VF_state, 0);
      if (b.last_damage == 0) {
        log_fail(id, FAIL_None);
        return BM_done;
      }

      const bool sub = vget(f.active, V_Substitute);
      applyDamage(p ^ 1, p ^ 1);
      if (stored(f).hp > 0 && !sub) buildRage(p ^ 1);
      residual = false;
      handleResidual(p);
      return BM_done;
    }

    if (vget(a, V_Thrashing)) {
      log_move(id, s.last_selected_move, ident(p ^ 1));
      vfset(a, VF_attacks, decrement<uint32_t>(A_attacking, p, vfget(a, VF_attacks)));
      ch_attacking(p, vfget(a, VF_attacks) == 0 ? OBS_ENDED : OBS_CONTINUING);
      if (vfget(a, VF_attacks) == 0) {
        const bool overwritten = vget(a, V_Confusion);
        vset(a, V_Thrashing, false);
        vset(a, V_Confusion, true);
        vfset(a, VF_confusion, roll_duration(p, 2, 6));
        log_start(id, START_ConfusionSilent);
        ch_observe(F_confusion, p, overwritten ? OBS_OVERWRITTEN : OBS_STARTED);
        if (!overwritten) calc_confusion(p);
      }
      s.last_used_move = s.last_selected_move;
      return BM_skip_can;
    }

    if (vget(a, V_Binding)) {
      vfset(a, VF_attacks, decrement<uint32_t>(A_binding, p, vfget(a, VF_attacks)));
      ch_binding(p, vfget(a, VF_attacks) == 0 ? OBS_ENDED : OBS_CONTINUING);
      log_move(id, s.last_selected_move, ident(p ^ 1));
      const bool sub = vget(f.active, V_Substitute);
      applyDamage(p ^ 1, p ^ 1);
      if (stored(f).hp > 0 && !sub) buildRage(p ^ 1);
      return BM_done;
    }

    return vget(a, V_Rage) ? BM_skip_pp : BM_ok;
  }

  bool canMove(int p, int mslot, bool autosel, bool skip_pp, bool& residual, uint8_t from = MOVE_None) {  // :746-795
    Side& s = side(p);
    const MoveData move = move_get(s.last_selected_move);
    if (vget(s.active, V_Charging)) {
      vset(s.active, V_Charging, false);
      vset(s.active, V_Invulnerable, false);
    } else if (move.effect == EFF_Charge) {
      log_move(ident(p), s.last_selected_move, 0, from);
      setCounterable(p, move);
      // Effects.charge :1867-1876
      vset(s.active, V_Charging, true);
      if (s.last_selected_move == MOVE_Fly || s.last_selected_move == MOVE_Dig) vset(s.active, V_Invulnerable, true);
      log_op(LG_LastStill);
      log_id_arg(LG_Prepare, ident(p), s.last_selected_move);
      return false;
    }
    s.last_used_move = s.last_selected_move;
    if (!skip_pp) decrementPP(s, mslot, autosel);
    log_move(ident(p), s.last_selected_move, ident(move.target_self ? p : p ^ 1), from);
    setCounterable(p, move);

    if (eff_onBegin(move.effect)) {
      onBegin(p, move, mslot, residual);
      return false;
    }
    if (move.effect == EFF_Thrashing) {
      // Effects.thrashing :2427-2436
      vset(s.active, V_Thrashing, true);
      vfset(s.active, VF_attacks, roll_duration(p, 2, 4));
      ch_observe(F_attacking, p, OBS_STARTED);
    } else if (move.effect == EFF_Binding) {
      vset(foe(p).active, V_Recharging, false);
    }
    return true;
  }

  void setCounterable(int p, const MoveData& move) {  // :797-806
    const bool c = side(p).last_selected_move != MOVE_Counter && move.bp > 0 &&
                   (move.type == TYPE_Normal || move.type == TYPE_Fighting);
    b.last_moves[p].counterable = c;
  }

  void decrementPP(Side& s, int mslot, bool /*autosel*/) {  // :808-826
    if (s.last_selected_move == MOVE_Struggle) return;
    Active& a = s.active;
    if (vget(a, V_Bide)) return;
    a.moves[mslot - 1].pp = (uint8_t)((a.moves[mslot - 1].pp - 1) & 63);
    if (vget(a, V_Transform)) return;
    MoveSlot& ms = stored(s).moves[mslot - 1];
    ms.pp = (uint8_t)((ms.pp - 1) & 63);
  }

  void incrementPP(Side& s, int p, int mslot) {  // :828-841
    Active& a = s.active;
    a.moves[mslot - 1].pp = (uint8_t)((a.moves[mslot - 1].pp + 1) & 63);
    if (vget(a, V_Transform)) return;
    const uint32_t ov = overridden(p, A_pp);
    const uint32_t n = ov ? ov : 1;
    MoveSlot& ms = stored(s).moves[mslot - 1];
    ms.pp = (uint8_t)((ms.pp + n) & 63);
  }

  // doMove (:844-1102) with mirrorMove (:1351-1376) and metronome (:1378-1396) folded into a loop:
  // the reference recurses canMove -> doMove with last_selected_move rewritten.
  int doMove(int p, int mslot, bool rewrap, bool autosel, bool& residual) {
    for (;;) {
      Side& s = side(p);
      Side& f = foe(p);
      const MoveData move = move_get(s.last_selected_move);
      const bool counter = s.last_selected_move == MOVE_Counter;
      const bool status = move.bp == 0 && move.effect != EFF_OHKO;

      bool crit = false, ohko = false, immune = false, mist = false;
      uint32_t hits = 1;
      uint32_t effectiveness = 100;  // Effectiveness.neutral (data.zig:546-557)

      if (!counter) b.last_damage = 0;
      if (move.effect == EFF_SuperFang || move.effect == EFF_SpecialDamage) return specialDamage(p, move);

      bool miss;
      {
        immune = !move.target_self && !status && !counter &&
                 (type_eff(move.type, f.active.types & 15) == 0 || type_eff(move.type, f.active.types >> 4) == 0);
        if (immune && move.effect != EFF_Binding) {
          miss = true;
        } else if (move.effect == EFF_OHKO && s.active.stats.spe < f.active.stats.spe) {
          b.last_damage = 0;
          miss = true;
        } else {
          miss = !move.target_self && !moveHit(p, move, immune, mist);
        }
      }

      bool late = move.effect != EFF_Explode;
      const bool skip = status || immune;
      if ((!skip || counter) && !miss) {
        if (eff_isMulti(move.effect)) {
          effect_multiHit(p, move);
          hits = vfget(s.active, VF_attacks);
          late = false;
        }
        const bool check = !counter && move.effect != EFF_OHKO;
        if (check) crit = checkCriticalHit(p, move);
        if (counter) return counterDamage(p);

        b.last_damage = 0;
        if (!skip) {
          if (move.effect == EFF_OHKO) {
            ohko = true;
            b.last_damage = 65535;
          } else {
            if (!calcDamage(p, p ^ 1, &move, crit)) return RES_ERROR;
            if (b.last_damage == 0) {
              immune = true;
              effectiveness = 0;
            } else {
              effectiveness = adjustDamage(p);
              immune = effectiveness == 0;
            }
            randomizeDamage(p);
          }
        }
      }

      if (!miss) {
        if (move.effect == EFF_MirrorMove) {
          // mirrorMove :1351-1376
          s.last_selected_move = f.last_used_move;
          if (f.last_used_move == MOVE_None || f.last_used_move == MOVE_MirrorMove) {
            log_fail(ident(p), FAIL_None);
            return NULLR;
          }
          if (f.last_used_move != MOVE_Struggle) incrementPP(s, p, mslot);
          if (!canMove(p, mslot, autosel, false, residual, MOVE_MirrorMove)) return NULLR;
          continue;
        } else if (move.effect == EFF_Metronome) {
          // metronome :1378-1396
          s.last_selected_move = roll_metronome(p);
          incrementPP(s, p, mslot);
          if (!canMove(p, mslot, autosel, false, residual, MOVE_Metronome)) return NULLR;
          continue;
        } else if (eff_onEnd(move.effect)) {
          onEnd(p, move);
          return NULLR;
        }
      }

      if (miss) {
        if (LOG) {  // :951-992
          const uint8_t foe_ident = ident(p ^ 1);
          const bool invulnerable = vget(f.active, V_Invulnerable) && move.effect != EFF_Swift;
          ohko = !immune && !invulnerable && move.effect == EFF_OHKO && s.active.stats.spe < f.active.stats.spe;
          if (ohko) {
            log_id_arg(LG_Immune, foe_ident, IMM_OHKO);
          } else if (immune && !invulnerable && move.effect != EFF_Binding) {
            log_id_arg(LG_Immune, foe_ident, IMM_None);
          } else if (mist) {
            if (!vget(f.active, V_Substitute)) log_activate(foe_ident, ACT_Mist);
            log_fail(foe_ident, FAIL_None);
          } else {
            log_lastmiss_miss(p);
          }
        }
        if (move.effect == EFF_JumpKick) {
          b.last_damage = 1;
          applyDamage(p, p ^ 1);
          if (stored(s).hp == 0) residual = false;
        } else if (move.effect == EFF_Explode) {
          effect_explode(p);
          buildRage(p ^ 1);
        } else if (move.effect == EFF_Disable) {
          buildRage(p ^ 1);
        }
        return NULLR;
      }

      const bool sub = vget(f.active, V_Substitute);
      bool nullified = false;
      uint32_t hit = 0;
      while (hit < hits) {
        if (LOG && hit == 0) {
          if (crit) log_id(LG_Crit, ident(p ^ 1));
          if (effectiveness > 100) {
            log_id(LG_SuperEffective, ident(p ^ 1));
          } else if (effectiveness < 100) {
            log_id(LG_Resisted, ident(p ^ 1));
          }
        }
        if (!skip) nullified = applyDamage(p ^ 1, p ^ 1);
        if (hit == 0 && ohko) log_op(LG_OHKO);
        hit++;
        if (stored(f).hp == 0) break;
        if (!late && (!sub || move.effect == EFF_Explode)) buildRage(p ^ 1);
        if (nullified) break;
      }

      if (vget(s.active, V_MultiHit)) {
        vset(s.active, V_MultiHit, false);
        vfset(s.active, VF_attacks, 0);
        if (move.effect == EFF_Twineedle && !nullified && stored(f).hp > 0) {
          effect_poison(p, move_get(MOVE_PoisonSting));
        }
        log_id_arg(LG_HitCount, ident(p ^ 1), (uint8_t)hit);
      } else if (move.effect == EFF_Binding) {
        effect_binding(p, rewrap);
        if (immune) {
          b.last_damage = 0;
          if (sub) {
            log_activate(ident(p ^ 1), ACT_Substitute);
          } else {
            log_damage(ident(p ^ 1), stored(f), DMG_None);
            buildRage(p ^ 1);
          }
          return NULLR;
        }
      }

      if (nullified) return NULLR;
      if (eff_alwaysHappens(move.effect)) alwaysHappens(p, move, residual);
      if (stored(f).hp == 0) return NULLR;
      if (late && !sub && move.effect != EFF_Disable) buildRage(p ^ 1);

      if (!eff_isSpecial(move.effect)) {
        if (move.effect == EFF_Twineedle) {
          // handled in the MultiHit cleanup above on Pokémon Showdown
        } else if (move.effect == EFF_Disable) {
          effect_disable(p);
          buildRage(p ^ 1);
          return NULLR;
        } else {
          moveEffect(p, move);
        }
      }
      return NULLR;
    }
  }

  bool checkCriticalHit(int p, const MoveData& move) {  // :1104-1116
    Side& s = side(p);
    uint32_t rate = GEN1_SPECIES_CRIT[stored(s).species];
    rate = vget(s.active, V_FocusEnergy) ? rate / 2 : (rate * 2 < 255 ? rate * 2 : 255);
    rate = move.effect == EFF_HighCritical ? (rate * 4 < 255 ? rate * 4 : 255) : rate / 2;
    return roll_criticalHit(p, (uint8_t)rate);
  }

  bool calcDamage(int p, int tp, const MoveData* m, bool crit) {  // :1118-1175
    const bool cfz = m == nullptr;
    const MoveData move = m ? *m : move_get(MOVE_Pound);
    Side& s = side(p);
    Side& t = side(tp);
    const bool special = type_special(move.type);

    uint32_t atk = crit ? (special ? stored(s).stats.spc : stored(s).stats.atk)
                        : (special ? s.active.stats.spc : s.active.stats.atk);
    uint32_t def;
    if (crit) {
      def = special ? stored(t).stats.spc : stored(t).stats.def;
    } else if (special) {
      def = (uint32_t)t.active.stats.spc * (vget(t.active, V_LightScreen) ? 2 : 1);
    } else {
      def = (uint32_t)t.active.stats.def * ((!cfz && vget(t.active, V_Reflect)) ? 2 : 1);
    }
    if (!cfz && (atk > 255 || def > 255)) {
      atk = ((atk / 4) & 255) > 1 ? ((atk / 4) & 255) : 1;
      def = ((def / 4) & 255) > 1 ? ((def / 4) & 255) : 1;
    }
    const uint32_t lvl = (uint32_t)stored(s).level * (crit ? 2 : 1);
    if (move.effect == EFF_Explode) def = def / 2 > 1 ? def / 2 : 1;
    if (def == 0) return false;

    uint32_t d = (lvl * 2 / 5) + 2;
    d *= move.bp;
    d *= atk;
    d /= def;
    d /= 50;
    if (d > 997) d = 997;
    d += 2;
    b.last_damage = (uint16_t)d;
    return true;
  }

  uint16_t adjustDamage(int p) {  // :1197-1226
    Side& s = side(p);
    const uint8_t types = foe(p).active.types;
    const MoveData move = move_get(s.last_selected_move);
    uint16_t d = b.last_damage;
    if (types_includes(s.active.types, move.type)) d = (uint16_t)(d + d / 2);
    const uint16_t eff1 = type_eff(move.type, types & 15), eff2 = type_eff(move.type, types >> 4);
    if (eff1 != 10) d = (uint16_t)((uint16_t)(d * eff1) / 10);
    if ((types & 15) != (types >> 4) && eff2 != 10) d = (uint16_t)((uint16_t)(d * eff2) / 10);
    b.last_damage = d;
    return (types & 15) == (types >> 4) ? (uint16_t)(eff1 * 10) : (uint16_t)(eff1 * eff2);
  }

  void randomizeDamage(int p) {  // :1228-1234
    if (b.last_damage <= 1) return;
    const uint32_t random = roll_damage(p);
    b.last_damage = (uint16_t)((uint32_t)b.last_damage * random / 255);
  }

  int specialDamage(int p, const MoveData& move) {  // :1236-1264
    Side& s = side(p);
    Side& f = foe(p);
    if (!checkHit(p, move)) return NULLR;
    switch (s.last_selected_move) {
      case MOVE_SuperFang: b.last_damage = (uint16_t)(stored(f).hp / 2 > 1 ? stored(f).hp / 2 : 1); break;
      case MOVE_SeismicToss:
      case MOVE_NightShade: b.last_damage = stored(s).level; break;
      case MOVE_SonicBoom: b.last_damage = 20; break;
      case MOVE_DragonRage: b.last_damage = 40; break;
      case MOVE_Psywave: {
        const uint8_t max = (uint8_t)((uint16_t)stored(s).level * 3 / 2);
        b.last_damage = roll_psywave(p, max);
        break;
      }
      default: ORACLE_ASSERT(false);
    }
    if (b.last_damage == 0) return NULLR;
    const bool sub = vget(f.active, V_Substitute);
    applyDamage(p ^ 1, p ^ 1);
    if (stored(f).hp > 0 && !sub) buildRage(p ^ 1);
    return NULLR;
  }

  int counterDamage(int p) {  // :1266-1312
    Side& f = foe(p);
    if (b.last_damage == 0) {
      log_fail(ident(p), FAIL_None);
      return NULLR;
    }
    const uint8_t fm = (f.last_selected_move == MOVE_None || f.last_selected_move == MOVE_SKIP_TURN)
                           ? (uint8_t)MOVE_Splash : f.last_selected_move;
    const MoveData fmd = move_get(fm);
    const bool used = b.last_moves[p ^ 1].counterable != 0;
    const bool selected = fmd.bp > 0 && f.last_selected_move != MOVE_Counter &&
                          (fmd.type == TYPE_Normal || fmd.type == TYPE_Fighting);
    if (!used || !selected) {  // both :1283-1286 and the Desync Clause Mod case :1288-1295
      log_fail(ident(p), FAIL_None);
      return NULLR;
    }
    b.last_damage = b.last_damage > 32767 ? 65535 : (uint16_t)(b.last_damage * 2);
    const bool sub = vget(f.active, V_Substitute);
    applyDamage(p ^ 1, p ^ 1);
    if (stored(f).hp > 0 && !sub) buildRage(p ^ 1);
    return NULLR;
  }

  bool applyDamage(int tp, int sp, uint8_t reason = DMG_None) {  // :1314-1349
    Side& t = side(tp);
    if (vget(t.active, V_Substitute)) {
      Side& sb = side(sp);
      if (!vget(sb.active, V_Substitute)) return false;
      if (b.last_damage >= vfget(sb.active, VF_substitute)) {
        vfset(sb.active, VF_substitute, 0);
        vset(sb.active, V_Substitute, false);
        log_end(ident(sp), END_Substitute);
        return true;
      }
      vfset(sb.active, VF_substitute, vfget(sb.active, VF_substitute) - b.last_damage);
      log_activate(ident(sp), ACT_Substitute);
      return false;
    }
    Pokemon& st = stored(t);
    if (b.last_damage > st.hp) b.last_damage = st.hp;
    st.hp -= b.last_damage;
    log_damage(ident(tp), st, reason);
    return false;
  }

  bool checkHit(int p, const MoveData& move) {  // :1398-1420
    bool immune = false, mist = false;
    if (moveHit(p, move, immune, mist)) return true;
    // `mist` is only ever raised for stat-down moves, which do not come through checkHit in showdown mode (:1408-1412)
    log_lastmiss_miss(p);
    return false;
  }

  bool moveHit(int p, const MoveData& move, bool& immune, bool& mist) {  // :1422-1505
    Side& s = side(p);
    Side& f = foe(p);
    bool miss;
    do {
      if (move.effect == EFF_Swift) return true;
      if (vget(f.active, V_Invulnerable)) {
        miss = true;
        break;
      }
      if (move.effect == EFF_DreamEater && (!status_is(stored(f).status, ST_SLP) || vget(f.active, V_Substitute))) {
        immune = true;
        return false;
      }
      if (move.effect == EFF_Sleep && vget(f.active, V_Recharging)) return true;
      if (vget(f.active, V_Mist) && eff_isStatDown(move.effect)) mist = true;

      const uint32_t state = vfget(s.active, VF_state);
      bool overwrite = move.effect == EFF_Rage || move.effect == EFF_Thrashing;
      const bool overwritten = overwrite && state > 0;
      uint32_t accuracy = overwritten ? state : move.accuracy;
      const uint8_t* bo = BOOSTS[bget(s.active, B_accuracy) + 6];
      accuracy = (uint16_t)(accuracy * bo[0] / bo[1]);
      bo = BOOSTS[-bget(f.active, B_evasion) + 6];
      accuracy = (uint16_t)(accuracy * bo[0] / bo[1]);
      accuracy = accuracy > 255 ? 255 : (accuracy < 1 ? 1 : accuracy);
      overwrite = vget(s.active, V_Rage) || vget(s.active, V_Thrashing);
      if (overwrite) vfset(s.active, VF_state, accuracy);
      miss = !roll_hit(p, (uint8_t)accuracy);
    } while (false);

    if (mist) {
      mist = !miss;
      miss = true;
    }
    if (!miss) return true;
    if (!vget(f.active, V_Invulnerable)) b.last_damage = 0;
    vset(s.active, V_Binding, false);
    ch_observe(F_binding, p, OBS_NONE);
    return false;
  }

  int checkFaint(int p) {  // :1507-1543
    Side& s = side(p);
    if (stored(s).hp > 0) return NULLR;
    Side& f = foe(p);
    const bool foe_fainted = stored(f).hp == 0;
    const bool player_out = findFirstAlive(s) == 0;
    const bool foe_out = foe_fainted && findFirstAlive(f) == 0;
    const bool more = player_out || foe_out;
    faint(p, !(more || foe_fainted));
    if (foe_fainted) faint(p ^ 1, !more);
    if (player_out && foe_out) {
      log_tie();
      return RES_TIE;
    }
    if (player_out) {
      log_win(p ^ 1);
      return p == P1 ? RES_LOSE : RES_WIN;
    }
    if (foe_out) {
      log_win(p);
      return p == P1 ? RES_WIN : RES_LOSE;
    }
    const int foe_choice = foe_fainted ? CH_SWITCH : CH_PASS;
    if (p == P1) return (CH_SWITCH << 4) | (foe_choice << 6);
    return (foe_choice << 4) | (CH_SWITCH << 6);
  }

  void faint(int p, bool done) {  // :1545-1583
    Side& s = side(p);
    Side& f = foe(p);
    if (vget(f.active, V_Bide)) vfset(f.active, VF_state, 0);
    s.active.volatiles = 0;
    s.last_used_move = MOVE_None;
    f.last_used_move = MOVE_None;
    Pokemon& st = stored(s);
    const uint8_t status = st.status;
    st.status = 0;
    s.active.stats.spe = status_is(status, ST_PAR) ? (uint16_t)(st.stats.spe / 4 > 1 ? st.stats.spe / 4 : 1) : st.stats.spe;
    log_faint(ident(p), done);
  }

  void handleResidual(int p) {  // :1585-1637
    Side& s = side(p);
    Pokemon& st = stored(s);
    Active& a = s.active;
    const bool brn = status_is(st.status, ST_BRN);
    if (brn || status_is(st.status, ST_PSN)) {
      uint32_t damage = st.stats.hp / 16 > 1 ? st.stats.hp / 16 : 1;
      if (vget(a, V_Toxic)) {
        vfset(a, VF_toxic, vfget(a, VF_toxic) + 1);
        damage *= vfget(a, VF_toxic);
      }
      if (damage > st.hp) damage = st.hp;
      if (damage > 0) {
        st.hp -= (uint16_t)damage;
        log_damage(ident(p), st, brn ? DMG_Burn : DMG_Poison);
      }
    }
    if (vget(a, V_LeechSeed)) {
      Pokemon& fs = stored(foe(p));
      if (fs.hp == 0) return;
      uint32_t damage = st.stats.hp / 16 > 1 ? st.stats.hp / 16 : 1;
      if (vget(a, V_Toxic)) {
        vfset(a, VF_toxic, vfget(a, VF_toxic) + 1);
        damage *= vfget(a, VF_toxic);
      }
      const uint32_t amount = damage < st.hp ? damage : st.hp;
      st.hp -= (uint16_t)amount;
      if (amount > 0) log_damage(ident(p), st, DMG_LeechSeed);
      const uint16_t before = fs.hp;
      const uint32_t h = fs.hp + damage;
      fs.hp = (uint16_t)(h < fs.stats.hp ? h : fs.stats.hp);
      if (fs.hp > before) log_heal(ident(p ^ 1), fs, HEAL_Silent);
    }
  }

  uint8_t endTurn() {  // :1639-1660
    b.turn += 1;
    if ((EBC && checkEBC()) || b.turn >= 1000) {
      log_tie();
      return RES_TIE;
    }
    log_turn(b.turn);
    return RESULT_DEFAULT;
  }

  static bool only_transform(const MoveSlot* ms) {
    for (int i = 0; i < 4; i++) {
      if (ms[i].id == MOVE_None) return true;
      if (ms[i].id != MOVE_Transform) return false;
    }
    return true;
  }

  bool checkEBC() {  // :1662-1714
    for (int i = 0; i < 2; i++) {
      const Side& s = b.sides[i];
      const Side& f = b.sides[i ^ 1];
      bool foe_all_ghosts = true, foe_all_transform = true;
      for (int j = 0; j < 6; j++) {
        const uint8_t id = f.order[j];
        if (id == 0) break;
        const bool active = j == 0;
        const Pokemon& pk = f.pokemon[id - 1];
        const bool ghost = pk.hp == 0 || types_includes(active ? f.active.types : pk.types, TYPE_Ghost);
        foe_all_ghosts = foe_all_ghosts && ghost;
        // Zig precedence: (foe_all_transform and pk.hp == 0) or transform
        foe_all_transform = (foe_all_transform && pk.hp == 0) || only_transform(active ? f.active.moves : pk.moves);
      }
      bool all_continue = true;
      for (int j = 0; j < 6; j++) {
        const uint8_t id = s.order[j];
        if (id == 0) break;
        const bool active = j == 0;
        const Pokemon& pk = s.pokemon[id - 1];
        if (pk.hp == 0 || status_is(pk.status, ST_FRZ)) continue;
        const MoveSlot* ms = active ? s.active.moves : pk.moves;
        if (foe_all_transform && only_transform(ms)) continue;
        bool no_pp = foe_all_ghosts;
        if (no_pp) {
          for (int k = 0; k < 4; k++)
            if (ms[k].pp != 0) no_pp = false;
        }
        if (no_pp) continue;
        all_continue = false;
        break;
      }
      if (!all_continue) return false;
      if (i == 1) return true;
    }
    return false;
  }

  bool disabled(Side& s, uint8_t id) {  // :1716-1732
    const uint32_t dm = vfget(s.active, VF_disable_move);
    if (dm != 0) {
      const uint8_t m = s.active.moves[dm - 1].id;
      const uint8_t last = (m == MOVE_Bide && vget(s.active, V_Bide)) ? m : s.last_selected_move;
      if (m != MOVE_None && m == last) {
        vset(s.active, V_Charging, false);
        log_cant_disable(id, last);
        return true;
      }
    }
    return false;
  }

  void buildRage(int who) {  // :1734-1739
    Side& s = side(who);
    if (vget(s.active, V_Rage) && bget(s.active, B_atk) < 6) effect_boost(who, move_get(MOVE_Rage));
  }

  // ------------------------------------------------------------------ effect dispatch (:1741-1819)
  void onBegin(int p, const MoveData& move, int mslot, bool& residual) {
    switch (move.effect) {
      case EFF_Confusion: effect_confusion(p, move); break;
      case EFF_Conversion: effect_conversion(p); break;
      case EFF_FocusEnergy:  // :2004-2011
        if (!vget(side(p).active, V_FocusEnergy)) {
          vset(side(p).active, V_FocusEnergy, true);
          log_start(ident(p), START_FocusEnergy);
        }
        break;
      case EFF_Haze: effect_haze(p); break;
      case EFF_Heal: effect_heal(p); break;
      case EFF_LeechSeed: effect_leechSeed(p, move); break;
      case EFF_LightScreen:  // :2136-2144
        if (vget(side(p).active, V_LightScreen)) {
          log_fail(ident(p), FAIL_None);
        } else {
          vset(side(p).active, V_LightScreen, true);
          log_start(ident(p), START_LightScreen);
        }
        break;
      case EFF_Mimic: effect_mimic(p, move, mslot); break;
      case EFF_Mist:  // :2182-2189
        if (!vget(side(p).active, V_Mist)) {
          vset(side(p).active, V_Mist, true);
          log_start(ident(p), START_Mist);
        }
        break;
      case EFF_Paralyze: effect_paralyze(p, move); break;
      case EFF_Poison: effect_poison(p, move); break;
      case EFF_Reflect:  // :2337-2345
        if (vget(side(p).active, V_Reflect)) {
          log_fail(ident(p), FAIL_None);
        } else {
          vset(side(p).active, V_Reflect, true);
          log_start(ident(p), START_Reflect);
        }
        break;
      case EFF_Splash: log_activate(ident(p), ACT_Splash); break;
      case EFF_Substitute: effect_substitute(p, residual); break;
      case EFF_SwitchAndTeleport: effect_switchAndTeleport(p, move); break;
      case EFF_Transform: effect_transform(p); break;
      default: ORACLE_ASSERT(false);
    }
  }
  void onEnd(int p, const MoveData& move) {
    switch (move.effect) {
      case EFF_Bide: {  // Effects.bide :1822-1833
        Active& a = side(p).active;
        vset(a, V_Bide, true);
        vfset(a, VF_state, 0);
        vfset(a, VF_attacks, roll_duration(p, 2, 4));
        ch_observe(F_attacking, p, OBS_STARTED);
        log_start(ident(p), START_Bide);
        break;
      }
      case EFF_AttackUp1: case EFF_AttackUp2: case EFF_DefenseUp1: case EFF_DefenseUp2:
      case EFF_EvasionUp1: case EFF_SpecialUp1: case EFF_SpecialUp2: case EFF_SpeedUp2:
        effect_boost(p, move);
        break;
      case EFF_AccuracyDown1: case EFF_AttackDown1: case EFF_DefenseDown1: case EFF_DefenseDown2:
      case EFF_SpeedDown1:
        effect_unboost(p, move);
        break;
      case EFF_Sleep: effect_sleep(p); break;
      default: ORACLE_ASSERT(false);
    }
  }
  void alwaysHappens(int p, const MoveData& move, bool& residual) {
    switch (move.effect) {
      case EFF_DrainHP: case EFF_DreamEater: effect_drainHP(p); break;
      case EFF_Explode: effect_explode(p); break;
      case EFF_PayDay:  // :2257-2261
        if (!vget(foe(p).active, V_Substitute)) log_op(LG_FieldActivate);
        break;
      case EFF_Rage: vset(side(p).active, V_Rage, true); break;
      case EFF_Recoil: effect_recoil(p, residual); break;
      default: break;  // JumpKick, Binding
    }
  }
  void moveEffect(int p, const MoveData& move) {
    switch (move.effect) {
      case EFF_BurnChance1: case EFF_BurnChance2: effect_burnChance(p, move); break;
      case EFF_ConfusionChance: effect_confusion(p, move); break;
      case EFF_FlinchChance1: case EFF_FlinchChance2: effect_flinchChance(p, move); break;
      case EFF_FreezeChance: effect_freezeChance(p, move); break;
      case EFF_HyperBeam:  // :2108-2111
        vset(side(p).active, V_Recharging, true);
        log_id(LG_MustRecharge, ident(p));
        break;
      case EFF_ParalyzeChance1: case EFF_ParalyzeChance2: effect_paralyzeChance(p, move); break;
      case EFF_PoisonChance1: case EFF_PoisonChance2: effect_poison(p, move); break;
      case EFF_AttackDownChance: case EFF_DefenseDownChance: case EFF_SpecialDownChance: case EFF_SpeedDownChance:
        effect_unboost(p, move);
        break;
      default: break;
    }
  }

  // ------------------------------------------------------------------ Effects (:1821-2690)
  void effect_burnChance(int p, const MoveData& move) {  // :1835-1865
    Side& f = foe(p);
    Pokemon& fs = stored(f);
    if (vget(f.active, V_Substitute)) return;
    if (fs.status != 0) {
      if (!types_includes(f.active.types, move.type)) advance(1);
      if (status_is(fs.status, ST_FRZ)) {
        log_curestatus(ident(p ^ 1), fs.status, CURE_Message);
        fs.status = 0;
      }
      return;
    }
    if (types_includes(f.active.types, move.type)) return;
    if (!roll_secondary(p, move.effect == EFF_BurnChance1 ? 26 : 77)) return;
    fs.status = status_init(ST_BRN);
    f.active.stats.atk = (uint16_t)(f.active.stats.atk / 2 > 1 ? f.active.stats.atk / 2 : 1);
    log_status(ident(p ^ 1), fs.status, STS_None);
  }
  void effect_confusion(int p, const MoveData& move) {  // :1878-1902
    Side& f = foe(p);
    const bool sub = vget(f.active, V_Substitute);
    if (move.effect == EFF_ConfusionChance) {
      if (!roll_secondary(p, 25)) return;
    } else {
      if (!checkHit(p, move)) return;
      if (sub) {
        log_fail(ident(p ^ 1), FAIL_None);
        return;
      }
    }
    if (vget(f.active, V_Confusion)) return;
    vset(f.active, V_Confusion, true);
    vfset(f.active, VF_confusion, roll_duration(p, 2, 6));
    ch_observe(F_confusion, p ^ 1, OBS_STARTED);
    calc_confusion(p ^ 1);
    log_start(ident(p ^ 1), START_Confusion);
  }
  void effect_conversion(int p) {  // :1904-1919
    Side& f = foe(p);
    if (vget(f.active, V_Invulnerable)) {
      log_lastmiss_miss(p);
      return;
    }
    side(p).active.types = f.active.types;
    log_start_typechange(ident(p), f.active.types, ident(p ^ 1));
  }
  void effect_disable(int p) {  // :1921-1961
    Side& f = foe(p);
    Active& fa = f.active;
    if (vfget(fa, VF_disable_move) != 0) {
      log_fail(ident(p ^ 1), FAIL_None);
      return;
    }
    uint32_t n = 0;
    bool err = true;
    for (int i = 0; i < 4; i++) {
      if (fa.moves[i].pp > 0) {
        n++;
        if (fa.moves[i].id != MOVE_None) err = false;
      }
    }
    if (n == 0 || err) {
      log_fail(ident(p ^ 1), FAIL_None);
      return;
    }
    vfset(fa, VF_disable_move, roll_moveSlot(p, fa.moves, n));
    vfset(fa, VF_disable_duration, roll_duration(p, 1, 9));
    ch_observe(F_disable, p ^ 1, OBS_STARTED);
    log_start_move(ident(p ^ 1), START_Disable, fa.moves[vfget(fa, VF_disable_move) - 1].id);
  }
  void effect_drainHP(int p) {  // :1963-1979
    Pokemon& st = stored(side(p));
    if (b.last_damage == 0) {
      if (!vget(foe(p).active, V_Substitute)) return;
    }
    const uint16_t drain = (uint16_t)(b.last_damage / 2 > 1 ? b.last_damage / 2 : 1);
    b.last_damage = drain;
    if (st.hp == st.stats.hp) return;
    const uint32_t h = (uint32_t)st.hp + drain;
    st.hp = (uint16_t)(h < st.stats.hp ? h : st.stats.hp);
    log_heal(ident(p), st, HEAL_Drain, ident(p ^ 1));
  }
  void effect_explode(int p) {  // :1981-1990
    Side& s = side(p);
    stored(s).hp = 0;
    vset(s.active, V_LeechSeed, false);
  }
  void effect_flinchChance(int p, const MoveData& move) {  // :1992-2002
    Active& fa = foe(p).active;
    if (vget(fa, V_Substitute)) return;
    if (!roll_secondary(p, move.effect == EFF_FlinchChance1 ? 26 : 77)) return;
    vset(fa, V_Flinch, true);
    vset(fa, V_Recharging, false);
  }
  void effect_freezeChance(int p, const MoveData& move) {  // :2013-2033
    Side& f = foe(p);
    Pokemon& fs = stored(f);
    if (vget(f.active, V_Substitute)) return;
    if (fs.status != 0) {
      if (!types_includes(f.active.types, move.type)) advance(1);
      return;
    }
    if (types_includes(f.active.types, move.type)) return;
    if (!roll_secondary(p, 26)) return;
    for (int i = 0; i < 6; i++)
      if (status_is(f.pokemon[i].status, ST_FRZ)) return;  // Freeze Clause Mod
    fs.status = status_init(ST_FRZ);
    log_status(ident(p ^ 1), fs.status, STS_None);
  }
  void effect_haze(int p) {  // :2035-2082
    Side& s = side(p);
    Side& f = foe(p);
    s.active.boosts = 0;
    f.active.boosts = 0;
    s.active.stats = stored(s).stats;
    f.active.stats = stored(f).stats;
    log_activate(ident(p), ACT_Haze);
    log_op(LG_ClearAllBoost);
    for (int i = 0; i < 2; i++) {
      Pokemon& st = stored(b.sides[i]);
      if (i != p && st.status != 0) {
        log_curestatus(ident(p ^ 1), st.status, CURE_Silent);
        st.status = 0;
        ch_observe(F_sleep, i, OBS_NONE);
      } else if (st.status == STATUS_TOX) {
        st.status = status_init(ST_PSN);
        log_status(ident(i), st.status, STS_None);
      }
      clearVolatiles(i);
    }
  }
  void effect_heal(int p) {  // :2084-2106
    Side& s = side(p);
    Pokemon& st = stored(s);
    const uint16_t delta = (uint16_t)(st.stats.hp - st.hp);
    if (delta == 0 || ((delta & 255) == 255 && st.hp % 256 != 0)) {
      log_fail(ident(p), FAIL_None);
      return;
    }
    const bool rest = s.last_selected_move == MOVE_Rest;
    if (rest) {
      advance(1);
      st.status = 0x80 | 2;
      log_status(ident(p), st.status, STS_From, MOVE_Rest);
      st.hp = st.stats.hp;
    } else {
      const uint32_t h = (uint32_t)st.hp + st.stats.hp / 2;
      st.hp = (uint16_t)(h < st.stats.hp ? h : st.stats.hp);
    }
    log_heal(ident(p), st, rest ? HEAL_Silent : HEAL_None);
  }
  void effect_leechSeed(int p, const MoveData& move) {  // :2113-2134
    Side& f = foe(p);
    if (!vget(f.active, V_Invulnerable) && types_includes(f.active.types, TYPE_Grass)) {
      log_id_arg(LG_Immune, ident(p ^ 1), IMM_None);
      return;
    }
    if (!checkHit(p, move)) return;
    if (vget(f.active, V_LeechSeed)) return;
    vset(f.active, V_LeechSeed, true);
    log_start(ident(p ^ 1), START_LeechSeed);
  }
  void effect_mimic(int p, const MoveData& move, int mslot) {  // :2146-2180
    Side& s = side(p);
    Side& f = foe(p);
    int oslot = mslot;
    bool has_mimic = false;
    for (int i = 0; i < 4; i++) {
      if (s.active.moves[i].id == MOVE_Mimic) {
        oslot = i + 1;
        has_mimic = true;
        break;
      }
    }
    if (!has_mimic) {
      if (!checkHit(p, move)) return;
      log_fail(ident(p ^ 1), FAIL_None);
      return;
    }
    if (!checkHit(p, move)) return;
    const uint32_t rslot = roll_moveSlot(p, f.active.moves, 0);
    s.active.moves[oslot - 1].id = f.active.moves[rslot - 1].id;
    log_start_move(ident(p), START_Mimic, s.active.moves[oslot - 1].id);
  }
  void effect_multiHit(int p, const MoveData& move) {  // :2191-2201
    Active& a = side(p).active;
    vset(a, V_MultiHit, true);
    vfset(a, VF_attacks, move.effect == EFF_MultiHit ? roll_distribution(p, false) : 2);
  }
  void effect_paralyze(int p, const MoveData& move) {  // :2203-2233
    Side& f = foe(p);
    Pokemon& fs = stored(f);
    const bool immune = move.type == TYPE_Electric &&
                        (type_eff(move.type, f.active.types & 15) == 0 || type_eff(move.type, f.active.types >> 4) == 0);
    if (immune && !vget(f.active, V_Invulnerable)) {
      log_id_arg(LG_Immune, ident(p ^ 1), IMM_None);
      return;
    }
    if (!checkHit(p, move)) return;
    if (fs.status != 0) {
      log_fail(ident(p ^ 1), status_is(fs.status, ST_PAR) ? FAIL_Paralysis : FAIL_None);
      return;
    }
    fs.status = status_init(ST_PAR);
    f.active.stats.spe = (uint16_t)(f.active.stats.spe / 4 > 1 ? f.active.stats.spe / 4 : 1);
    log_status(ident(p ^ 1), fs.status, STS_None);
  }
  void effect_paralyzeChance(int p, const MoveData& move) {  // :2235-2255
    Side& f = foe(p);
    Pokemon& fs = stored(f);
    if (vget(f.active, V_Substitute)) return;
    if (fs.status != 0) {
      if (!types_includes(f.active.types, move.type)) advance(1);
      return;
    }
    if (types_includes(f.active.types, move.type)) return;
    if (!roll_secondary(p, move.effect == EFF_ParalyzeChance1 ? 26 : 77)) return;
    fs.status = status_init(ST_PAR);
    f.active.stats.spe = (uint16_t)(f.active.stats.spe / 4 > 1 ? f.active.stats.spe / 4 : 1);
    log_status(ident(p ^ 1), fs.status, STS_None);
  }
  void effect_poison(int p, const MoveData& move) {  // :2263-2309
    Side& f = foe(p);
    Pokemon& fs = stored(f);
    const bool toxic = side(p).last_selected_move == MOVE_Toxic;
    if (move.effect == EFF_Poison && !checkHit(p, move)) {
      return;
    } else if (vget(f.active, V_Substitute)) {
      if (move.effect == EFF_Poison) log_fail(ident(p ^ 1), FAIL_None);
      return;
    } else if (fs.status != 0) {
      if (move.effect != EFF_Poison) {
        advance(1);
        return;
      }
      uint8_t why = FAIL_None;  // :2276-2286
      if (status_is(fs.status, ST_PSN) && toxic == (fs.status == STATUS_TOX)) why = toxic ? FAIL_Toxic : FAIL_Poison;
      log_fail(ident(p ^ 1), why);
      return;
    } else if (types_includes(f.active.types, TYPE_Poison)) {
      if (move.effect != EFF_Poison) {
        advance(1);
        return;
      }
      log_id_arg(LG_Immune, ident(p ^ 1), IMM_None);
      return;
    }
    if (move.effect != EFF_Poison) {
      if (!roll_secondary(p, move.effect == EFF_PoisonChance1 ? 52 : 103)) return;
    }
    fs.status = status_init(ST_PSN);
    if (toxic) {
      fs.status = STATUS_TOX;
      vset(f.active, V_Toxic, true);
      vfset(f.active, VF_toxic, 0);
    }
    log_status(ident(p ^ 1), fs.status, STS_None);
  }
  void effect_recoil(int p, bool& residual) {  // :2317-2335
    Side& s = side(p);
    Pokemon& st = stored(s);
    if (b.last_damage == 0) return;
    const int div = s.last_selected_move == MOVE_Struggle ? 2 : 4;
    const int damage = b.last_damage / div > 1 ? b.last_damage / div : 1;
    const int hp = (int)st.hp - damage;
    st.hp = (uint16_t)(hp > 0 ? hp : 0);
    log_damage(ident(p), st, DMG_RecoilOf, ident(p ^ 1));
    if (st.hp == 0) residual = false;
  }
  void effect_sleep(int p) {  // :2347-2380
    Side& f = foe(p);
    Pokemon& fs = stored(f);
    if (!vget(f.active, V_Recharging)) {
      if (fs.status != 0) {
        log_fail(ident(p ^ 1), status_is(fs.status, ST_SLP) ? FAIL_Sleep : FAIL_None);
        return;
      }
    }
    for (int i = 0; i < 6; i++) {
      const uint8_t st = f.pokemon[i].status;
      if (status_is(st, ST_SLP) && !status_is(st, ST_EXT)) return;  // Sleep Clause Mod
    }
    vset(f.active, V_Recharging, false);
    fs.status = (uint8_t)roll_duration(p, 1, 8);
    ch_observe(F_sleep, p ^ 1, OBS_STARTED);
    log_status(ident(p ^ 1), fs.status, STS_From, side(p).last_selected_move);
  }
  void effect_substitute(int p, bool& residual) {  // :2386-2413
    Side& s = side(p);
    Pokemon& st = stored(s);
    if (vget(s.active, V_Substitute)) {
      log_fail(ident(p), FAIL_Substitute);
      return;
    }
    const uint8_t hp = (uint8_t)(st.stats.hp / 4);
    const uint16_t required = (uint16_t)((st.stats.hp - 1) / 4 + 1);
    if (st.hp < required) {
      log_fail(ident(p), FAIL_Weak);
      return;
    }
    st.hp -= hp;
    vfset(s.active, VF_substitute, (uint32_t)hp + 1);
    vset(s.active, V_Substitute, true);
    log_start(ident(p), START_Substitute);
    if (hp > 0) log_damage(ident(p), st, DMG_None);
    if (hp > 0 && st.hp == 0) residual = false;
  }
  void effect_switchAndTeleport(int p, const MoveData& move) {  // :2415-2425
    if (side(p).last_selected_move == MOVE_Teleport) return;
    (void)checkHit(p, move);
    b.last_damage = 0;
  }
  void effect_transform(int p) {  // :2438-2465
    Side& s = side(p);
    Side& f = foe(p);
    vset(s.active, V_Transform, true);
    const uint32_t ft = vfget(f.active, VF_transform);
    vfset(s.active, VF_transform, ft != 0 ? ft : ident(p ^ 1));
    s.active.stats.atk = f.active.stats.atk;
    s.active.stats.def = f.active.stats.def;
    s.active.stats.spe = f.active.stats.spe;
    s.active.stats.spc = f.active.stats.spc;
    s.active.species = f.active.species;
    s.active.types = f.active.types;
    s.active.boosts = f.active.boosts;
    for (int i = 0; i < 4; i++) {
      s.active.moves[i].id = f.active.moves[i].id;
      s.active.moves[i].pp = f.active.moves[i].id != MOVE_None ? 5 : 0;
    }
    log_id_arg(LG_Transform, ident(p), ident(p ^ 1));
  }
  void effect_binding(int p, bool rewrap) {  // :2467-2482
    Side& s = side(p);
    Side& f = foe(p);
    if (vget(s.active, V_Binding)) return;
    vset(s.active, V_Binding, true);
    if (stored(f).hp == 0) {
      if (!rewrap) advance(1);
      return;
    }
    vfset(s.active, VF_attacks, roll_distribution(p, true) - 1);
    ch_observe(F_binding, p, OBS_STARTED);
  }

  Stats& unmodifiedStats(Side& s) {  // :2692-2696
    if (!vget(s.active, V_Transform)) return stored(s).stats;
    const uint32_t id = vfget(s.active, VF_transform);
    return b.sides[id >> 3].pokemon[(id & 7) - 1].stats;
  }

  static uint16_t* stat_ptr(Stats& st, int which) {
    switch (which) {
      case B_atk: return &st.atk;
      case B_def: return &st.def;
      case B_spe: return &st.spe;
      default: return &st.spc;
    }
  }

  void effect_boost(int p, const MoveData& move) {  // :2484-2582
    Side& s = side(p);
    Active& a = s.active;
    int which, n;
    switch (move.effect) {
      case EFF_AttackUp1: case EFF_Rage: which = B_atk; n = 1; break;
      case EFF_AttackUp2: which = B_atk; n = 2; break;
      case EFF_DefenseUp1: which = B_def; n = 1; break;
      case EFF_DefenseUp2: which = B_def; n = 2; break;
      case EFF_SpeedUp2: which = B_spe; n = 2; break;
      case EFF_SpecialUp1: which = B_spc; n = 1; break;
      case EFF_SpecialUp2: which = B_spc; n = 2; break;
      default: which = B_evasion; n = 1; break;  // EvasionUp1
    }
    const uint8_t id = ident(p);
    if (bget(a, which) == 6) {
      log_fail(id, FAIL_None);
      return;
    }
    const int nb = bget(a, which) + n;
    bset(a, which, nb > 6 ? 6 : nb);
    // log reason(s): Special raises SpecialAttack and SpecialDefense, Rage has its own (:2499,2558-2569)
    static const uint8_t REASON[6] = {BST_Attack, BST_Defense, BST_Speed, BST_SpecialAttack, BST_Accuracy, BST_Evasion};
    const uint8_t reason = move.effect == EFF_Rage ? (uint8_t)BST_Rage : REASON[which];
    if (which != B_evasion) {
      uint16_t* stat = stat_ptr(a.stats, which);
      if (*stat == 999) {
        bset(a, which, bget(a, which) - 1);
        log_boost(id, reason, n);
        log_boost(id, REASON[which], -1);
        if (which == B_spc) {
          log_boost(id, BST_SpecialDefense, n);
          log_boost(id, BST_SpecialDefense, -1);
        }
        if (move.effect != EFF_Rage) log_fail(id, FAIL_None);
        return;
      }
      const uint8_t* mod = BOOSTS[bget(a, which) + 6];
      const uint32_t base = *stat_ptr(unmodifiedStats(s), which);
      const uint32_t v = base * mod[0] / mod[1];
      *stat = (uint16_t)(v < 999 ? v : 999);
      log_boost(id, reason, n);
      if (which == B_spc) log_boost(id, BST_SpecialDefense, n);
      if (move.effect == EFF_Rage) return;
    } else {
      log_boost(id, BST_Evasion, 1);
    }
    Side& f = foe(p);
    statusModify(stored(f).status, f.active.stats);
  }

  void effect_unboost(int p, const MoveData& move) {  // :2584-2689
    Side& f = foe(p);
    Active& fa = f.active;
    const bool secondary = eff_isStatDownChance(move.effect);
    const uint8_t id = ident(p ^ 1);
    if (vget(fa, V_Substitute)) {
      if (!secondary) log_fail(id, FAIL_None);
      return;
    }
    if (secondary) {
      const bool proc = roll_secondary(p, 85);
      if (!proc || vget(fa, V_Invulnerable)) return;
    }
    int which, n = 1;
    switch (move.effect) {
      case EFF_AttackDown1: case EFF_AttackDownChance: which = B_atk; break;
      case EFF_DefenseDown1: case EFF_DefenseDownChance: which = B_def; break;
      case EFF_DefenseDown2: which = B_def; n = 2; break;
      case EFF_SpeedDown1: case EFF_SpeedDownChance: which = B_spe; break;
      case EFF_SpecialDownChance: which = B_spc; break;
      default: which = B_accuracy; break;  // AccuracyDown1
    }
    if (bget(fa, which) == -6) {
      log_fail(id, FAIL_None);  // `fail` = showdown or !secondary (:2603)
      return;
    }
    const int nb = bget(fa, which) - n;
    bset(fa, which, nb < -6 ? -6 : nb);
    static const uint8_t REASON[6] = {BST_Attack, BST_Defense, BST_Speed, BST_SpecialAttack, BST_Accuracy, BST_Evasion};
    if (which != B_accuracy) {
      uint16_t* stat = stat_ptr(fa.stats, which);
      if (*stat == 1) {
        bset(fa, which, bget(fa, which) + 1);
        log_boost(id, REASON[which], -n);
        log_boost(id, REASON[which], 1);
        if (which == B_spc) {
          log_boost(id, BST_SpecialDefense, -1);
          log_boost(id, BST_SpecialDefense, 1);
        }
        log_fail(id, FAIL_None);
        return;
      }
      const uint8_t* mod = BOOSTS[bget(fa, which) + 6];
      const uint32_t base = *stat_ptr(unmodifiedStats(f), which);
      const uint32_t v = base * mod[0] / mod[1];
      *stat = (uint16_t)(v > 1 ? v : 1);
      log_boost(id, REASON[which], -n);
      if (which == B_spc) log_boost(id, BST_SpecialDefense, -1);
    } else {
      log_boost(id, BST_Accuracy, -1);
    }
    statusModify(stored(f).status, fa.stats);
  }

  void clearVolatiles(int who) {  // :2711-2759
    Active& a = side(who).active;
    const uint8_t id = ident(who);
    if (vfget(a, VF_disable_move) != 0) {
      vfset(a, VF_disable_move, 0);
      vfset(a, VF_disable_duration, 0);
      ch_observe(F_disable, who, OBS_NONE);
      log_end(id, END_DisableSilent);
    }
    if (vget(a, V_Confusion)) {
      vset(a, V_Confusion, false);
      ch_observe(F_confusion, who, OBS_NONE);
      log_end(id, END_ConfusionSilent);
    }
    static const struct { int bit; uint8_t end; } FLAGS[5] = {{V_Mist, END_MistSilent}, {V_FocusEnergy, END_FocusEnergySilent},
        {V_LeechSeed, END_LeechSeedSilent}, {V_LightScreen, END_LightScreenSilent}, {V_Reflect, END_ReflectSilent}};
    for (const auto& fl : FLAGS) {
      if (vget(a, fl.bit)) {
        vset(a, fl.bit, false);
        log_end(id, fl.end);
      }
    }
    if (vget(a, V_Toxic)) {
      vset(a, V_Toxic, false);
      vfset(a, VF_toxic, 0);
      log_end(id, END_ToxicSilent);
    }
  }

  // ------------------------------------------------------------------ choices (:3139-3251)
  uint8_t choices(int p, int request, uint8_t* out) {
    uint8_t n = 0;
    if (request == CH_PASS) {
      out[n++] = 0;
      return n;
    }
    const Side& s = b.sides[p];
    if (request == CH_SWITCH) {
      for (int slot = 2; slot <= 6; slot++) {
        const uint8_t id = s.order[slot - 1];
        if (id == 0 || s.pokemon[id - 1].hp == 0) continue;
        out[n++] = (uint8_t)((slot << 2) | CH_SWITCH);
      }
      if (n == 0) out[n++] = 0;
      return n;
    }
    const Active& a = s.active;
    if (vget(a, V_Recharging) || vget(a, V_Rage) || vget(a, V_Thrashing) || vget(a, V_Charging)) {
      out[n++] = (1 << 2) | CH_MOVE;
      return n;
    }
    for (int slot = 2; slot <= 6; slot++) {
      const uint8_t id = s.order[slot - 1];
      if (id == 0 || s.pokemon[id - 1].hp == 0) continue;
      out[n++] = (uint8_t)((slot << 2) | CH_SWITCH);
    }
    const bool limited = vget(a, V_Bide) || vget(a, V_Binding);
    if (limited) {
      for (int slot = 1; slot <= 4; slot++) {
        const MoveSlot m = a.moves[slot - 1];
        if (m.id == MOVE_None) break;
        if (m.id == (vget(a, V_Bide) ? (uint8_t)MOVE_Bide : s.last_selected_move)) {
          const bool struggle = m.id == MOVE_Bide && (m.pp == 0 || vfget(a, VF_disable_move) == (uint32_t)slot);
          out[n++] = (uint8_t)(((struggle ? 0 : slot) << 2) | CH_MOVE);
          return n;
        }
      }
    }
    const uint8_t before = n;
    for (int slot = 1; slot <= 4; slot++) {
      const MoveSlot m = a.moves[slot - 1];
      if (m.id == MOVE_None) break;
      if (m.pp == 0) continue;
      if (vfget(a, VF_disable_move) == (uint32_t)slot) continue;
      out[n++] = (uint8_t)((slot << 2) | CH_MOVE);
    }
    if (n == before) out[n++] = (0 << 2) | CH_MOVE;
    return n;
  }
};

// Oak's build of libpkmn (reference dev/libpkmn:9): -Dshowdown miss=false advance=false ebc=false key=true
using OakEngine = Engine<false, false, false, true, false>;
using OakLogEngine = Engine<false, false, false, true, false, true>;
// libpkmn's default -Dshowdown options with a scripted roll list: what gen1/test.zig is written for
using KatEngine = Engine<true, true, true, false, true>;
using KatLogEngine = Engine<true, true, true, false, true, true>;
// libpkmn default -Dshowdown options with the live PRNG
using DefaultEngine = Engine<true, true, true, false, false>;

}  // namespace oracle
