// B200 Generation I battle engine: one battle per thread, state in a shared-memory column
// (gen1_state.cuh), PRNG / last_damage / chance side-cars in registers.
//
// Behavioural contract = the reference's libpkmn built the way Oak builds it
// (reference dev/libpkmn:9: -Dshowdown, miss=false, advance=false, ebc=false, key=true, chance, calc):
//   update      engine/src/lib/gen1/mechanics.zig:53-79   choices  :3139-3251
//   Rolls.*     :2774-3115 (64-bit LCG of common/rng.zig:34-59,149-164; `advance` is a no-op)
//   Chance.*    gen1/chance.zig:367-756 (durations + actions, no probability)
//   Calc        gen1/calc.zig:94-116 (overrides), enabled by the CALC template flag
// Differences from a scalar restatement, chosen for the SIMT machine:
//   * both players run through ONE copy of every routine with the player as data, so lanes whose
//     first mover differs stay converged; doTurn is a 2-trip loop with a single executeMove site;
//   * mirrorMove/metronome recursion (mechanics.zig:1351-1396) is an explicit loop in doMove;
//   * boolean state is tested with masks on the packed volatile words instead of field reads;
//   * legal choices are packed 5 bits each into one 64-bit register (no per-thread arrays).
#pragma once

#include "gen1_mem_ops.cuh"

namespace oakb200 {

enum { OBSF_SLEEP = 0, OBSF_CONFUSION, OBSF_DISABLE, OBSF_ATTACKING, OBSF_BINDING };
enum { BM_DONE = 0, BM_SKIP_CAN, BM_SKIP_PP, BM_OK, BM_ERR };
static const int NO_RESULT = -1;

// LOG: write the binary protocol log of every update into `sink` (set sink.buf / cap / pos before update()); a separate
// instantiation, so the kernels built without it carry none of it.
// SPLIT: the on-chip column holds only the fields an update can write (gen1_state.cuh); call m.set_const(record) after begin().
template <int STRIDE, bool CALC, bool CHANCE, bool LOG = false, bool SPLIT = false>
struct Engine {
  typedef typename MemFor<STRIDE, LOG, SPLIT>::type M;
  M m;
  LogSink sink;
  uint64_t rng;
  uint32_t last_damage;
  uint64_t act[2];  // chance actions (per-update observation)
  uint32_t dur[2];  // chance durations (carried)
  uint64_t ovr[2];  // calc overrides (CALC only)

  // ---------------------------------------------------------------- load / store of register copies
  template <bool S>
  OAK_HD void attach_sink(Mem<STRIDE, S>&) {}
  OAK_HD void attach_sink(MemL<STRIDE>& w) { w.sink = &sink; }
  OAK_HD void begin(uint8_t* column_base, uint32_t d0, uint32_t d1) {
    m.set(column_base);
    attach_sink(m);
    rng = m.ld64(B_RNG);
    last_damage = m.ld16(B_LAST_DAMAGE);
    act[0] = act[1] = 0;
    dur[0] = d0;
    dur[1] = d1;
    ovr[0] = ovr[1] = 0;
  }
  OAK_HD void end() {
    m.st64(B_RNG, rng);
    m.st16(B_LAST_DAMAGE, last_damage);
  }

  // ---------------------------------------------------------------- field helpers (forward to Mem)
  static OAK_HD H sb(int p) { return M::sb(p); }
  static OAK_HD PH pokemon(int p, int idx) { return M::pokemon(p, idx); }
  OAK_HD uint32_t ident(int p) const { return m.ident(p); }
  OAK_HD void log_lastmiss_miss(int p) const {  // checkHit's miss report (:1413-1415)
    if (!LOG) return;
    m.lg1(LG_LastMiss);
    m.lg2(LG_Miss, ident(p));
  }
  OAK_HD PH stored(int p) const { return m.stored(p); }
  OAK_HD uint32_t vlo(int p) const { return m.vlo(p); }
  OAK_HD void set_vlo(int p, uint32_t v) const { m.set_vlo(p, v); }
  OAK_HD uint32_t vhi(int p) const { return m.vhi(p); }
  OAK_HD void set_vhi(int p, uint32_t v) const { m.set_vhi(p, v); }
  OAK_HD bool vflag(int p, uint32_t mask) const { return m.vflag(p, mask); }
  OAK_HD void vclear(int p, uint32_t mask) const { m.vclear(p, mask); }
  OAK_HD void vraise(int p, uint32_t mask) const { m.vraise(p, mask); }
  OAK_HD uint32_t vlo_field(int p, int shift, int bits) const { return m.vlo_field(p, shift, bits); }
  OAK_HD void set_vlo_field(int p, int shift, int bits, uint32_t x) const { m.set_vlo_field(p, shift, bits, x); }
  OAK_HD uint32_t vhi_field(int p, int shift, int bits) const { return m.vhi_field(p, shift, bits); }
  OAK_HD void set_vhi_field(int p, int shift, int bits, uint32_t x) const { m.set_vhi_field(p, shift, bits, x); }
  OAK_HD uint32_t vstate(int p) const { return m.vstate(p); }
  OAK_HD void set_vstate(int p, uint32_t s) const { m.set_vstate(p, s); }
  OAK_HD uint32_t status_of(PH pk) const { return m.status_of(pk); }
  OAK_HD uint32_t hp_of(PH pk) const { return m.hp_of(pk); }
  OAK_HD uint32_t last_selected(int p) const { return m.last_selected(p); }
  OAK_HD uint32_t active_types(int p) const { return m.active_types(p); }
  OAK_HD uint32_t active_stat(int p, int st) const { return m.active_stat(p, st); }
  OAK_HD void set_active_stat(int p, int st, uint32_t v) const { m.set_active_stat(p, st, v); }
  // applyDamage through the shared routine; returns true when a Substitute broke
  OAK_HD bool applyDamage(int t, int subp, uint32_t reason = DMG_None) {
    const uint32_t r = LOG ? mem_apply_damage_reason(m, t, subp, last_damage, reason) : mem_apply_damage(m, t, subp, last_damage);
    last_damage = r & 0x7FFFFFFFu;
    return (r >> 31) != 0;
  }
  OAK_HD void buildRage(int who) { mem_build_rage(m, who); }
  // applyDamage to the foe with the current last_damage + Rage build-up unless a Substitute took it
  OAK_HD void hitFoeWithLastDamage(int p) {
    const int f = p ^ 1;
    const bool sub = vflag(f, VL_SUBSTITUTE);
    applyDamage(f, f);
    if (hp_of(stored(f)) > 0 && !sub) buildRage(f);
  }
  OAK_HD uint64_t choices(int p, int request, int* n) const {
    const uint64_t l = mem_choices(m, p, request);
    *n = choice_count(l);
    return l;
  }
  static OAK_HD uint32_t choice_at(uint64_t list, int i) { return oakb200::choice_at(list, i); }

  // ---------------------------------------------------------------- PRNG (rng.zig:34-59,149-164)
  OAK_HD uint32_t next() {
    rng = 0x5D588B656C078965ull * rng + 0x269EC3ull;
    return (uint32_t)(rng >> 32);
  }
  OAK_HD uint32_t range(uint32_t from, uint32_t to) {
#ifdef __CUDA_ARCH__
    return __umulhi(next(), to - from) + from;
#else
    return (uint32_t)(((uint64_t)next() * (to - from)) >> 32) + from;
#endif
  }
  OAK_HD bool chance(uint32_t num, uint32_t den) { return range(0, den) < num; }

  // ---------------------------------------------------------------- calc overrides (calc.zig:94-116)
  OAK_HD uint32_t overridden(int p, int shift, int bits) const {
    if (!CALC) return 0;
    return (uint32_t)((ovr[p] >> shift) & ((1ull << bits) - 1));
  }
  OAK_HD void calc_confusion(int p) {
    if (!CALC) return;
    if (overridden(p, AF_CONFUSION) == OBS_STARTED) ovr[p] &= ~(15ull << 27);
  }

  // ---------------------------------------------------------------- chance recorders (chance.zig:435-756)
  OAK_HD void a_set(int p, int shift, int bits, uint32_t x) {
    if (!CHANCE) return;
    const uint64_t msk = ((1ull << bits) - 1) << shift;
    act[p] = (act[p] & ~msk) | (((uint64_t)x << shift) & msk);
  }
  OAK_HD void a_bool(int p, int shift, int bits, bool v) { a_set(p, shift, bits, v ? OPT_TRUE : OPT_FALSE); }
  OAK_HD void ch_switched(int p, int slot) {
    if (!CHANCE) return;
    uint32_t d = dur[p];
    const uint32_t s0 = d & 7u, sn = (d >> (3 * (slot - 1))) & 7u;
    d = (d & ~7u) | sn;
    d = (d & ~(7u << (3 * (slot - 1)))) | (s0 << (3 * (slot - 1)));
    dur[p] = d & 0x0003FFFFu;  // confusion, disable, attacking, binding <- 0
    dur[p ^ 1] = bf_set(dur[p ^ 1], DF_BINDING, 0);
  }
  OAK_HD void ch_duration(int p) { a_set(p, AF_DURATION, 2); }  // key=true records the constant 2 (chance.zig:550-554)
  OAK_HD void dur_put(int p, int field, uint32_t v) {
    uint32_t d = dur[p];
    switch (field) {
      case OBSF_SLEEP: d = (d & ~7u) | v; break;
      case OBSF_CONFUSION: d = bf_set(d, DF_CONFUSION, v); break;
      case OBSF_DISABLE: d = bf_set(d, DF_DISABLE, v); break;
      case OBSF_ATTACKING: d = bf_set(d, DF_ATTACKING, v); break;
      default: d = bf_set(d, DF_BINDING, v); break;
    }
    dur[p] = d;
  }
  OAK_HD void a_put_obs(int p, int field, uint32_t obs) {
    switch (field) {
      case OBSF_SLEEP: a_set(p, AF_SLEEP, obs); break;
      case OBSF_CONFUSION: a_set(p, AF_CONFUSION, obs); break;
      case OBSF_DISABLE: a_set(p, AF_DISABLE, obs); break;
      case OBSF_ATTACKING: a_set(p, AF_ATTACKING, obs); break;
      default: a_set(p, AF_BINDING, obs); break;
    }
  }
  // chance.zig:556-600: obs is None / started / overwritten
  OAK_HD void ch_observe(int field, int p, uint32_t obs) {
    if (!CHANCE) return;
    a_put_obs(p, field, obs);
    dur_put(p, field, obs == OBS_NONE ? 0u : 1u);
  }
  // chance.zig:602-733: per-turn tick of a tracked duration; obs is None(sleep only) / continuing / ended
  OAK_HD void ch_tick(int field, int p, uint32_t obs) {
    if (!CHANCE) return;
    a_put_obs(p, field, obs);
    if (obs == OBS_CONTINUING) {
      uint32_t d = dur[p], n;
      switch (field) {
        case OBSF_SLEEP: n = d & 7u; break;
        case OBSF_CONFUSION: n = bf_get(d, DF_CONFUSION); break;
        case OBSF_DISABLE: n = bf_get(d, DF_DISABLE); break;
        case OBSF_ATTACKING: n = bf_get(d, DF_ATTACKING); break;
        default: n = bf_get(d, DF_BINDING); break;
      }
      dur_put(p, field, n + 1);
    } else {
      dur_put(p, field, 0);
    }
  }

  // ---------------------------------------------------------------- Rolls (mechanics.zig:2774-3115)
  OAK_HD bool roll_speedTie() {
    const uint32_t ov = overridden(0, AF_SPEED_TIE);
    const bool p1 = ov ? (ov == 1) : (range(0, 2) == 0);
    a_set(0, AF_SPEED_TIE, p1 ? 1 : 2);
    a_set(1, AF_SPEED_TIE, p1 ? 1 : 2);
    return p1;
  }
  OAK_HD bool roll_bool(int p, int shift, int bits, uint32_t num) {  // chance(num, 256) with override + record
    const uint32_t ov = overridden(p, shift, bits);
    const bool v = ov ? (ov == OPT_TRUE) : chance(num, 256);
    a_bool(p, shift, bits, v);
    return v;
  }
  OAK_HD bool roll_crit(int p, uint32_t rate) { return roll_bool(p, AF_CRIT, rate); }
  OAK_HD bool roll_hit(int p, uint32_t accuracy) { return roll_bool(p, AF_HIT, accuracy + 1); }  // miss=false (:2827-2841)
  OAK_HD bool roll_paralyzed(int p) { return roll_bool(p, AF_PARALYZED, 63); }
  OAK_HD bool roll_secondary(int p, uint32_t rate) { return roll_bool(p, AF_SECONDARY, rate); }
  OAK_HD bool roll_confused(int p) {
    const uint32_t ov = overridden(p, AF_CONFUSED);
    const bool v = ov ? (ov == OPT_TRUE) : !chance(128, 256);
    a_bool(p, AF_CONFUSED, v);
    return v;
  }
  OAK_HD uint32_t roll_damage(int p) {
    const uint32_t ov = overridden(p, AF_DAMAGE);
    const uint32_t roll = ov ? ov : range(217, 256);
    a_set(p, AF_DAMAGE, roll);
    return roll;
  }
  OAK_HD uint32_t roll_metronome(int p) {
    const uint32_t ov = overridden(p, AF_METRONOME);
    uint32_t mv = ov;
    if (!ov) {
      const uint32_t r = range(0, 163);  // Move.METRONOME: 1..164 without Metronome (moves.zig:1840-1851)
      mv = r + 1 < (uint32_t)MOVE_Metronome ? r + 1 : r + 2;
    }
    a_set(p, AF_METRONOME, mv);
    return mv;
  }
  OAK_HD uint32_t roll_psywave(int p, uint32_t max) {
    const uint32_t ov = overridden(p, AF_PSYWAVE);
    const uint32_t power = ov ? ov - 1 : range(0, max);
    a_set(p, AF_PSYWAVE, power + 1);
    return power;
  }
  OAK_HD uint32_t roll_duration(int p, uint32_t lo, uint32_t hi_excl) {
    const uint32_t ov = overridden(p, AF_DURATION);
    const uint32_t d = (ov >= lo && ov < hi_excl) ? ov : range(lo, hi_excl);
    ch_duration(p);
    return d;
  }
  OAK_HD uint32_t roll_distribution(int p, bool binding) {  // {2,2,2,3,3,3,4,5}[range(0,8)]
    const uint32_t ov = binding ? overridden(p, AF_DURATION) : overridden(p, AF_MULTI_HIT);
    uint32_t n = ov;
    if (!(ov >= 2 && ov <= 5)) n = (0x54333222u >> (4 * range(0, 8))) & 15u;
    if (binding) {
      ch_duration(p);
    } else {
      a_set(p, AF_MULTI_HIT, n);
    }
    return n;
  }
  // moves_off: handle of a (word-aligned) 4 x {id, pp} array
  OAK_HD uint32_t roll_moveSlot(int p, H moves_off, uint32_t check_pp) {
    const uint32_t ov = overridden(p, AF_MOVE_SLOT);
    uint32_t slot = 0;
    if (ov && m.ld8(moves_off, 2 * (ov - 1)) != MOVE_None && (check_pp == 0 || m.ld8(moves_off, 2 * (ov - 1) + 1) > 0)) {
      slot = ov;
    } else if (check_pp == 0) {
      int last = 0;
      for (int i = 0; i < 4; i++)
        if (m.ld8(moves_off, 2 * i) != MOVE_None) last = i + 1;
      if (last) slot = range(0, (uint32_t)last) + 1;
    } else {
      uint32_t r = range(0, check_pp) + 1;
      for (int i = 0; i < 4; i++) {
        if (r > 0 && m.ld8(moves_off, 2 * i + 1) > 0) {
          r--;
          if (r == 0) slot = (uint32_t)i + 1;
        }
      }
    }
    a_set(p, AF_MOVE_SLOT, slot);
    return slot;
  }
  OAK_HD uint32_t decrement(int p, int shift, int bits, uint32_t n) {  // mechanics.zig:2761-2772
    const uint32_t obs = overridden(p, shift, bits);
    if (obs == OBS_NONE) return n - 1;
    if (obs == OBS_STARTED || obs == OBS_ENDED) return 0;
    return n > 1 ? n - 1 : n;
  }

#include "gen1_engine_turn.inc"
#include "gen1_engine_move.inc"
#include "gen1_engine_effects.inc"
};

}  // namespace oakb200
