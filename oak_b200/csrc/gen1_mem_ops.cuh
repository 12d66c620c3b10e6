// Pure state transformations of the battle engine: routines that read and write only the battle
// column (no PRNG, no chance/calc registers). They are compiled as real (non-inlined) device
// functions taking the column pointer by value, so each exists ONCE in the kernel image however many
// call sites use it: the first B200 profile (profiles/r1a_*) showed the fully-inlined engine at
// 142 KB of SASS stalling ~17 cycles per issue on instruction fetch.
// Reference line numbers are engine/src/lib/gen1/mechanics.zig.
#pragma once

#include "gen1_state.cuh"

namespace oakb200 {

// :2698-2704; stats = handle of a 5 x u16 stats block
template <class M>
OAK_HD void mem_status_modify(M m, uint32_t status, H stats) {
  if (status & SB_PAR) {
    const uint32_t v = m.ld16(stats, ST_SPE) / 4;
    m.st16(stats, ST_SPE, v > 1 ? v : 1);
  } else if (status & SB_BRN) {
    const uint32_t v = m.ld16(stats, ST_ATK) / 2;
    m.st16(stats, ST_ATK, v > 1 ? v : 1);
  }
}

// applyDamage :1314-1349. Returns new last_damage | (substitute broke ? 1u << 31 : 0). The common case (no Substitute up)
// is inline at the call sites — a call costs about as much as the four instructions it would save — the Substitute branch
// stays out of line.
template <class M>
OAK_HD_NOINLINE uint32_t mem_apply_damage_substitute(M m, int subp, uint32_t last_damage) {
  const uint32_t lo = m.vlo(subp);
  if (!(lo & VL_SUBSTITUTE)) return last_damage;
  const uint32_t hi = m.vhi(subp);
  const uint32_t shp = bf_get(hi, VHI_SUBSTITUTE);
  if (last_damage >= shp) {
    m.set_vhi(subp, bf_set(hi, VHI_SUBSTITUTE, 0));
    m.set_vlo(subp, lo & ~VL_SUBSTITUTE);
    m.lg3(LG_End, m.ident(subp), END_Substitute);
    return last_damage | 0x80000000u;
  }
  m.set_vhi(subp, bf_set(hi, VHI_SUBSTITUTE, shp - last_damage));
  m.lg3(LG_Activate, m.ident(subp), ACT_Substitute);
  return last_damage;
}
template <class M>
OAK_HD uint32_t mem_apply_damage_impl(M m, int t, int subp, uint32_t last_damage, uint32_t reason) {
  if (m.vflag(t, VL_SUBSTITUTE)) return mem_apply_damage_substitute(m, subp, last_damage);
  const PH pk = m.stored(t);
  const uint32_t hp = m.ld16(pk, PK_HP);
  if (last_damage > hp) last_damage = hp;
  m.st16(pk, PK_HP, hp - last_damage);
  m.lg_hp(LG_Damage, m.ident(t), pk, reason);
  return last_damage;
}
template <class M>
OAK_HD uint32_t mem_apply_damage(M m, int t, int subp, uint32_t last_damage) {
  return mem_apply_damage_impl(m, t, subp, last_damage, DMG_None);
}
// with the |-damage| reason of the log (only the confusion self-hit has one, mechanics.zig:624); LOG builds only
template <class M>
OAK_HD uint32_t mem_apply_damage_reason(M m, int t, int subp, uint32_t last_damage, uint32_t reason) {
  return mem_apply_damage_impl(m, t, subp, last_damage, reason);
}

// unmodifiedStats :2692-2696
template <class M>
OAK_HD PH mem_unmodified_stats(M m, int p) {  // handle of the Pokemon whose (PK_STATS) block applies
  if (!m.vflag(p, VL_TRANSFORM)) return m.stored(p);
  const uint32_t id = m.vhi_field(p, VHI_TRANSFORM);
  return M::pokemon((int)(id >> 3), (int)(id & 7u) - 1);
}

// Effects.boost :2484-2582 (also Rage build-up via buildRage :1734-1739)
template <class M>
OAK_HD_NOINLINE void mem_boost(M m, int p, uint32_t eff) {
  int which, n;
  switch (eff) {
    case EFF_AttackUp1: case EFF_Rage: which = BO_ATK; n = 1; break;
    case EFF_AttackUp2: which = BO_ATK; n = 2; break;
    case EFF_DefenseUp1: which = BO_DEF; n = 1; break;
    case EFF_DefenseUp2: which = BO_DEF; n = 2; break;
    case EFF_SpeedUp2: which = BO_SPE; n = 2; break;
    case EFF_SpecialUp1: which = BO_SPC; n = 1; break;
    case EFF_SpecialUp2: which = BO_SPC; n = 2; break;
    default: which = BO_EVA; n = 1; break;  // EvasionUp1
  }
  const H s = M::sb(p);
  const uint32_t b = m.ld32(s, A_BOOSTS);
  const int cur = boost_get(b, which);
  const uint32_t id = M::LOGGING ? m.ident(p) : 0u;
  if (cur == 6) {
    m.lg3(LG_Fail, id, FAIL_None);
    return;
  }
  const int nb = cur + n > 6 ? 6 : cur + n;
  // |-boost| reasons (protocol.zig Boost): Attack 1, Defense 2, Speed 3, SpecialAttack 4 (+ SpecialDefense 5), Evasion 7
  const uint32_t stat_reason = which == BO_EVA ? (uint32_t)BST_Evasion : (uint32_t)which + 1u;
  const uint32_t reason = eff == EFF_Rage ? (uint32_t)BST_Rage : stat_reason;
  if (which != BO_EVA) {
    const int so = A_STATS + ST_ATK + 2 * which;
    if (m.ld16(s, so) == 999) {
      m.st32(s, A_BOOSTS, boost_set(b, which, nb - 1));
      if (M::LOGGING) {  // :2498-2506 and the same shape for the other stats
        m.lg_boost(id, reason, n);
        m.lg_boost(id, stat_reason, -1);
        if (which == BO_SPC) {
          m.lg_boost(id, BST_SpecialDefense, n);
          m.lg_boost(id, BST_SpecialDefense, -1);
        }
        if (eff != EFF_Rage) m.lg3(LG_Fail, id, FAIL_None);
      }
      return;
    }
    const uint32_t v = boost_scale(m.stat_at(mem_unmodified_stats(m, p), ST_ATK + 2 * which), nb);
    m.st16(s, so, v < 999 ? v : 999);
  }
  m.st32(s, A_BOOSTS, boost_set(b, which, nb));
  m.lg_boost(id, reason, n);
  if (which == BO_SPC) m.lg_boost(id, BST_SpecialDefense, n);
  if (eff == EFF_Rage) return;
  // "stat modification errors" glitch: the foe's PAR/BRN drop is applied again (:2581)
  mem_status_modify(m, m.status_of(m.stored(p ^ 1)), M::hadd(M::sb(p ^ 1), A_STATS));
}

// Effects.unboost :2600-2689 — the part after the Substitute / secondary-chance / hit gates
template <class M>
OAK_HD_NOINLINE void mem_unboost(M m, int f, uint32_t eff) {
  int which, n = 1;
  switch (eff) {
    case EFF_AttackDown1: case EFF_AttackDownChance: which = BO_ATK; break;
    case EFF_DefenseDown1: case EFF_DefenseDownChance: which = BO_DEF; break;
    case EFF_DefenseDown2: which = BO_DEF; n = 2; break;
    case EFF_SpeedDown1: case EFF_SpeedDownChance: which = BO_SPE; break;
    case EFF_SpecialDownChance: which = BO_SPC; break;
    default: which = BO_ACC; break;  // AccuracyDown1
  }
  const H s = M::sb(f);
  const uint32_t b = m.ld32(s, A_BOOSTS);
  const int cur = boost_get(b, which);
  const uint32_t id = M::LOGGING ? m.ident(f) : 0u;
  if (cur == -6) {
    m.lg3(LG_Fail, id, FAIL_None);  // showdown: also for secondary effects (:2603)
    return;
  }
  const int nb = cur - n < -6 ? -6 : cur - n;
  const uint32_t stat_reason = which == BO_ACC ? (uint32_t)BST_Accuracy : (uint32_t)which + 1u;
  if (which != BO_ACC) {
    const int so = A_STATS + ST_ATK + 2 * which;
    if (m.ld16(s, so) == 1) {
      m.st32(s, A_BOOSTS, boost_set(b, which, nb + 1));
      if (M::LOGGING) {
        m.lg_boost(id, stat_reason, -n);
        m.lg_boost(id, stat_reason, 1);
        if (which == BO_SPC) {
          m.lg_boost(id, BST_SpecialDefense, -1);
          m.lg_boost(id, BST_SpecialDefense, 1);
        }
        m.lg3(LG_Fail, id, FAIL_None);
      }
      return;
    }
    const uint32_t v = boost_scale(m.stat_at(mem_unmodified_stats(m, f), ST_ATK + 2 * which), nb);
    m.st16(s, so, v > 1 ? v : 1);
  }
  m.st32(s, A_BOOSTS, boost_set(b, which, nb));
  m.lg_boost(id, stat_reason, -n);
  if (which == BO_SPC) m.lg_boost(id, BST_SpecialDefense, -1);
  mem_status_modify(m, m.status_of(m.stored(f)), M::hadd(s, A_STATS));
}

// buildRage :1734-1739
template <class M>
OAK_HD void mem_build_rage(M m, int who) {
  if (m.vflag(who, VL_RAGE) && boost_get(m.ld32(M::sb(who), A_BOOSTS), BO_ATK) < 6) mem_boost(m, who, EFF_Rage);
}

// handleResidual :1585-1637
template <class M>
OAK_HD_NOINLINE void mem_handle_residual_slow(M m, int p) {
  const PH pk = m.stored(p);
  const uint32_t st = m.ld8(pk, PK_STATUS);
  const uint32_t lo = m.vlo(p);
  if (!(st & (SB_BRN | SB_PSN)) && !(lo & VL_LEECH_SEED)) return;
  const uint32_t maxhp = m.ld16(pk, PK_STATS + ST_HP);
  const uint32_t tick = maxhp / 16 > 1 ? maxhp / 16 : 1;
  if (st & (SB_BRN | SB_PSN)) {
    uint32_t damage = tick;
    if (lo & VL_TOXIC) {
      const uint32_t t = (m.vhi_field(p, VHI_TOXIC) + 1) & 31u;
      m.set_vhi_field(p, VHI_TOXIC, t);
      damage *= t;
    }
    const uint32_t hp = m.ld16(pk, PK_HP);
    if (damage > hp) damage = hp;
    m.st16(pk, PK_HP, hp - damage);
    if (M::LOGGING && damage > 0) m.lg_hp(LG_Damage, m.ident(p), pk, (st & SB_BRN) ? (uint32_t)DMG_Burn : (uint32_t)DMG_Poison);
  }
  if (lo & VL_LEECH_SEED) {
    const PH fk = m.stored(p ^ 1);
    const uint32_t fhp = m.ld16(fk, PK_HP);
    if (fhp == 0) return;
    uint32_t damage = tick;
    if (lo & VL_TOXIC) {
      const uint32_t t = (m.vhi_field(p, VHI_TOXIC) + 1) & 31u;
      m.set_vhi_field(p, VHI_TOXIC, t);
      damage *= t;
    }
    const uint32_t hp = m.ld16(pk, PK_HP);
    m.st16(pk, PK_HP, hp - (damage < hp ? damage : hp));
    if (M::LOGGING && damage > 0 && hp > 0) m.lg_hp(LG_Damage, m.ident(p), pk, DMG_LeechSeed);
    const uint32_t fmax = m.ld16(fk, PK_STATS + ST_HP);
    const uint32_t nhp = fhp + damage < fmax ? fhp + damage : fmax;
    m.st16(fk, PK_HP, nhp);
    if (M::LOGGING && nhp > fhp) m.lg_hp(LG_Heal, m.ident(p ^ 1), fk, HEAL_Silent);
  }
}

// the common case (no burn / poison / Leech Seed) decided at the call site: no call, no argument moves
template <class M>
OAK_HD void mem_handle_residual(M m, int p) {
  const uint32_t st = m.ld8(m.stored(p), PK_STATUS);
  if (!(st & (SB_BRN | SB_PSN)) && !m.vflag(p, VL_LEECH_SEED)) return;
  mem_handle_residual_slow(m, p);
}

template <class M>
OAK_HD int mem_find_first_alive(M m, int p) {  // :99-102
  for (int i = 0; i < 6; i++)
    if (m.ld16(M::pokemon(p, i), PK_HP) > 0) return (int)m.ld8(M::sb(p), S_ORDER + i);
  return 0;
}

// checkFaint :1507-1543 + faint :1545-1583. Returns the result byte or -1.
template <class M>
OAK_HD_NOINLINE int mem_check_faint_slow(M m, int p) {
  const bool foe_fainted = m.hp_of(m.stored(p ^ 1)) == 0;
  const bool player_out = mem_find_first_alive(m, p) == 0;
  const bool foe_out = foe_fainted && mem_find_first_alive(m, p ^ 1) == 0;
  const bool more = player_out || foe_out;
  for (int k = 0; k < 2; k++) {  // faint(p), then faint(foe) when it fainted too
    const int w = p ^ k;
    if (k == 1 && !foe_fainted) break;
    const int f = w ^ 1;
    if (m.vflag(f, VL_BIDE)) m.set_vstate(f, 0);
    m.set_vlo(w, 0);
    m.set_vhi(w, 0);
    m.st8(M::sb(w), S_LAST_USED, MOVE_None);
    m.st8(M::sb(f), S_LAST_USED, MOVE_None);
    const PH pk = m.stored(w);
    const uint32_t st = m.ld8(pk, PK_STATUS);
    m.st8(pk, PK_STATUS, 0);
    uint32_t spe = m.ld16(pk, PK_STATS + ST_SPE);
    if (st & SB_PAR) spe = spe / 4 > 1 ? spe / 4 : 1;
    m.set_active_stat(w, ST_SPE, spe);
    if (M::LOGGING) {  // faint(.., done): :1522-1523, protocol.zig:396-401
      m.lg2(LG_Faint, m.ident(w));
      if (k == 0 ? !(more || foe_fainted) : !more) m.lg1(LG_None);
    }
  }
  if (player_out && foe_out) {
    m.lg2(LG_Tie, LG_None);
    return RES_TIE;
  }
  if (player_out) {
    m.lg3(LG_Win, p ^ 1, LG_None);
    return p == 0 ? RES_LOSE : RES_WIN;
  }
  if (foe_out) {
    m.lg3(LG_Win, p, LG_None);
    return p == 0 ? RES_WIN : RES_LOSE;
  }
  const int fc = foe_fainted ? CH_SWITCH : CH_PASS;
  return p == 0 ? ((CH_SWITCH << 4) | (fc << 6)) : ((fc << 4) | (CH_SWITCH << 6));
}

// the common case (the active Pokémon has not fainted) decided at the call site
template <class M>
OAK_HD int mem_check_faint(M m, int p) {
  if (m.hp_of(m.stored(p)) > 0) return -1;
  return mem_check_faint_slow(m, p);
}

// switchIn :225-263 without the chance bookkeeping (done by the caller)
template <class M>
OAK_HD_NOINLINE void mem_switch_in(M m, int p, int slot) {
  const H s = M::sb(p);
  const uint32_t out = m.ld8(s, S_ORDER);
  const uint32_t in = m.ld8(s, S_ORDER + slot - 1);
  m.st8(s, S_ORDER, in);
  m.st8(s, S_ORDER + slot - 1, out);
  const PH pk = M::pokemon(p, (int)in - 1);
  m.st8(B_LAST_MOVES + 2 * p, 1);
  m.st8(s, S_LAST_USED, MOVE_None);
  m.st8(M::sb(p ^ 1), S_LAST_USED, MOVE_None);
#ifdef __CUDA_ARCH__
  if (M::IS_SPLIT) {
    // Random playouts switch on about every other choice, so this copy is hot (3.2 % of the playout kernel's instructions
    // as 11 scattered loads of constant fields + 14 stores): the 24-byte record (8-byte aligned) as three 64-bit loads, the
    // active block composed by byte permutes (data.zig:185-203 -> :205-232) and written as eight words.
    const uint2 a = m.cld64(pk.c), b = m.cld64(pk.c + 8), c = m.cld64(pk.c + 16);
    // a.x a.y = hp atk | def spe; b.x = spc id0 pp0; b.y = id1 pp1 id2 pp2; c.x = id3 pp3 hp; c.y = status species types level
    const uint32_t ppw = m.ld32a(m.atv(pk.v));                        // the four PP bytes (the record's own are stale)
    const uint32_t idw = __byte_perm(__byte_perm(b.x, b.y, 0x0642), c.x, 0x4210);  // id0 id1 id2 id3
    m.st32(s, A_STATS, a.x);
    m.st32(s, A_STATS + 4, a.y);
    m.st32(s, A_STATS + 8, __byte_perm(b.x, c.y, 0x6510));            // spc | species | types
    m.st32(s, A_BOOSTS, 0);
    m.st32(s, A_VOL, 0);
    m.st32(s, A_VOL + 4, 0);
    m.st32(s, A_MOVES, __byte_perm(idw, ppw, 0x5140));                // id0 pp0 id1 pp1
    m.st32(s, A_MOVES + 4, __byte_perm(idw, ppw, 0x7362));            // id2 pp2 id3 pp3
  } else
#endif
  {
    for (int i = 0; i < 10; i += 2) m.st16(s, A_STATS + i, m.ld16(pk, PK_STATS + i));
    m.st8(s, A_SPECIES, m.ld8(pk, PK_SPECIES));
    m.st8(s, A_TYPES, m.ld8(pk, PK_TYPES));
    m.st32(s, A_BOOSTS, 0);
    m.st32(s, A_VOL, 0);
    m.st32(s, A_VOL + 4, 0);
    for (int i = 0; i < 8; i += 2) m.st16(s, A_MOVES + i, m.ld16(pk, PK_MOVES + i));
  }
  const uint32_t st = m.ld8(pk, PK_STATUS);
  mem_status_modify(m, st, M::hadd(s, A_STATS));
  m.vclear(p ^ 1, VL_BINDING);
  m.lg_switched(((uint32_t)p << 3) | in, pk);
  if (st == SB_TOX) {
    m.st8(pk, PK_STATUS, SB_PSN);
    m.lg4(LG_Status, ((uint32_t)p << 3) | in, SB_PSN, STS_Silent);
  }
}

// choices :3139-3251. Legal choices packed 5 bits each (type | data << 2), first in the low bits;
// the count (1..9) in bits 60..63. Called twice per update by every lane (7.5 % of the playout kernel's instructions), so
// the party order and the four {move, pp} pairs are read as two words each and the two partial lists (switches: at most
// 25 bits, moves: at most 20) are built in 32-bit registers and joined by one 64-bit shift.
template <class M>
OAK_HD_NOINLINE uint64_t mem_choices(M m, int p, int request) {
  if (request == CH_PASS) return 1ull << 60;
  const H s = M::sb(p);
  const uint32_t lo = m.vlo(p);
  if (request == CH_MOVE && (lo & (VL_RECHARGING | VL_RAGE | VL_THRASHING | VL_CHARGING))) {
    return (1ull << 60) | (uint64_t)((1u << 2) | CH_MOVE);
  }
  const uint32_t o0 = m.ld32(s, S_ORDER), o1 = m.ld32(s, S_ORDER + 4);  // order[0..3] | order[4], order[5], last_selected, last_used
  // No branch per entry: lanes differ in which slots are legal, and every divergent branch makes the warp run the same few
  // instructions once per group of lanes (the branchy form was 9.7 % of the playout kernel's warp instructions).
  uint32_t sw = 0, k = 0;
#pragma unroll
  for (int slot = 2; slot <= 6; slot++) {
    const uint32_t id = (slot <= 4 ? o0 >> (8 * (slot - 1)) : o1 >> (8 * (slot - 5))) & 255u;
    const uint32_t hp = m.ld16(M::pokemon(p, id ? (int)id - 1 : 0), PK_HP);  // an empty slot reads member 0 and is masked
    const uint32_t ok = (id != 0 && hp != 0) ? 1u : 0u;
    sw |= (ok ? (((uint32_t)slot << 2) | CH_SWITCH) : 0u) << (5 * k);
    k += ok;
  }
  if (request == CH_SWITCH) return (uint64_t)sw | ((uint64_t)(k == 0 ? 1u : k) << 60);
  const uint32_t dm = m.vhi_field(p, VHI_DISABLE_MOVE);
  const uint32_t mv0 = m.ld32(s, A_MOVES), mv1 = m.ld32(s, A_MOVES + 4);  // id0 pp0 id1 pp1 | id2 pp2 id3 pp3
  if (lo & (VL_BIDE | VL_BINDING)) {
    const uint32_t want = (lo & VL_BIDE) ? (uint32_t)MOVE_Bide : (o1 >> 16) & 255u;  // last_selected
#pragma unroll
    for (int slot = 1; slot <= 4; slot++) {
      const uint32_t e = (slot <= 2 ? mv0 >> (16 * (slot - 1)) : mv1 >> (16 * (slot - 3))) & 0xFFFFu;
      const uint32_t id = e & 255u;
      if (id == MOVE_None) break;
      if (id == want) {
        const bool struggle = id == MOVE_Bide && ((e >> 8) == 0 || dm == (uint32_t)slot);
        const uint32_t entry = ((struggle ? 0u : (uint32_t)slot) << 2) | CH_MOVE;
        return (uint64_t)sw | ((uint64_t)entry << (5 * k)) | ((uint64_t)(k + 1) << 60);
      }
    }
  }
  uint32_t mvl = 0, km = 0, open = 1;  // open: no empty move slot met yet (the reference stops at the first)
#pragma unroll
  for (int slot = 1; slot <= 4; slot++) {
    const uint32_t e = (slot <= 2 ? mv0 >> (16 * (slot - 1)) : mv1 >> (16 * (slot - 3))) & 0xFFFFu;
    open &= (e & 255u) != MOVE_None ? 1u : 0u;
    const uint32_t ok = ((e >> 8) != 0 && dm != (uint32_t)slot) ? open : 0u;
    mvl |= (ok ? (((uint32_t)slot << 2) | CH_MOVE) : 0u) << (5 * km);
    km += ok;
  }
  if (km == 0) {  // move 0 = Struggle
    mvl = CH_MOVE;
    km = 1;
  }
  return (uint64_t)sw | ((uint64_t)mvl << (5 * k)) | ((uint64_t)(k + km) << 60);
}
OAK_HD uint32_t choice_at(uint64_t list, int i) { return (uint32_t)(list >> (5 * i)) & 31u; }
OAK_HD int choice_count(uint64_t list) { return (int)(list >> 60); }

// Effects.haze :2035-2082 + clearVolatiles :2711-2759, state part (chance bookkeeping by the caller)
template <class M>
OAK_HD_NOINLINE void mem_haze(M m, int p) {
  m.lg3(LG_Activate, M::LOGGING ? m.ident(p) : 0u, ACT_Haze);
  m.lg1(LG_ClearAllBoost);
  for (int i = 0; i < 2; i++) {
    const H s = M::sb(i);
    const PH pk = m.stored(i);
    m.st32(s, A_BOOSTS, 0);
    for (int k = 0; k < 10; k += 2) m.st16(s, A_STATS + k, m.ld16(pk, PK_STATS + k));
    const uint32_t st = m.ld8(pk, PK_STATUS);
    if (i != p && st != 0) {
      m.lg4(LG_CureStatus, M::LOGGING ? m.ident(i) : 0u, st, CURE_Silent);
      m.st8(pk, PK_STATUS, 0);
    } else if (st == SB_TOX) {
      m.st8(pk, PK_STATUS, SB_PSN);
      m.lg4(LG_Status, M::LOGGING ? m.ident(i) : 0u, SB_PSN, STS_None);
    }
    uint32_t lo = m.vlo(i), hi = m.vhi(i);
    if (M::LOGGING) {  // clearVolatiles' |-end| messages in its order (:2711-2759)
      const uint32_t id = m.ident(i);
      if (bf_get(hi, VHI_DISABLE_MOVE) != 0) m.lg3(LG_End, id, END_DisableSilent);
      if (lo & VL_CONFUSION) m.lg3(LG_End, id, END_ConfusionSilent);
      if (lo & VL_MIST) m.lg3(LG_End, id, END_MistSilent);
      if (lo & VL_FOCUS_ENERGY) m.lg3(LG_End, id, END_FocusEnergySilent);
      if (lo & VL_LEECH_SEED) m.lg3(LG_End, id, END_LeechSeedSilent);
      if (lo & VL_LIGHT_SCREEN) m.lg3(LG_End, id, END_LightScreenSilent);
      if (lo & VL_REFLECT) m.lg3(LG_End, id, END_ReflectSilent);
      if (lo & VL_TOXIC) m.lg3(LG_End, id, END_ToxicSilent);
    }
    if (bf_get(hi, VHI_DISABLE_MOVE) != 0) hi = bf_set(bf_set(hi, VHI_DISABLE_MOVE, 0), VHI_DISABLE_DURATION, 0);
    if (lo & VL_TOXIC) hi = bf_set(hi, VHI_TOXIC, 0);
    lo &= ~(VL_CONFUSION | VL_MIST | VL_FOCUS_ENERGY | VL_LEECH_SEED | VL_LIGHT_SCREEN | VL_REFLECT | VL_TOXIC);
    m.set_vlo(i, lo);
    m.set_vhi(i, hi);
  }
}

// Effects.heal :2084-2106
template <class M>
OAK_HD void mem_heal(M m, int p) {  // one call site, and a frequent one in random play (Recover / Softboiled / Rest): inline
  const PH pk = m.stored(p);
  const uint32_t hp = m.ld16(pk, PK_HP), maxhp = m.ld16(pk, PK_STATS + ST_HP);
  const uint32_t delta = (maxhp - hp) & 0xFFFFu;
  const uint32_t id = M::LOGGING ? m.ident(p) : 0u;
  if (delta == 0 || ((delta & 255u) == 255u && (hp & 255u) != 0)) {
    m.lg3(LG_Fail, id, FAIL_None);
    return;
  }
  const bool rest = m.last_selected(p) == MOVE_Rest;
  if (rest) {
    m.st8(pk, PK_STATUS, SB_EXT | 2u);
    m.lg4(LG_Status, id, SB_EXT | 2u, STS_From);
    m.lg1(MOVE_Rest);
    m.st16(pk, PK_HP, maxhp);
  } else {
    m.st16(pk, PK_HP, hp + maxhp / 2 < maxhp ? hp + maxhp / 2 : maxhp);
  }
  m.lg_hp(LG_Heal, id, pk, rest ? (uint32_t)HEAL_Silent : (uint32_t)HEAL_None);
}

// Effects.substitute :2386-2413; returns true when the user is left with 0 HP (residual suppressed)
template <class M>
OAK_HD_NOINLINE bool mem_substitute(M m, int p) {
  const uint32_t lo = m.vlo(p);
  const uint32_t id = M::LOGGING ? m.ident(p) : 0u;
  if (lo & VL_SUBSTITUTE) {
    m.lg3(LG_Fail, id, FAIL_Substitute);
    return false;
  }
  const PH pk = m.stored(p);
  const uint32_t maxhp = m.ld16(pk, PK_STATS + ST_HP), hp = m.ld16(pk, PK_HP);
  const uint32_t cost = (maxhp / 4) & 255u;
  if (hp < (maxhp - 1) / 4 + 1) {
    m.lg3(LG_Fail, id, FAIL_Weak);
    return false;
  }
  m.st16(pk, PK_HP, hp - cost);
  m.set_vlo(p, lo | VL_SUBSTITUTE);
  m.set_vhi_field(p, VHI_SUBSTITUTE, cost + 1);
  m.lg3(LG_Start, id, START_Substitute);
  if (M::LOGGING && cost > 0) m.lg_hp(LG_Damage, id, pk, DMG_None);
  return cost > 0 && hp == cost;
}

// Effects.transform :2438-2465
template <class M>
OAK_HD_NOINLINE void mem_transform(M m, int p) {
  const H s = M::sb(p), fs = M::sb(p ^ 1);
  const uint32_t ft = m.vhi_field(p ^ 1, VHI_TRANSFORM);
  m.set_vlo(p, m.vlo(p) | VL_TRANSFORM);
  m.set_vhi_field(p, VHI_TRANSFORM, ft != 0 ? ft : m.ident(p ^ 1));
  for (int k = ST_ATK; k <= ST_SPC; k += 2) m.st16(s, A_STATS + k, m.ld16(fs, A_STATS + k));
  m.st8(s, A_SPECIES, m.ld8(fs, A_SPECIES));
  m.st8(s, A_TYPES, m.ld8(fs, A_TYPES));
  m.st32(s, A_BOOSTS, m.ld32(fs, A_BOOSTS));
  for (int i = 0; i < 4; i++) {
    const uint32_t id = m.ld8(fs, A_MOVES + 2 * i);
    m.st8(s, A_MOVES + 2 * i, id);
    m.st8(s, A_MOVES + 2 * i + 1, id != MOVE_None ? 5 : 0);
  }
  if (M::LOGGING) m.lg3(LG_Transform, m.ident(p), m.ident(p ^ 1));
}

}  // namespace oakb200
