// ORACLE — TEST INFRASTRUCTURE ONLY.
// A pybind11 module around the REFERENCE'S OWN training-row builder, compiled in place from
// /root/reference/cpp/include/py/battle/encoded-frames.h (never copied): Py::Battle::EncodedFrames::from_bytes replays one
// stored game and writes the dense rows the learner trains on. The reference links that header against libpkmn (a Zig
// build that cannot be made here); the handful of libpkmn C symbols it calls are supplied below by the oracle's engine
// (gen1_oracle.hpp, itself pinned to libpkmn's traces), so what this module pins is the reference's encoders, policy
// index, frame reader and replay glue. Built into oracle/_ref/ref_pyframes*.so by oracle/Makefile where the reference
// tree exists; tools/gen_training_rows_golden.py uses it to write tests/golden/training_rows_golden.npz.
#include <cstdint>
#include <cstring>

#include "gen1_oracle.hpp"

#include <pkmn.h>

namespace {
struct Opts {  // our layout of the opaque 128-byte options
  oracle::Options o;
};
static_assert(sizeof(Opts) <= PKMN_GEN1_BATTLE_OPTIONS_SIZE, "fits");
}  // namespace

extern "C" {
const size_t PKMN_GEN1_MAX_CHOICES = 9;
const size_t PKMN_GEN1_CHOICES_SIZE = 9;
const size_t PKMN_GEN1_MAX_LOGS = 180;
const size_t PKMN_GEN1_LOGS_SIZE = 180;
pkmn_result_kind pkmn_result_type(pkmn_result r) { return (pkmn_result_kind)(r & 15u); }
pkmn_choice_kind pkmn_result_p1(pkmn_result r) { return (pkmn_choice_kind)((r >> 4) & 3u); }
pkmn_choice_kind pkmn_result_p2(pkmn_result r) { return (pkmn_choice_kind)(r >> 6); }
void pkmn_gen1_battle_options_set(pkmn_gen1_battle_options* options, const pkmn_gen1_log_options*, const pkmn_gen1_chance_options* chance,
                                  const pkmn_gen1_calc_options* calc) {
  Opts* p = reinterpret_cast<Opts*>(options->bytes);
  if (chance) {
    memcpy(p->o.actions, chance->actions.bytes, 16);
    memcpy(p->o.durations, chance->durations.bytes, 8);
  } else {
    memset(p->o.actions, 0, 16);  // durations are kept (chance.zig:367-373)
  }
  if (calc) memcpy(p->o.overrides, calc->overrides.bytes, 16);
  else memset(p->o.overrides, 0, 16);
}
pkmn_gen1_chance_durations* pkmn_gen1_battle_options_chance_durations(const pkmn_gen1_battle_options* options) {
  Opts* p = reinterpret_cast<Opts*>(const_cast<uint8_t*>(options->bytes));
  return reinterpret_cast<pkmn_gen1_chance_durations*>(p->o.durations);
}
pkmn_gen1_chance_actions* pkmn_gen1_battle_options_chance_actions(const pkmn_gen1_battle_options* options) {
  Opts* p = reinterpret_cast<Opts*>(const_cast<uint8_t*>(options->bytes));
  return reinterpret_cast<pkmn_gen1_chance_actions*>(p->o.actions);
}
pkmn_result pkmn_gen1_battle_update(pkmn_gen1_battle* battle, pkmn_choice c1, pkmn_choice c2, pkmn_gen1_battle_options* options) {
  oracle::Battle& b = *reinterpret_cast<oracle::Battle*>(battle->bytes);
  Opts scratch{};
  Opts* p = options ? reinterpret_cast<Opts*>(options->bytes) : &scratch;
  oracle::OakEngine e(b, p->o);
  return e.update(c1, c2);
}
uint8_t pkmn_gen1_battle_choices(const pkmn_gen1_battle* battle, pkmn_player player, pkmn_choice_kind request, pkmn_choice out[], size_t len) {
  oracle::Battle b;
  memcpy(&b, battle->bytes, 384);
  oracle::Options o{};
  oracle::OakEngine e(b, o);
  uint8_t ch[9];
  const uint8_t n = e.choices(player, request, ch);
  for (uint8_t i = 0; i < n && i < len; i++) out[i] = ch[i];
  return n;
}
}

#include <py/battle/encoded-frames.h>
#include <py/battle/frames.h>

namespace py = pybind11;

PYBIND11_MODULE(ref_pyframes, m) {
  m.doc() = "reference EncodedFrames::from_bytes (test oracle)";
  // one stored game -> dict of the arrays EncodedBattleFrames exposes (pyoak.cc:655-673)
  m.def("from_bytes", [](const py::bytes& data, size_t sz) {
    auto f = Py::Battle::EncodedFrames::from_bytes(data, sz);
    py::dict d;
    d["k"] = f.k;
    d["choice"] = f.choice;
    d["iterations"] = f.iterations;
    d["empirical_policies"] = f.empirical_policies;
    d["nash_policies"] = f.nash_policies;
    d["empirical_value"] = f.empirical_value;
    d["nash_value"] = f.nash_value;
    d["score"] = f.score;
    d["pokemon"] = f.pokemon;
    d["active"] = f.active;
    d["hp"] = f.hp;
    d["choice_indices"] = f.choice_indices;
    return d;
  });
  // the uncompressed form (pyoak BattleFrames, pyoak.cc:638-653)
  m.def("frames_from_bytes", [](const py::bytes& data, size_t sz) {
    auto f = Py::Battle::Frames::from_bytes(data, sz);
    py::dict d;
    d["k"] = f.k;
    d["choice"] = f.choice;
    d["iterations"] = f.iterations;
    d["empirical_policies"] = f.empirical_policies;
    d["nash_policies"] = f.nash_policies;
    d["empirical_value"] = f.empirical_value;
    d["nash_value"] = f.nash_value;
    d["score"] = f.score;
    d["battle"] = f.battle;
    d["durations"] = f.durations;
    d["result"] = f.result;
    d["choices"] = f.choices;
    return d;
  });
}
