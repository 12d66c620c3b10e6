// libpkmn's C interface (include/oakb200_pkmn.h; reference engine/src/lib/c.zig:4-46, bindings/c.zig) on top of the
// batched engine: a batch of one on a context owned by the calling thread. Same kernels as the batched path; no CPU
// engine is involved. The pure helpers (choice / result accessors, PSRNG, rational) are host arithmetic.
#include "../../include/oakb200_pkmn.h"

#include <cstdlib>
#include <cstring>

#include "context.h"

namespace {

// our layout of the caller-owned, 128-byte, alignment-1 pkmn_gen1_battle_options (fields accessed with memcpy)
enum : size_t {
  O_LOG_BUF = 0,     // uint8_t*
  O_LOG_LEN = 8,     // size_t
  O_LOG_POS = 16,    // size_t: the writer's position (protocol.zig:831-863)
  O_LOG_TAG = 24,    // uint64_t: LOG_TAG ^ buf when a log buffer is installed (Oak leaves `options` uninitialised, mcts.h:135)
  O_PROB = 32,       // pkmn_rational {f64 p, f64 q}
  O_ACTIONS = 48,    // 16
  O_DURATIONS = 64,  // 8
  O_OVERRIDES = 72,  // 16
  O_SUMMARIES = 88,  // 12 (gen1/calc.zig:38-46), always zero
  O_HAS_CALC = 100,
  O_END = 101,
};
static_assert(O_END <= PKMN_GEN1_BATTLE_OPTIONS_SIZE, "fits pkmn_gen1_battle_options");
const uint64_t LOG_TAG = 0x4F414B4C4F470001ull;

template <class T>
T get(const uint8_t* o, size_t off) {
  T v;
  memcpy(&v, o + off, sizeof(T));
  return v;
}
template <class T>
void put(uint8_t* o, size_t off, T v) {
  memcpy(o + off, &v, sizeof(T));
}
bool has_log(const uint8_t* o) { return get<uint64_t>(o, O_LOG_TAG) == (LOG_TAG ^ (uint64_t)(uintptr_t)get<uint8_t*>(o, O_LOG_BUF)); }

// one context per calling thread, destroyed with the thread: the reference is re-entrant on disjoint battles
// (SURVEY §8b Threading), so N threads must not serialise on one context
struct ThreadCtx {
  oakb200_ctx* ctx = nullptr;
  bool failed = false;
  ~ThreadCtx() {
    if (ctx) oakb200_destroy(ctx);
  }
  oakb200_ctx* get() {
    if (ctx || failed) return ctx;
    const char* dev = getenv("OAKB200_SCALAR_DEVICE");
    if (oakb200_create(dev ? atoi(dev) : 0, 1, &ctx) != OAKB200_OK) {
      if (ctx) oakb200_destroy(ctx);
      ctx = nullptr;
      failed = true;
    }
    return ctx;
  }
};
thread_local ThreadCtx t_ctx;

double gcd(double a, double b) {  // rational.zig: operands are integers held in f64
  while (b != 0) {
    const double t = a - b * (double)(long long)(a / b);
    a = b;
    b = t;
  }
  return a;
}

}  // namespace

extern "C" {

// ---- exported constants (engine/src/lib/c.zig:4-14; gen1/data.zig:20-35 in a ReleaseFast build)
const pkmn_options PKMN_OPTIONS = {true, true, true, true};
const size_t PKMN_MAX_CHOICES = 9;
const size_t PKMN_CHOICES_SIZE = 16;
const size_t PKMN_MAX_LOGS = 180;
const size_t PKMN_LOGS_SIZE = 256;
const size_t PKMN_GEN1_MAX_CHOICES = 9;
const size_t PKMN_GEN1_CHOICES_SIZE = 16;
const size_t PKMN_GEN1_MAX_LOGS = 180;
const size_t PKMN_GEN1_LOGS_SIZE = 256;

// ---- bindings/c.zig:26-56
pkmn_choice pkmn_choice_init(pkmn_choice_kind type, uint8_t data) { return (pkmn_choice)(((unsigned)type & 3u) | ((unsigned)data << 2)); }
pkmn_choice_kind pkmn_choice_type(pkmn_choice choice) { return (pkmn_choice_kind)(choice & 3u); }
uint8_t pkmn_choice_data(pkmn_choice choice) { return (uint8_t)(choice >> 2); }
pkmn_result_kind pkmn_result_type(pkmn_result result) { return (pkmn_result_kind)(result & 15u); }
pkmn_choice_kind pkmn_result_p1(pkmn_result result) { return (pkmn_choice_kind)((result >> 4) & 3u); }
pkmn_choice_kind pkmn_result_p2(pkmn_result result) { return (pkmn_choice_kind)(result >> 6); }
bool pkmn_error(pkmn_result result) { return result == 0x0C; }

// ---- bindings/c.zig:58-64, common/rng.zig:34-59,149-164
void pkmn_psrng_init(pkmn_psrng* psrng, uint64_t seed) { memcpy(psrng->bytes, &seed, 8); }
uint32_t pkmn_psrng_next(pkmn_psrng* psrng) {
  uint64_t s;
  memcpy(&s, psrng->bytes, 8);
  s = 0x5D588B656C078965ull * s + 0x269EC3ull;
  memcpy(psrng->bytes, &s, 8);
  return (uint32_t)(s >> 32);
}

// ---- bindings/c.zig:66-80
void pkmn_rational_init(pkmn_rational* r) {
  const double one = 1.0;
  memcpy(r->bytes, &one, 8);
  memcpy(r->bytes + 8, &one, 8);
}
void pkmn_rational_reduce(pkmn_rational* r) {
  double p, q;
  memcpy(&p, r->bytes, 8);
  memcpy(&q, r->bytes + 8, 8);
  const double d = gcd(p, q);
  if (d > 1) {
    p /= d;
    q /= d;
    memcpy(r->bytes, &p, 8);
    memcpy(r->bytes + 8, &q, 8);
  }
}
float64_t pkmn_rational_numerator(const pkmn_rational* r) { return get<double>(r->bytes, 0); }
float64_t pkmn_rational_denominator(const pkmn_rational* r) { return get<double>(r->bytes, 8); }

// ---- bindings/c.zig:116-148
void pkmn_gen1_battle_options_set(pkmn_gen1_battle_options* options, const pkmn_gen1_log_options* log,
                                  const pkmn_gen1_chance_options* chance, const pkmn_gen1_calc_options* calc) {
  uint8_t* o = options->bytes;
  if (log) {
    put<uint8_t*>(o, O_LOG_BUF, log->buf);
    put<size_t>(o, O_LOG_LEN, log->len);
    put<size_t>(o, O_LOG_POS, 0);
    put<uint64_t>(o, O_LOG_TAG, LOG_TAG ^ (uint64_t)(uintptr_t)log->buf);
  } else if (has_log(o)) {
    put<size_t>(o, O_LOG_POS, 0);  // writer.reset()
  }
  if (chance) {
    memcpy(o + O_PROB, chance->probability.bytes, 16);
    memcpy(o + O_ACTIONS, chance->actions.bytes, 16);
    memcpy(o + O_DURATIONS, chance->durations.bytes, 8);
  } else {  // chance.reset(): probability 1, actions cleared, durations kept (chance.zig:367-373)
    pkmn_rational_init(reinterpret_cast<pkmn_rational*>(o + O_PROB));
    memset(o + O_ACTIONS, 0, 16);
  }
  if (calc) {
    memcpy(o + O_OVERRIDES, calc->overrides.bytes, 16);
    o[O_HAS_CALC] = 1;
  } else {
    memset(o + O_OVERRIDES, 0, 16);
    o[O_HAS_CALC] = 0;
  }
  memset(o + O_SUMMARIES, 0, 12);
}

pkmn_rational* pkmn_gen1_battle_options_chance_probability(const pkmn_gen1_battle_options* options) {
  return reinterpret_cast<pkmn_rational*>(const_cast<uint8_t*>(options->bytes) + O_PROB);
}
pkmn_gen1_chance_actions* pkmn_gen1_battle_options_chance_actions(const pkmn_gen1_battle_options* options) {
  return reinterpret_cast<pkmn_gen1_chance_actions*>(const_cast<uint8_t*>(options->bytes) + O_ACTIONS);
}
pkmn_gen1_chance_durations* pkmn_gen1_battle_options_chance_durations(const pkmn_gen1_battle_options* options) {
  return reinterpret_cast<pkmn_gen1_chance_durations*>(const_cast<uint8_t*>(options->bytes) + O_DURATIONS);
}
pkmn_gen1_calc_summaries* pkmn_gen1_battle_options_calc_summaries(const pkmn_gen1_battle_options* options) {
  return reinterpret_cast<pkmn_gen1_calc_summaries*>(const_cast<uint8_t*>(options->bytes) + O_SUMMARIES);
}

// ---- bindings/c.zig:174-184
pkmn_result pkmn_gen1_battle_update(pkmn_gen1_battle* battle, pkmn_choice c1, pkmn_choice c2, pkmn_gen1_battle_options* options) {
  oakb200_ctx* ctx = t_ctx.get();
  if (!ctx) return PKMN_RESULT_ERROR;
  uint8_t* o = options ? options->bytes : nullptr;
  uint8_t result = 0, actions[16];
  uint8_t* log = nullptr;
  uint32_t cap = 0, len = 0;
  if (o && has_log(o)) {
    const size_t total = get<size_t>(o, O_LOG_LEN), pos = get<size_t>(o, O_LOG_POS);
    log = get<uint8_t*>(o, O_LOG_BUF) + pos;
    cap = (uint32_t)(total - pos < 256 ? total - pos : 256);
  }
  if (oakb200_scalar_update(ctx, battle->bytes, c1, c2, o ? o + O_DURATIONS : nullptr, (o && o[O_HAS_CALC]) ? o + O_OVERRIDES : nullptr,
                            actions, &result, log, cap, &len) != OAKB200_OK)
    return PKMN_RESULT_ERROR;
  if (o) {
    memcpy(o + O_ACTIONS, actions, 16);
    if (log) put<size_t>(o, O_LOG_POS, get<size_t>(o, O_LOG_POS) + len);
  }
  return result;
}

// ---- bindings/c.zig:186-198
uint8_t pkmn_gen1_battle_choices(const pkmn_gen1_battle* battle, pkmn_player player, pkmn_choice_kind request, pkmn_choice out[],
                                 size_t len) {
  if ((unsigned)player > 1 || (unsigned)request > 2 || !out || len == 0) return 0;
  oakb200_ctx* ctx = t_ctx.get();
  if (!ctx) return 0;
  // the batched kernel derives both players' request kinds from a result byte: put `request` in this player's field
  const uint8_t result = (uint8_t)((unsigned)request << (player ? 6 : 4));
  uint8_t choices[18], counts[2];
  if (oakb200_scalar_choices(ctx, battle->bytes, result, choices, counts) != OAKB200_OK) return 0;
  const uint8_t n = counts[player];
  memcpy(out, choices + 9 * player, n < len ? n : len);  // at most len entries; the return value is how many there are
  return n;
}

}  // extern "C"
