// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference's training-frame serialisation and of the
// policy post-processing that picks the recorded choices. Pinned against the reference's own header compiled in place
// (oracle/ref_shim.cpp: ref_frames_write_game) by tests/test_oracle_frames.py and the committed golden stream.
//   compress_probs            cpp/include/train/battle/compressed-frame.h:12-24   (x * 65535, truncated to u16)
//   Update::write / n_bytes   compressed-frame.h:58-112
//   CompressedFrames::write   compressed-frame.h:173-199
//   softmax                   cpp/include/search/util/softmax.h:5-15
//   get_policy tail + sample  cpp/include/util/policy.h:76-107, cpp/include/util/random.h:41-50
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>

namespace frames_oracle {

inline uint16_t compress(double x) { return (uint16_t)(x * 65535.0); }

// -> bytes written
inline uint32_t write_game(const uint8_t* battle, uint8_t result, uint32_t n_updates, const uint8_t* k, const uint8_t* choice,
                           const uint32_t* iterations, const double* emp_value, const double* nash_value, const double* emp,
                           const double* nash, uint8_t* out) {
  uint32_t at = 4 + 2;
  memcpy(out + at, battle, 384);
  at += 384;
  out[at++] = result;
  auto put16 = [&](uint16_t v) { memcpy(out + at, &v, 2); at += 2; };
  for (uint32_t u = 0; u < n_updates; u++) {
    const uint32_t m = k[2 * u], n = k[2 * u + 1];
    out[at++] = (uint8_t)((m - 1) | ((n - 1) << 4));
    out[at++] = choice[2 * u];
    out[at++] = choice[2 * u + 1];
    memcpy(out + at, &iterations[u], 4);
    at += 4;
    put16(compress(emp_value[u]));
    put16(compress(nash_value[u]));
    for (uint32_t i = 0; i < m; i++) put16(compress(emp[(u * 2 + 0) * 9 + i]));
    for (uint32_t i = 0; i < m; i++) put16(compress(nash[(u * 2 + 0) * 9 + i]));
    for (uint32_t i = 0; i < n; i++) put16(compress(emp[(u * 2 + 1) * 9 + i]));
    for (uint32_t i = 0; i < n; i++) put16(compress(nash[(u * 2 + 1) * 9 + i]));
  }
  memcpy(out, &at, 4);
  const uint16_t nu = (uint16_t)n_updates;
  memcpy(out + 4, &nu, 2);
  return at;
}

// softmax over k logits (fp32), then temperature / floor / renormalise in double and sample_pdf with the given u
inline int policy_sample(const float* logits, int k, double temp, double floor_p, double u, float* probs9) {
  float sum = 0;
  for (int i = 0; i < 9; i++) probs9[i] = 0;
  for (int i = 0; i < k; i++) {
    probs9[i] = std::exp(logits[i]);
    sum += probs9[i];
  }
  double pol[9] = {};
  for (int i = 0; i < k; i++) {
    probs9[i] /= sum;
    pol[i] = probs9[i];
  }
  if (temp != 1) {
    double s = 0;
    for (auto& x : pol) {
      x = std::pow(x, temp);
      s += x;
    }
    for (auto& x : pol) x /= s;
  }
  double s = 0;
  for (auto& x : pol) {
    if (x < floor_p) x = 0;
    s += x;
  }
  if (s == 0) return 0;
  for (auto& x : pol) x /= s;
  for (int i = 0; i < 9; i++) {
    u -= pol[i];
    if (u <= 0) return i < k ? i : 0;
  }
  return 0;
}

}  // namespace frames_oracle
