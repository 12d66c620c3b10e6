// ORACLE — TEST INFRASTRUCTURE ONLY. Sequential CPU restatement of the reference's tree search for one root:
//   MCTS::Search::run / run_root_iteration / run_iteration / init_stats_and_rollout  cpp/include/search/mcts.h:154-496
//   Search::battle_options_set + roll_byte (clamped damage rolls)                     mcts.h:569-604
//   UCB::Bandit, Joint<Bandit>                                                        cpp/include/search/bandit/ucb.h:15-66, joint.h
//   Node heap keyed by (p1 index, p2 index, chance observation)                       mcts.h:93-99,355-361
//   process_output's empirical part                                                   mcts.h:623-661
// Recursive like the reference, on the oracle's own engine. The reference draws its random numbers from one
// std::mt19937 (util/random.h); here, as in the product's ABI, iteration n of a root with seed s uses
// battle rng = splitmix64(s ^ K(2n+1)) and playout seed = splitmix64(s ^ K(2n+2)), K = 0xD1B54A32D192ED03, and the
// playout picks with oak_pick. Parity unpinned against the reference binary (its statistical KATs,
// cpp/src/search-test.cc, need the unbuildable programs); two of those KAT positions are checked for their analytic
// values through the product in tests/test_gpu_search.py.
#pragma once

#include <cmath>
#include <cstring>
#include <map>
#include <tuple>
#include <array>

#include "gen1_oracle.hpp"
#include "oak_oracle.hpp"
#include "nn_oracle.hpp"

namespace mcts_oracle {

using namespace oracle;

struct Bandit {
  std::array<float, 9> scores{};
  std::array<uint32_t, 9> visits{};
  uint8_t k = 0;
  void init(int n) {
    k = (uint8_t)n;
    std::fill(scores.begin(), scores.begin() + n, 0.5f);
    std::fill(visits.begin(), visits.begin() + n, 1u);
  }
  int select(float c) const {
    int index = 0;
    if (k == 1) return 0;
    uint64_t N = 0;
    for (int i = 0; i < k; ++i) N += visits[i];
    float sqrtN = std::sqrt(N);
    float max = 0;
    for (int i = 0; i < k; ++i) {
      float e = c * sqrtN / k;
      float a = (e + scores[i]) / visits[i];
      if (a > max) {
        max = a;
        index = i;
      }
    }
    return index;
  }
  void update(int index, float value) {
    scores[index] += value;
    ++visits[index];
  }
};

struct Node {
  using Key = std::tuple<uint8_t, uint8_t, std::array<uint8_t, 16>>;
  Bandit p1, p2;
  std::map<Key, Node> children;
  bool is_init() const { return p1.k != 0; }
};

struct Search {
  float c;
  uint32_t root_rolls, other_rolls;
  uint64_t seed, iterations = 0;
  Node root;
  uint32_t visit[81] = {0};
  double value[81] = {0};
  uint64_t nodes = 1;
  // per-iteration
  Options options{};
  uint64_t playout_seed = 0;
  // leaf evaluation: null = MonteCarlo playout, else the network's value (mcts.h:409-424, UCB is not a contextual bandit:
  // value_inference only) with the durations the search has tracked down to the leaf
  const BattleNet* net = nullptr;
  bool net_discrete = false;

  static uint8_t roll_byte(uint32_t n_rolls, uint8_t s) {
    if (n_rolls == 1) return 236;
    return (uint8_t)(217 + (38 / (n_rolls - 1)) * (s % n_rolls));
  }
  void options_set(const Battle& b, size_t depth) {
    options.actions[0] = options.actions[1] = 0;  // chance reset keeps the durations
    options.overrides[0] = options.overrides[1] = 0;
    const uint32_t nr = depth == 0 ? root_rolls : other_rolls;
    if (root_rolls == 39 && other_rolls == 39) return;  // !Options.clamping
    uint8_t rb[8];
    memcpy(rb, &b.rng, 8);
    options.overrides[0] = roll_byte(nr, rb[6]);
    options.overrides[1] = roll_byte(nr, rb[7]);
  }

  std::pair<float, float> run_iteration(Node& heap, Battle& battle, uint8_t& result, size_t depth) {
    if (heap.is_init()) {
      const int i = heap.p1.select(c), j = heap.p2.select(c);
      uint8_t ch1[9], ch2[9];
      {
        OakEngine e(battle, options);
        e.choices(P1, (result >> 4) & 3, ch1);
        e.choices(P2, (result >> 6) & 3, ch2);
      }
      options_set(battle, depth);
      {
        OakEngine e(battle, options);
        result = e.update(ch1[i], ch2[j]);
      }
      std::array<uint8_t, 16> obs;
      memcpy(obs.data(), options.actions, 16);
      auto it = heap.children.find({(uint8_t)i, (uint8_t)j, obs});
      if (it == heap.children.end()) {
        it = heap.children.emplace(Node::Key{(uint8_t)i, (uint8_t)j, obs}, Node{}).first;
        nodes++;
      }
      const auto v = run_iteration(it->second, battle, result, depth + 1);
      heap.p1.update(i, v.first);
      heap.p2.update(j, v.second);
      if (depth == 0) {
        ++visit[i * 9 + j];
        value[i * 9 + j] += v.first;
      }
      return v;
    }
    switch (result & 15) {
      case 0: {
        uint8_t ch[9];
        OakEngine e(battle, options);
        const int m = e.choices(P1, (result >> 4) & 3, ch), n = e.choices(P2, (result >> 6) & 3, ch);
        heap.p1.init(m);
        heap.p2.init(n);
        uint8_t dur[8];
        memcpy(dur, options.durations, 8);
        if (net) {
          float l1[9], l2[9];
          const float v = net_discrete ? net->eval_discrete(battle, dur, 0, 0, nullptr, nullptr, l1, l2) : net->eval(battle, dur, 0, 0, nullptr, nullptr, l1, l2);
          return {v, 1 - v};
        }
        oak_rollout(battle, dur, &result, playout_seed, 100000);
        const float v = (result & 15) == RES_WIN ? 1.0f : ((result & 15) == RES_LOSE ? 0.0f : 0.5f);
        return {v, 1 - v};
      }
      case RES_WIN: return {1, 0};
      case RES_LOSE: return {0, 1};
      default: return {.5, .5};
    }
  }

  void run(const uint8_t* battle384, const uint8_t* durations8, uint8_t result0, uint32_t n_iter) {
    if (!root.is_init()) {  // mcts.h:186-188
      Battle b;
      memcpy(&b, battle384, 384);
      Options o{};
      OakEngine e(b, o);
      uint8_t ch[9];
      const int m = e.choices(P1, (result0 >> 4) & 3, ch), n = e.choices(P2, (result0 >> 6) & 3, ch);
      root.p1.init(m);
      root.p2.init(n);
    }
    for (uint32_t t = 0; t < n_iter; t++) {
      Battle b;
      memcpy(&b, battle384, 384);
      const uint64_t nth = iterations;
      b.rng = splitmix64(seed ^ (0xD1B54A32D192ED03ull * (2 * nth + 1)));
      playout_seed = splitmix64(seed ^ (0xD1B54A32D192ED03ull * (2 * nth + 2)));
      oak_randomize_hidden_variables(b, durations8);
      options = Options{};
      memcpy(options.durations, durations8, 8);
      uint8_t result = result0;
      run_iteration(root, b, result, 0);
      iterations++;
    }
  }
};

}  // namespace mcts_oracle
