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

// The five bandits of cpp/include/search/bandit/: ucb.h:15-66, ucb1.h:15-75, pucb.h:15-78, exp3.h:17-75, pexp3.h:17-82.
enum Kind { K_UCB = 0, K_UCB1 = 1, K_PUCB = 2, K_EXP3 = 3, K_PEXP3 = 4 };
struct Params {
  int kind = K_UCB;
  float c = 1.0f, gamma = 1.0f, alpha = 0.05f;
};
struct Bandit {
  std::array<float, 9> scores{};  // gains for the Exp3 family
  std::array<float, 9> priors{};
  std::array<uint32_t, 9> visits{};
  uint8_t k = 0;
  void init(int n, const Params& pr) {
    k = (uint8_t)n;
    if (pr.kind == K_UCB || pr.kind == K_PUCB) {
      std::fill(scores.begin(), scores.begin() + n, 0.5f);
      std::fill(visits.begin(), visits.begin() + n, 1u);
      std::fill(priors.begin(), priors.begin() + n, 1.0f / n);  // until softmax_logits arrives (same evaluation)
    } else if (pr.kind == K_UCB1) {
      std::fill(scores.begin(), scores.begin() + n, 0.0f);
      std::fill(visits.begin(), visits.begin() + n, 0u);
    } else {
      std::fill(scores.begin(), scores.begin() + n, 0.0f);
      std::fill(scores.begin() + n, scores.end(), -INFINITY);
    }
  }
  void softmax_logits(const Params& pr, const float* logits) {
    if (pr.kind == K_PUCB) {  // util/softmax.h:5-15
      float sum = 0;
      for (int i = 0; i < k; ++i) {
        priors[i] = std::exp(logits[i]);
        sum += priors[i];
      }
      for (int i = 0; i < k; ++i) priors[i] /= sum;
    } else if (pr.kind == K_PEXP3) {  // pexp3.h:44-48
      const float eta = pr.gamma / k;
      for (int i = 0; i < k; ++i) scores[i] = logits[i] / eta;
    }
  }
  int select(const Params& pr, double u, float& prob) const {
    prob = 1;
    int index = 0;
    if (k == 1) return 0;
    if (pr.kind == K_UCB || pr.kind == K_PUCB) {
      uint64_t N = 0;
      for (int i = 0; i < k; ++i) N += visits[i];
      float sqrtN = std::sqrt(N);
      float max = 0;
      for (int i = 0; i < k; ++i) {
        float e = pr.kind == K_UCB ? pr.c * sqrtN / k : pr.c * priors[i] * sqrtN;
        float a = (e + scores[i]) / visits[i];
        if (a > max) {
          max = a;
          index = i;
        }
      }
    } else if (pr.kind == K_UCB1) {
      std::array<float, 9> q{};
      uint64_t N{};
      for (int i = k - 1; i >= 0; --i) {
        if (visits[i] == 0) return i;
        q[i] = (float)scores[i] / visits[i];
        N += visits[i];
      }
      float lnN = std::log(N);
      float max = 0;
      for (int i = 0; i < k; ++i) {
        float e = std::sqrt(pr.c * lnN / visits[i]);
        float a = e + q[i];
        if (a > max) {
          max = a;
          index = i;
        }
      }
    } else {
      std::array<float, 9> policy;
      const float eta{pr.gamma / k};
      const float delta{pr.alpha / k};
      float sum = 0;  // util/softmax.h:17-27 over all nine gains
      for (int i = 0; i < 9; ++i) {
        policy[i] = std::exp(scores[i] * eta);
        sum += policy[i];
      }
      for (int i = 0; i < 9; ++i) policy[i] = (1 - pr.alpha) * (policy[i] / sum) + delta;
      double p = u;  // sample_pdf, util/random.h:39-49
      uint32_t pick = 0;
      for (uint32_t i = 0; i < 9; ++i) {
        p -= static_cast<double>(policy[i]);
        if (p <= 0) {
          pick = i;
          break;
        }
      }
      index = std::min<int>((int)pick, k - 1);
      prob = policy[index];
    }
    return index;
  }
  void update(const Params& pr, int index, float value, float prob) {
    if (pr.kind <= K_PUCB) {
      scores[index] += value;
      ++visits[index];
    } else if ((scores[index] += (value - 0.5f) / prob) > 0) {
      const float max = scores[index];
      for (auto& v : scores) v -= max;
    }
  }
};

struct Node {
  using Key = std::tuple<uint8_t, uint8_t, std::array<uint8_t, 16>>;
  Bandit p1, p2;
  std::map<Key, Node> children;
  bool is_init() const { return p1.k != 0; }
};

struct Search {
  float c = 1.0f;  // UCB's c when `params` is left at its default
  Params params;
  bool params_set = false;
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

  double exp3_u(size_t depth, int player) const {  // the product's documented stream (search.cu header)
    const uint64_t x = splitmix64(seed ^ (0xD1B54A32D192ED03ull * (2 * iterations + 1)) ^ (0x9FB21C651E98DF25ull * (2 * (uint64_t)depth + player + 1)));
    return (double)(x >> 11) * (1.0 / 9007199254740992.0);
  }
  bool contextual() const { return params.kind == K_PUCB || params.kind == K_PEXP3; }

  std::pair<float, float> run_iteration(Node& heap, Battle& battle, uint8_t& result, size_t depth) {
    if (heap.is_init()) {
      float prob1, prob2;
      const int i = heap.p1.select(params, exp3_u(depth, 0), prob1), j = heap.p2.select(params, exp3_u(depth, 1), prob2);
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
      heap.p1.update(params, i, v.first, prob1);
      heap.p2.update(params, j, v.second, prob2);
      if (depth == 0) {
        ++visit[i * 9 + j];
        value[i * 9 + j] += v.first;
      }
      return v;
    }
    switch (result & 15) {
      case 0: {
        uint8_t ch1[9], ch2[9];
        OakEngine e(battle, options);
        const int m = e.choices(P1, (result >> 4) & 3, ch1), n = e.choices(P2, (result >> 6) & 3, ch2);
        heap.p1.init(m, params);
        heap.p2.init(n, params);
        uint8_t dur[8];
        memcpy(dur, options.durations, 8);
        if (net) {
          float l1[9], l2[9];
          const bool ctx = contextual();  // value_policy_inference + softmax_logits (mcts.h:415-421), else value_inference
          const float v = net_discrete ? net->eval_discrete(battle, dur, ctx ? m : 0, ctx ? n : 0, ch1, ch2, l1, l2)
                                       : net->eval(battle, dur, ctx ? m : 0, ctx ? n : 0, ch1, ch2, l1, l2);
          if (ctx) {
            heap.p1.softmax_logits(params, l1);
            heap.p2.softmax_logits(params, l2);
          }
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
    if (!params_set) {
      params.kind = K_UCB;
      params.c = c;
    }
    if (!root.is_init()) {  // mcts.h:186-209
      Battle b;
      memcpy(&b, battle384, 384);
      Options o{};
      OakEngine e(b, o);
      uint8_t ch1[9], ch2[9];
      const int m = e.choices(P1, (result0 >> 4) & 3, ch1), n = e.choices(P2, (result0 >> 6) & 3, ch2);
      root.p1.init(m, params);
      root.p2.init(n, params);
      if (contextual() && net) {
        float l1[9], l2[9];
        if (net_discrete) net->eval_discrete(b, durations8, m, n, ch1, ch2, l1, l2);
        else net->eval(b, durations8, m, n, ch1, ch2, l1, l2);
        root.p1.softmax_logits(params, l1);
        root.p2.softmax_logits(params, l2);
      }
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
