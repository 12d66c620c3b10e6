// Batched-leaf tree search: host trees, GPU transitions and leaves (BASELINE configs[3]).
//
// Reference algorithm: MCTS::Search::{run, run_root_iteration, run_iteration, init_stats_and_rollout}
// (cpp/include/search/mcts.h:154-496) with its joint bandits (cpp/include/search/bandit/{ucb,ucb1,pucb,exp3,pexp3}.h,
// cpp/include/search/joint.h; the contextual ones primed with the network's policy logits, mcts.h:190-209,409-421) on a Node heap keyed by (p1 index, p2 index, 16-byte chance observation)
// (mcts.h:93-99,355-361), MonteCarlo evaluation, Search::battle_options_set's clamped damage rolls
// (mcts.h:569-604; default_search = {root_rolls 3, other_rolls 1}, mcts.h:133).
//
// What stays on the host (as the north star says): the trees and the bandit arithmetic. What moves to the GPU: every
// battle operation. R roots x P particles descend in lockstep; per tree level one search_step_kernel launch advances
// all of them (choices, clamped update, observation, child sizes), then the hidden-variable draw and the playouts of
// all leaves are one launch each. With P = 1 the search of each root is the reference's sequential algorithm
// exactly (same selections, same statistics) for the documented random streams below; with P > 1 the particles of a
// wave select with virtual visits (the visit of an in-flight descent is counted at selection time, its value at
// back-propagation), which changes the order of selections, not the statistics kept.
//
// Random streams (the reference draws them from one std::mt19937 per thread, cpp/include/util/random.h): for root
// seed s and that root's n-th iteration, battle rng = splitmix64(s ^ K * (2n + 1)), playout seed =
// splitmix64(s ^ K * (2n + 2)), K = 0xD1B54A32D192ED03; the playout then picks as oakb200_rollout does; the Exp3 family's
// selection at depth d for player q draws u = (splitmix64(s ^ K(2n+1) ^ 0x9FB21C651E98DF25 (2d + q + 1)) >> 11) 2^-53.
// Not built: the root matrix-UCB variant (mcts.h:498-566), which needs the LRSNash solver inside the iteration loop.
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cmath>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "context.h"

namespace {

uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
const uint64_t STREAM_K = 0xD1B54A32D192ED03ull;

// The reference's bandits (cpp/include/search/bandit/{ucb,ucb1,pucb,exp3,pexp3}.h) as one record with the algorithm as
// data: the trees of one search all use the same one. a[] = scores (UCB family) or gains (Exp3 family).
enum BanditKind { B_UCB = 0, B_UCB1 = 1, B_PUCB = 2, B_EXP3 = 3, B_PEXP3 = 4 };
struct BanditParams {
  int kind = B_UCB;
  float c = 1.0f;                  // UCB / UCB1 / PUCB
  float gamma = 1.0f, alpha = 0.05f;  // Exp3 / PExp3 (cpp/src/search.cc:279-300)
  bool contextual() const { return kind == B_PUCB || kind == B_PEXP3; }
  bool counts_visits() const { return kind <= B_PUCB; }
};
struct Bandit {
  uint8_t k = 0;  // first: a node's key, both counts and P1's scores share the node's first cache line
  float a[9];
  uint32_t visits[9];
  // PUCB's priors live in Root::priors (one float[18] per node, allocated only for the contextual kinds)
  void init(int n, const BanditParams& bp, float* priors) {
    k = (uint8_t)n;
    switch (bp.kind) {
      case B_UCB: case B_PUCB:  // ucb.h:28-32, pucb.h:30-34
        for (int i = 0; i < n; i++) { a[i] = 0.5f; visits[i] = 1; }
        if (bp.kind == B_PUCB)
          for (int i = 0; i < n; i++) priors[i] = 1.0f / n;  // until the leaf evaluation of this node arrives
        break;
      case B_UCB1:  // ucb1.h:31-35
        for (int i = 0; i < n; i++) { a[i] = 0.0f; visits[i] = 0; }
        break;
      default:  // exp3.h:36-40
        for (int i = 0; i < 9; i++) a[i] = i < n ? 0.0f : -INFINITY;
        break;
    }
  }
  // softmax_logits: pucb.h:38-40 (priors = softmax of the legal logits), pexp3.h:44-48 (gains = logits / eta)
  void set_logits(const float* logits, const BanditParams& bp, float* priors) {
    if (bp.kind == B_PUCB) {
      float sum = 0;
      for (int i = 0; i < k; i++) { priors[i] = std::exp(logits[i]); sum += priors[i]; }
      for (int i = 0; i < k; i++) priors[i] /= sum;
    } else if (bp.kind == B_PEXP3) {
      const float eta = bp.gamma / k;
      for (int i = 0; i < k; i++) a[i] = logits[i] / eta;
    }
  }
  // u: the uniform draw of this selection (Exp3 family only); *prob: probability of the selected arm (Exp3 family)
  int select(const BanditParams& bp, double u, float* prob, const float* priors) const {
    *prob = 1.0f;
    if (k == 1) return 0;
    int index = 0;
    if (bp.kind == B_UCB || bp.kind == B_PUCB) {  // ucb.h:41-61, pucb.h:47-67
      uint64_t N = 0;
      for (int i = 0; i < k; i++) N += visits[i];
      const float sqrtN = (float)std::sqrt((double)N);
      // the same operations in the same order as the reference, arm by arm; values first (a loop the compiler
      // vectorises), then the first maximum
      float v[9];
      if (bp.kind == B_UCB) {
        const float e = bp.c * sqrtN / k;
        for (int i = 0; i < k; i++) v[i] = (e + a[i]) / (float)visits[i];
      } else {
        for (int i = 0; i < k; i++) v[i] = (bp.c * priors[i] * sqrtN + a[i]) / (float)visits[i];
      }
      float max = 0;
      for (int i = 0; i < k; i++)
        if (v[i] > max) { max = v[i]; index = i; }
    } else if (bp.kind == B_UCB1) {  // ucb1.h:44-71
      float q[9];
      uint64_t N = 0;
      for (int i = k - 1; i >= 0; --i) {
        if (visits[i] == 0) return i;
        q[i] = a[i] / visits[i];
        N += visits[i];
      }
      const float lnN = std::log((float)N);
      float max = 0;
      for (int i = 0; i < k; i++) {
        const float v = std::sqrt(bp.c * lnN / visits[i]) + q[i];
        if (v > max) { max = v; index = i; }
      }
    } else {  // exp3.h:44-61, pexp3.h:50-67; softmax over all nine gains (the unused ones are -inf), util/softmax.h:17-27
      const float eta = bp.gamma / k, delta = bp.alpha / k;
      float policy[9], sum = 0;
      for (int i = 0; i < 9; i++) { policy[i] = std::exp(a[i] * eta); sum += policy[i]; }
      for (int i = 0; i < 9; i++) policy[i] = (1.0f - bp.alpha) * (policy[i] / sum) + delta;
      double p = u;  // sample_pdf, util/random.h:39-49
      for (int i = 0; i < 9; i++) {
        p -= (double)policy[i];
        if (p <= 0) { index = i; break; }
      }
      if (index > k - 1) index = k - 1;
      *prob = policy[index];
    }
    return index;
  }
  void update(int index, float value, float prob, const BanditParams& bp) {
    if (bp.counts_visits()) {  // the visit itself was counted at selection (virtual visit)
      a[index] += value;
    } else if ((a[index] += (value - 0.5f) / prob) > 0) {  // exp3.h:63-70
      const float max = a[index];
      for (int i = 0; i < 9; i++) a[i] -= max;
    }
  }
};

// One tree per root, nodes in one arena; a child is found through an open-addressed table keyed by
// (parent, p1 index, p2 index, 16-byte chance observation) — the reference's std::map key (mcts.h:93-99,355-361).
struct alignas(64) Node {  // 192 bytes = three lines; line 0 = key + P1's count and scores
  int32_t parent = -1;
  uint8_t i = 0, j = 0;
  uint8_t obs[16] = {0};
  Bandit p1, p2;
  bool is_init() const { return p1.k != 0; }
  void prefetch() const {
    const char* c = reinterpret_cast<const char*>(this);
    __builtin_prefetch(c, 1, 3);
    __builtin_prefetch(c + 64, 1, 3);
    __builtin_prefetch(c + 128, 1, 3);
  }
};
static_assert(sizeof(Node) == 192, "Node is three cache lines");
struct Slot {      // child table entry: the node and 32 bits of its key's hash, so that a probe that is not the key's
  int32_t node;    // rarely touches the node itself
  uint32_t tag;
};
struct PathEntry {
  int32_t node;
  uint8_t i, j;
  float prob1, prob2;
};

struct Root {
  std::vector<Node> nodes;     // nodes[0] = root
  std::vector<float> priors;   // contextual bandits only: [node][player][9]
  std::vector<Slot> table;     // open addressing, node -1 = empty; size a power of two > 2 x nodes
  float* pri(int32_t node, int player) { return priors.empty() ? nullptr : priors.data() + ((size_t)node * 2 + player) * 9; }
  // back to an empty tree over the same arenas (their pages stay mapped: a search object is reused turn after turn)
  void reset(const BanditParams& bp) {
    nodes.clear();
    priors.clear();
    std::fill(table.begin(), table.end(), Slot{-1, 0});
    nodes.emplace_back();
    if (bp.contextual()) priors.resize(18);
    nodes[0].p1.init(k1, bp, pri(0, 0));
    nodes[0].p2.init(k2, bp, pri(0, 1));
    memset(visit, 0, sizeof(visit));
    memset(value, 0, sizeof(value));
    iterations = target = total_depth = 0;
  }
  std::vector<PathEntry> path; // [level][particle] of the wave in flight
  uint8_t k1 = 0, k2 = 0;
  uint8_t choices[18] = {0};
  uint32_t visit[81] = {0};
  double value[81] = {0};
  uint64_t iterations = 0, target = 0;
  uint64_t seed = 0;
  uint64_t total_depth = 0;

  static uint64_t key_hash(int32_t parent, uint8_t i, uint8_t j, const uint8_t* obs) {
    uint64_t a, b;
    memcpy(&a, obs, 8);
    memcpy(&b, obs + 8, 8);
    uint64_t h = (uint64_t)(uint32_t)parent * 0x9E3779B97F4A7C15ull + i * 0x100000001B3ull + j * 0xD6E8FEB86659FD93ull;
    h ^= a * 0xBF58476D1CE4E5B9ull;
    h = (h << 31) | (h >> 33);
    h ^= b * 0x94D049BB133111EBull;
    return h ^ (h >> 29);
  }
  void rehash(size_t size) {
    table.assign(size, Slot{-1, 0});
    for (size_t n = 1; n < nodes.size(); n++) {
      const uint64_t h = key_hash(nodes[n].parent, nodes[n].i, nodes[n].j, nodes[n].obs);
      size_t at = h & (size - 1);
      while (table[at].node >= 0) at = (at + 1) & (size - 1);
      table[at] = Slot{(int32_t)n, (uint32_t)(h >> 32)};
    }
  }
  void reserve_table() {
    if (table.size() < 2 * (nodes.size() + 1)) rehash(table.empty() ? 256 : table.size() * 2);
  }
  // the lines child(h) will read first: the key's home slot now, and (once that line is here) the node it names
  void prefetch_slot(uint64_t h) const { __builtin_prefetch(&table[h & (table.size() - 1)], 0, 3); }
  void prefetch_child(uint64_t h) const {
    const Slot& e = table[h & (table.size() - 1)];
    if (e.node >= 0) nodes[e.node].prefetch();
  }
  // the child of `parent` along (i, j, obs), created if new; h = key_hash of the four
  int32_t child(int32_t parent, uint8_t i, uint8_t j, const uint8_t* obs, uint64_t h) {
    reserve_table();
    const size_t mask = table.size() - 1;
    const uint32_t tag = (uint32_t)(h >> 32);
    size_t at = h & mask;
    for (;; at = (at + 1) & mask) {
      const Slot e = table[at];
      if (e.node < 0) break;
      if (e.tag != tag) continue;
      const Node& c = nodes[e.node];
      if (c.parent == parent && c.i == i && c.j == j && memcmp(c.obs, obs, 16) == 0) return e.node;
    }
    const int32_t n = (int32_t)nodes.size();
    nodes.emplace_back();
    if (!priors.empty()) priors.resize(priors.size() + 18);
    Node& c = nodes.back();
    c.parent = parent;
    c.i = i;
    c.j = j;
    memcpy(c.obs, obs, 16);
    table[at] = Slot{n, tag};
    return n;
  }
};

struct Particle {
  int32_t cur = 0;
  int32_t fresh = -1;     // the node this descent initialised (its priors arrive with the leaf evaluation)
  uint16_t len = 0;       // path entries
  bool active = false, spawned = false;
};

// persistent worker pool: parallel_for over roots
class Pool {
 public:
  explicit Pool(int threads) {
    if (threads < 1) threads = 1;
    for (int t = 1; t < threads; t++) workers_.emplace_back([this, t] { loop((size_t)t); });
  }
  ~Pool() {
    {
      std::lock_guard<std::mutex> l(mu_);
      stop_ = true;
      gen_++;
    }
    cv_.notify_all();
    for (auto& w : workers_) w.join();
  }
  void parallel_for(size_t n, const std::function<void(size_t)>& fn) {
    if (workers_.empty() || n < 2) {
      for (size_t i = 0; i < n; i++) fn(i);
      return;
    }
    {
      std::lock_guard<std::mutex> l(mu_);
      fn_ = &fn;
      n_ = n;
      pending_ = (int)workers_.size();
      gen_++;
    }
    cv_.notify_all();
    work(0);
    std::unique_lock<std::mutex> l(mu_);
    done_.wait(l, [this] { return pending_ == 0; });
  }

 private:
  // Static slices: thread w (0 = the caller) always gets items [n w / T, n (w + 1) / T). The items are roots, whose trees
  // (~200 KB each) then stay in the cache of the core that walked them last, phase after phase; handing roots out one by one
  // from a shared counter moved every tree between cores at every phase (12.4 -> see profiles/r2h_search_host.txt).
  void work(size_t slot) {
    const size_t T = workers_.size() + 1;
    const size_t lo = n_ * slot / T, hi = n_ * (slot + 1) / T;
    for (size_t i = lo; i < hi; i++) (*fn_)(i);
  }
  void loop(size_t slot) {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [&] { return gen_ != seen; });
        seen = gen_;
        if (stop_) return;
      }
      work(slot);
      {
        std::lock_guard<std::mutex> l(mu_);
        if (--pending_ == 0) done_.notify_one();
      }
    }
  }
  std::vector<std::thread> workers_;
  std::mutex mu_;
  std::condition_variable cv_, done_;
  const std::function<void(size_t)>* fn_ = nullptr;
  size_t n_ = 0;
  int pending_ = 0;
  uint64_t gen_ = 0;
  bool stop_ = false;
};

}  // namespace

struct oakb200_search {
  oakb200_ctx* ctx = nullptr;
  size_t roots_max = 0, n_roots = 0;
  uint32_t P = 1;
  std::vector<Root> roots;
  std::vector<Particle> particles;
  // device
  uint8_t *d_root_b = nullptr, *d_root_d = nullptr, *d_root_r = nullptr;
  uint8_t *d_b = nullptr, *d_d = nullptr, *d_r = nullptr, *d_idx1 = nullptr, *d_idx2 = nullptr, *d_rolls = nullptr, *d_act = nullptr,
          *d_cnt = nullptr, *d_score = nullptr;
  uint64_t *d_rng = nullptr, *d_pseed = nullptr;
  // network leaves
  int leaf_mode = 0;  // 0 MonteCarlo playout, 1 network value
  float *d_value = nullptr, *d_logits = nullptr, *h_value = nullptr, *h_logits = nullptr;
  uint8_t* d_ecnt = nullptr;
  BanditParams bandit;
  bool bandit_set = false;   // false: joint UCB with the c given to oakb200_search_run
  bool root_primed = false;  // contextual bandits: the roots' priors have been evaluated (mcts.h:190-209)
  // pinned host mirrors
  uint8_t *h_idx1 = nullptr, *h_idx2 = nullptr, *h_rolls = nullptr, *h_res = nullptr, *h_act = nullptr, *h_cnt = nullptr, *h_score = nullptr;
  uint64_t *h_rng = nullptr, *h_pseed = nullptr;
  Pool* pool = nullptr;
  int pool_threads = 0;
  uint64_t levels = 0, waves = 0;
  double last_host_ms = 0, last_idle_ms = 0, last_total_ms = 0;  // of the last oakb200_search_run
  uint32_t last_waves = 0;
};

namespace {

void search_free(oakb200_search* s) {
  if (!s) return;
  if (s->ctx) cudaSetDevice(s->ctx->device);
  void* dev[] = {s->d_root_b, s->d_root_d, s->d_root_r, s->d_b, s->d_d, s->d_r, s->d_idx1, s->d_idx2, s->d_rolls, s->d_act, s->d_cnt,
                 s->d_score, s->d_rng, s->d_pseed, s->d_value, s->d_logits, s->d_ecnt};
  for (void* p : dev)
    if (p) cudaFree(p);
  void* host[] = {s->h_idx1, s->h_idx2, s->h_rolls, s->h_res, s->h_act, s->h_cnt, s->h_score, s->h_rng, s->h_pseed, s->h_value, s->h_logits};
  for (void* p : host)
    if (p) cudaFreeHost(p);
  delete s->pool;
  delete s;
}

template <class T>
cudaError_t dmalloc(T** p, size_t bytes) { return cudaMalloc(reinterpret_cast<void**>(p), bytes); }
template <class T>
cudaError_t hmalloc(T** p, size_t bytes) { return cudaMallocHost(reinterpret_cast<void**>(p), bytes); }

}  // namespace

extern "C" {

int oakb200_search_create(oakb200_ctx* ctx, size_t roots, uint32_t particles, oakb200_search** out) {
  if (!ctx || !out || roots == 0 || particles == 0 || particles > 4096 || roots * (size_t)particles > 0x7FFFFFFFull)
    return fail(ctx, OAKB200_EINVAL, "search_create: roots >= 1, 1 <= particles <= 4096");
  *out = nullptr;
  CU(ctx, cudaSetDevice(ctx->device));
  oakb200_search* s = new oakb200_search();
  s->ctx = ctx;
  s->roots_max = roots;
  s->P = particles;
  const size_t B = roots * (size_t)particles;
  cudaError_t e = cudaSuccess;
  auto ok = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
  ok(dmalloc(&s->d_root_b, roots * 384)); ok(dmalloc(&s->d_root_d, roots * 8)); ok(dmalloc(&s->d_root_r, roots));
  ok(dmalloc(&s->d_b, B * 384)); ok(dmalloc(&s->d_d, B * 8)); ok(dmalloc(&s->d_r, B));
  ok(dmalloc(&s->d_idx1, B)); ok(dmalloc(&s->d_idx2, B)); ok(dmalloc(&s->d_rolls, B));
  ok(dmalloc(&s->d_act, B * 18)); ok(dmalloc(&s->d_cnt, B * 2)); ok(dmalloc(&s->d_score, B));
  ok(dmalloc(&s->d_rng, B * 8)); ok(dmalloc(&s->d_pseed, B * 8));
  ok(hmalloc(&s->h_idx1, B)); ok(hmalloc(&s->h_idx2, B)); ok(hmalloc(&s->h_rolls, B)); ok(hmalloc(&s->h_res, B));
  ok(hmalloc(&s->h_act, B * 16)); ok(hmalloc(&s->h_cnt, B * 2)); ok(hmalloc(&s->h_score, B));
  ok(hmalloc(&s->h_rng, B * 8)); ok(hmalloc(&s->h_pseed, B * 8));
  if (e != cudaSuccess) {
    search_free(s);
    return fail(ctx, OAKB200_ECUDA, "search_create: allocation", e);
  }
  s->particles.resize(B);
  *out = s;
  return OAKB200_OK;
}

void oakb200_search_destroy(oakb200_search* s) { search_free(s); }

int oakb200_search_timing(oakb200_search* s, double* out4) {
  if (!s || !out4) return OAKB200_EINVAL;
  out4[0] = s->last_total_ms;
  out4[1] = s->last_host_ms;
  out4[2] = s->last_idle_ms;
  out4[3] = (double)s->last_waves;
  return OAKB200_OK;
}

int oakb200_search_set_roots(oakb200_search* s, size_t n, const uint8_t* battles, const uint8_t* durations, const uint8_t* results,
                             const uint64_t* seeds) {
  if (!s) return OAKB200_EINVAL;
  oakb200_ctx* ctx = s->ctx;
  if (!battles || !durations || !results || !seeds || n == 0 || n > s->roots_max) return fail(ctx, OAKB200_EINVAL, "search_set_roots: bad argument");
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemcpyAsync(s->d_root_b, battles, n * 384, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(s->d_root_d, durations, n * 8, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(s->d_root_r, results, n, cudaMemcpyHostToDevice, ctx->stream));
  // the roots' legal choices (mcts.h:161-166) come from the same kernel every caller uses
  int rc = oakb200_choices_device(ctx, n, s->d_root_b, s->d_root_r, s->d_act, s->d_cnt);
  if (rc) return rc;
  std::vector<uint8_t> ch(n * 18), cnt(n * 2);
  CU(ctx, cudaMemcpyAsync(ch.data(), s->d_act, n * 18, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaMemcpyAsync(cnt.data(), s->d_cnt, n * 2, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  s->n_roots = n;
  if (s->roots.size() < n) s->roots.resize(n);  // the roots' arenas (nodes, table, paths) are kept from search to search
  for (size_t r = 0; r < n; r++) {
    if ((results[r] & 15u) != 0) return fail(ctx, OAKB200_EINVAL, "search_set_roots: a root battle is already over");
    Root& R = s->roots[r];
    R.seed = seeds[r];
    R.k1 = cnt[2 * r];
    R.k2 = cnt[2 * r + 1];
    memcpy(R.choices, ch.data() + 18 * r, 18);
    if (R.nodes.capacity() < 256) R.nodes.reserve(256);
    R.reset(s->bandit);
  }
  s->root_primed = false;
  return OAKB200_OK;
}

// Which bandit the trees use (cpp/src/search.cc:236-300: "ucb-<c>", "ucb1-<c>", "pucb-<c>", "exp3-<gamma>[-<alpha>]",
// "pexp3-<gamma>[-<alpha>]"): kind 0..4 in that order, f1 = c or gamma, f2 = alpha (Exp3 family; the reference's default is
// 0.05). The contextual ones (pucb, pexp3) take their priors from the network's policy head at the root and at every new
// leaf (mcts.h:190-209,409-421), so they need leaf mode 1. Resets the trees of roots already set.
int oakb200_search_set_bandit(oakb200_search* s, int kind, float f1, float f2) {
  if (!s) return OAKB200_EINVAL;
  oakb200_ctx* ctx = s->ctx;
  if (kind < B_UCB || kind > B_PEXP3 || !(f1 > 0) || (kind >= B_EXP3 && !(f2 >= 0 && f2 < 1)))
    return fail(ctx, OAKB200_EINVAL, "search_set_bandit: kind 0..4 (ucb, ucb1, pucb, exp3, pexp3), f1 > 0, 0 <= alpha < 1");
  s->bandit.kind = kind;
  s->bandit.c = s->bandit.gamma = f1;
  s->bandit.alpha = f2;
  s->bandit_set = true;
  s->root_primed = false;
  for (size_t r = 0; r < s->n_roots; r++) s->roots[r].reset(s->bandit);
  return OAKB200_OK;
}

int oakb200_search_run(oakb200_search* s, uint32_t iterations, float ucb_c, uint32_t root_rolls, uint32_t other_rolls, int host_threads) {
  if (!s) return OAKB200_EINVAL;
  oakb200_ctx* ctx = s->ctx;
  auto rolls_ok = [](uint32_t r) { return r == 1 || r == 2 || r == 3 || r == 20 || r == 39; };
  if (s->n_roots == 0 || !rolls_ok(root_rolls) || !rolls_ok(other_rolls))
    return fail(ctx, OAKB200_EINVAL, "search_run: set roots first; rolls must be 1, 2, 3, 20 or 39");
  CU(ctx, cudaSetDevice(ctx->device));
  if (host_threads < 1) {
    // one process per GPU: the ranks of a node share its cores (torchrun exports LOCAL_WORLD_SIZE); 8 ranks x every core
    // each would oversubscribe the host 8 times over
    host_threads = (int)std::thread::hardware_concurrency();
    if (const char* lw = getenv("LOCAL_WORLD_SIZE")) {
      const int ranks = atoi(lw);
      if (ranks > 1) host_threads /= ranks;
    }
    if (host_threads < 1) host_threads = 1;
  }
  if (!s->pool || s->pool_threads != host_threads) {
    delete s->pool;
    s->pool = new Pool(host_threads);
    s->pool_threads = host_threads;
  }
  const uint32_t P = s->P;
  const size_t R = s->n_roots;
  const uint32_t MAX_DEPTH = 100;  // mcts.h:309
  if (iterations == 0) return OAKB200_OK;
  const uint32_t waves = (iterations + P - 1) / P;
  if (!s->bandit_set) {
    s->bandit.kind = B_UCB;
    s->bandit.c = ucb_c;
  }
  const BanditParams bp = s->bandit;
  if (bp.contextual() && s->leaf_mode != 1)
    return fail(ctx, OAKB200_EINVAL, "search_run: pucb / pexp3 take their priors from the network: oakb200_search_set_leaf(search, 1) first");
  // exactly `iterations` more per root (the reference runs its budget to the iteration, mcts.h:237-241): the last wave
  // spawns only what is left
  for (size_t r = 0; r < R; r++) s->roots[r].target = s->roots[r].iterations + iterations;

  // contextual bandits: the root's priors from one batched evaluation of the root positions (mcts.h:190-209)
  if (bp.contextual() && !s->root_primed) {
    int rc = oakb200_eval_device(ctx, R, s->d_root_b, s->d_root_d, s->d_root_r, s->d_value, s->d_logits, s->d_ecnt);
    if (rc) return rc;
    CU(ctx, cudaMemcpyAsync(s->h_logits, s->d_logits, R * 18 * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t r = 0; r < R; r++) {
      s->roots[r].nodes[0].p1.set_logits(s->h_logits + 18 * r, bp, s->roots[r].pri(0, 0));
      s->roots[r].nodes[0].p2.set_logits(s->h_logits + 18 * r + 9, bp, s->roots[r].pri(0, 1));
    }
    s->root_primed = true;
  }

  // The roots are cut into groups, each with its own stream and its own position in the per-wave phases (spawn, then per
  // tree level select -> device step -> expand, then the leaves, then back-propagation). The scheduler below advances
  // whichever group's device work has finished: while the device runs one group's kernels the host threads select /
  // expand / back-propagate in another group's trees, and one group's level kernels run on the SMs another group's
  // playout launch frees in its tail. Trees of different roots never interact and every random stream is a function of
  // (root seed, iteration), so the grouping does not change any result.
  struct Group {
    size_t r0, r1;           // roots [r0, r1)
    cudaStream_t stream;
    uint32_t wave = 0, depth = 0;
    int phase = 0;           // 0 spawn, 1 select + step, 2 expand, 3 back-propagate, 4 finished
    bool reached_leaves = false;  // has launched a leaf phase at least once (the next group starts then: see below)
  };
  // Measured on B200, 4 096 roots x 64 (tools/search_timeline.py): groups that start together stay in lockstep — all of
  // them in the host-bound tree levels, then all of them waiting for their playouts — so group g + 1 starts its first wave
  // only when group g has launched its first leaves, and from then on one group's leaves run under another's tree levels.
  // 4 096 roots x 64, 16 waves on B200 + 16 host threads: 1 group 9.05 M iterations/s, 2: 8.7, 3: 9.9, 4: 10.8, 6: 9.6
  int n_groups = 4;
  if (const char* g = getenv("OAKB200_SEARCH_GROUPS")) n_groups = atoi(g);
  if (n_groups < 1) n_groups = 1;
  if (n_groups > 8) n_groups = 8;
  if ((size_t)n_groups > R) n_groups = (int)R;
  std::vector<Group> groups(n_groups);
  for (int g = 0; g < n_groups; g++) {
    groups[g].r0 = R * g / n_groups;
    groups[g].r1 = R * (g + 1) / n_groups;
    groups[g].stream = nullptr;
    if (n_groups == 1) groups[g].stream = ctx->stream;
    else CU(ctx, cudaStreamCreateWithFlags(&groups[g].stream, cudaStreamNonBlocking));
  }
  if (n_groups > 1) CU(ctx, cudaStreamSynchronize(ctx->stream));  // set_roots' uploads are on the context's stream
  cudaStream_t saved_stream = ctx->stream;
  const int saved_slot = ctx->ticket_slot;
  const int saved_cap = ctx->rollout_cta_cap;
  // a playout launch that takes every CTA slot of every SM would hold back the other groups' small level kernels for its
  // whole body: leave room for them
  if (n_groups > 1) {
    ctx->rollout_cta_cap = 6;
    if (const char* c = getenv("OAKB200_SEARCH_ROLLOUT_CTAS")) ctx->rollout_cta_cap = atoi(c);
  }
  int rc_out = OAKB200_OK;
  auto fail_cuda = [&](cudaError_t e, const char* what) {
    if (e != cudaSuccess && rc_out == OAKB200_OK) rc_out = fail(ctx, OAKB200_ECUDA, what, e);
    return e != cudaSuccess;
  };
  // the uniform draw of one Exp3-family selection: a function of (root seed, iteration, depth, player)
  auto exp3_u = [](uint64_t seed, uint64_t nth, uint32_t depth, int player) {
    const uint64_t x = splitmix64(seed ^ (STREAM_K * (2 * nth + 1)) ^ (0x9FB21C651E98DF25ull * (2 * (uint64_t)depth + player + 1)));
    return (double)(x >> 11) * (1.0 / 9007199254740992.0);
  };

  long long host_ns = 0, idle_ns = 0;  // diagnostics: OAKB200_SEARCH_DEBUG=1 prints them
  // one host phase of group G (its previous device work has finished) + the device work that follows it
  auto advance = [&](Group& G, int gi) {
    const size_t r0 = G.r0, nr = G.r1 - G.r0, b0 = r0 * P, nb = nr * P;
    ctx->stream = G.stream;
    ctx->ticket_slot = 8 + gi;
    if (fail_cuda(cudaStreamSynchronize(G.stream), "search_run: stream")) return;
    const auto t_host = std::chrono::steady_clock::now();
    struct HostTimer {
      std::chrono::steady_clock::time_point t0;
      long long& acc;
      ~HostTimer() { acc += (std::chrono::steady_clock::now() - t0).count(); }
    } host_timer{t_host, host_ns};
    if (G.phase == 0) {
      // ---- spawn: fresh rng per particle, hidden variables redrawn from the root's observed durations
      s->pool->parallel_for(nr, [&](size_t k) {
        const size_t r = r0 + k;
        Root& root = s->roots[r];
        const uint64_t left = root.target - root.iterations;
        for (uint32_t p = 0; p < P; p++) {
          const size_t b = r * P + p;
          const uint64_t nth = root.iterations + p;
          s->h_rng[b] = splitmix64(root.seed ^ (STREAM_K * (2 * nth + 1)));
          s->h_pseed[b] = splitmix64(root.seed ^ (STREAM_K * (2 * nth + 2)));
          Particle& q = s->particles[b];
          q.cur = 0;
          q.fresh = -1;
          q.len = 0;
          q.active = q.spawned = p < left;
        }
      });
      if (fail_cuda(cudaMemcpyAsync(s->d_rng + b0, s->h_rng + b0, nb * 8, cudaMemcpyHostToDevice, G.stream), "search_run: copy")) return;
      if (fail_cuda(cudaMemcpyAsync(s->d_pseed + b0, s->h_pseed + b0, nb * 8, cudaMemcpyHostToDevice, G.stream), "search_run: copy")) return;
      if (fail_cuda(oakb200_launch_search_spawn(ctx, s->d_root_b + r0 * 384, s->d_root_d + r0 * 8, s->d_root_r + r0, s->d_rng + b0, P,
                                                s->d_b + b0 * 384, s->d_d + b0 * 8, s->d_r + b0, (uint32_t)nb), "search_run: spawn")) return;
      const int rc = oakb200_randomize_hidden_device(ctx, nb, s->d_b + b0 * 384, s->d_d + b0 * 8);
      if (rc) { rc_out = rc; return; }
      G.depth = 0;
      G.phase = 1;
      // fall through to the first selection: it needs nothing from the device
    }
    if (G.phase == 2) {
      // ---- expand: match the observation to a child, create it if new (mcts.h:355-361)
      const uint32_t depth = G.depth;
      s->pool->parallel_for(nr, [&](size_t k) {
        const size_t r = r0 + k;
        Root& root = s->roots[r];
        const PathEntry* level = root.path.data() + (size_t)depth * P;
        // the trees do not fit any cache (4 096 roots x ~200 KB): the lookups of the particles a few places ahead are
        // started early — first the table slot of the key, then the node that slot names
        root.reserve_table();
        uint64_t hashes[256];
        const uint32_t PH = P <= 256 ? P : 256;
        for (uint32_t p = 0; p < PH; p++) {
          const PathEntry& pe = level[p];
          hashes[p] = s->particles[r * P + p].active ? Root::key_hash(pe.node, pe.i, pe.j, s->h_act + (r * P + p) * 16) : 0;
        }
        for (uint32_t p = 0; p < 8 && p < PH; p++) root.prefetch_slot(hashes[p]);
        for (uint32_t p = 0; p < 4 && p < PH; p++) root.prefetch_child(hashes[p]);
        for (uint32_t p = 0; p < P; p++) {
          const size_t b = r * P + p;
          Particle& q = s->particles[b];
          if (p + 8 < PH) root.prefetch_slot(hashes[p + 8]);
          if (p + 4 < PH) root.prefetch_child(hashes[p + 4]);
          if (!q.active) continue;
          const PathEntry& pe = level[p];
          const uint64_t h = p < PH ? hashes[p] : Root::key_hash(pe.node, pe.i, pe.j, s->h_act + b * 16);
          const int32_t child = root.child(pe.node, pe.i, pe.j, s->h_act + b * 16, h);
          q.cur = child;
          Node& cn = root.nodes[child];
          const bool over = (s->h_res[b] & 15u) != 0;
          if (cn.is_init() && !over) continue;  // keep descending (mcts.h:328)
          if (!over) {                           // new leaf: init its bandits, then evaluate it (mcts.h:399-424, :463-465)
            cn.p1.init(s->h_cnt[2 * b], bp, root.pri(child, 0));
            cn.p2.init(s->h_cnt[2 * b + 1], bp, root.pri(child, 1));
            q.fresh = child;
          }
          root.total_depth += depth + 1;
          q.active = false;
        }
      });
      G.depth++;
      G.phase = 1;
    }
    if (G.phase == 1) {
      // ---- select: one tree level per launch
      const uint32_t depth = G.depth;
      std::atomic<size_t> n_active{0};
      s->pool->parallel_for(nr, [&](size_t k) {
        const size_t r = r0 + k;
        Root& root = s->roots[r];
        if (root.path.size() < (size_t)(depth + 1) * P) root.path.resize((size_t)(depth + 1) * P);
        PathEntry* level = root.path.data() + (size_t)depth * P;
        size_t cnt = 0;
        for (uint32_t p = 0; p < 4 && p < P; p++)
          if (s->particles[r * P + p].active) root.nodes[s->particles[r * P + p].cur].prefetch();
        for (uint32_t p = 0; p < P; p++) {
          const size_t b = r * P + p;
          Particle& q = s->particles[b];
          if (p + 4 < P && s->particles[b + 4].active) root.nodes[s->particles[b + 4].cur].prefetch();
          if (!q.active || depth >= MAX_DEPTH) {
            q.active = false;
            s->h_idx1[b] = s->h_idx2[b] = 255;
            continue;
          }
          Node& node = root.nodes[q.cur];
          double u1 = 0, u2 = 0;
          if (!bp.counts_visits()) {
            const uint64_t nth = root.iterations + p;
            u1 = exp3_u(root.seed, nth, depth, 0);
            u2 = exp3_u(root.seed, nth, depth, 1);
          }
          PathEntry& pe = level[p];
          const int i = node.p1.select(bp, u1, &pe.prob1, root.pri(q.cur, 0)), j = node.p2.select(bp, u2, &pe.prob2, root.pri(q.cur, 1));
          if (bp.counts_visits()) {  // the visit is counted now (virtual visit), the value at back-propagation
            node.p1.visits[i]++;
            node.p2.visits[j]++;
          }
          pe.node = q.cur;
          pe.i = (uint8_t)i;
          pe.j = (uint8_t)j;
          q.len = (uint16_t)(depth + 1);
          s->h_idx1[b] = (uint8_t)i;
          s->h_idx2[b] = (uint8_t)j;
          s->h_rolls[b] = (uint8_t)(depth == 0 ? root_rolls : other_rolls);
          cnt++;
        }
        n_active += cnt;
      });
      if (n_active.load() != 0) {
        if (gi == 0) s->levels++;
        bool bad = fail_cuda(cudaMemcpyAsync(s->d_idx1 + b0, s->h_idx1 + b0, nb, cudaMemcpyHostToDevice, G.stream), "search_run: copy");
        bad = bad || fail_cuda(cudaMemcpyAsync(s->d_idx2 + b0, s->h_idx2 + b0, nb, cudaMemcpyHostToDevice, G.stream), "search_run: copy");
        bad = bad || fail_cuda(cudaMemcpyAsync(s->d_rolls + b0, s->h_rolls + b0, nb, cudaMemcpyHostToDevice, G.stream), "search_run: copy");
        bad = bad || fail_cuda(oakb200_launch_search_step(ctx, s->d_b + b0 * 384, s->d_d + b0 * 8, s->d_r + b0, s->d_idx1 + b0, s->d_idx2 + b0,
                                                          s->d_rolls + b0, s->d_act + b0 * 16, s->d_cnt + b0 * 2, (uint32_t)nb), "search_run: step");
        bad = bad || fail_cuda(cudaMemcpyAsync(s->h_res + b0, s->d_r + b0, nb, cudaMemcpyDeviceToHost, G.stream), "search_run: copy");
        bad = bad || fail_cuda(cudaMemcpyAsync(s->h_act + b0 * 16, s->d_act + b0 * 16, nb * 16, cudaMemcpyDeviceToHost, G.stream), "search_run: copy");
        bad = bad || fail_cuda(cudaMemcpyAsync(s->h_cnt + b0 * 2, s->d_cnt + b0 * 2, nb * 2, cudaMemcpyDeviceToHost, G.stream), "search_run: copy");
        if (bad) return;
        G.phase = 2;
        return;
      }
      // ---- every descent of the group has reached a leaf
      if (s->leaf_mode == 1) {
        // the network's value of every leaf (mcts.h:409-424) with the durations tracked down the path, and for the
        // contextual bandits its policy logits over the leaf's legal choices; a finished battle's score replaces the
        // value at back-propagation
        const int rc = oakb200_eval_device(ctx, nb, s->d_b + b0 * 384, s->d_d + b0 * 8, s->d_r + b0, s->d_value + b0, s->d_logits + b0 * 18,
                                           s->d_ecnt + b0 * 2);
        if (rc) { rc_out = rc; return; }
        if (fail_cuda(cudaMemcpyAsync(s->h_value + b0, s->d_value + b0, nb * 4, cudaMemcpyDeviceToHost, G.stream), "search_run: copy")) return;
        if (fail_cuda(cudaMemcpyAsync(s->h_res + b0, s->d_r + b0, nb, cudaMemcpyDeviceToHost, G.stream), "search_run: copy")) return;
        if (bp.contextual() &&
            fail_cuda(cudaMemcpyAsync(s->h_logits + b0 * 18, s->d_logits + b0 * 18, nb * 18 * 4, cudaMemcpyDeviceToHost, G.stream), "search_run: copy"))
          return;
        G.phase = 3;
        G.reached_leaves = true;
        return;
      }
      // playouts (a finished battle returns its score at once)
      const int rc = oakb200_rollout_device(ctx, nb, s->d_b + b0 * 384, s->d_d + b0 * 8, s->d_r + b0, s->d_pseed + b0, 100000, 0,
                                            s->d_score + b0, nullptr, nullptr);
      if (rc) { rc_out = rc; return; }
      if (fail_cuda(cudaMemcpyAsync(s->h_score + b0, s->d_score + b0, nb, cudaMemcpyDeviceToHost, G.stream), "search_run: copy")) return;
      G.phase = 3;
      G.reached_leaves = true;
      return;
    }
    if (G.phase == 3) {
      // ---- back-propagation (mcts.h:363-384), particle by particle in index order
      s->pool->parallel_for(nr, [&](size_t k) {
        const size_t r = r0 + k;
        Root& root = s->roots[r];
        auto prefetch_path = [&](uint32_t p) {  // the nodes particle p's value will be added to
          const Particle& q = s->particles[r * P + p];
          if (!q.spawned) return;
          for (int lv = (int)q.len - 1; lv >= 0; lv--) {
            const Node& n = root.nodes[root.path[(size_t)lv * P + p].node];
            __builtin_prefetch(&n, 1, 3);                                     // p1.a
            __builtin_prefetch(reinterpret_cast<const char*>(&n) + 64, 1, 3);  // p2.a starts in the second line
            __builtin_prefetch(reinterpret_cast<const char*>(&n) + 128, 1, 3);
          }
        };
        for (uint32_t p = 0; p < 3 && p < P; p++) prefetch_path(p);
        for (uint32_t p = 0; p < P; p++) {
          const size_t b = r * P + p;
          const Particle& q = s->particles[b];
          if (p + 3 < P) prefetch_path(p + 3);
          if (!q.spawned) continue;
          float v1;
          if (s->leaf_mode == 1) {
            const uint32_t t = s->h_res[b] & 15u;
            v1 = t == 0 ? s->h_value[b] : (t == 1 ? 1.0f : (t == 2 ? 0.0f : 0.5f));
            if (bp.contextual() && q.fresh >= 0 && t == 0) {  // stats.softmax_logits at the new leaf (mcts.h:415-421)
              root.nodes[q.fresh].p1.set_logits(s->h_logits + b * 18, bp, root.pri(q.fresh, 0));
              root.nodes[q.fresh].p2.set_logits(s->h_logits + b * 18 + 9, bp, root.pri(q.fresh, 1));
            }
          } else {
            v1 = (float)s->h_score[b] * 0.5f;
          }
          const float v2 = 1.0f - v1;
          for (int lv = (int)q.len - 1; lv >= 0; lv--) {  // deepest first, as the recursion unwinds
            const PathEntry& pe = root.path[(size_t)lv * P + p];
            Node& node = root.nodes[pe.node];
            node.p1.update(pe.i, v1, pe.prob1, bp);
            node.p2.update(pe.j, v2, pe.prob2, bp);
          }
          if (q.len) {
            const PathEntry& pe = root.path[p];
            root.visit[pe.i * 9 + pe.j]++;
            root.value[pe.i * 9 + pe.j] += v1;
          }
          root.iterations++;
        }
      });
      if (gi == 0) s->waves++;
      G.wave++;
      G.phase = G.wave < waves ? 0 : 4;
    }
  };

  // OAKB200_SEARCH_DEBUG=2: a timeline of (start ms, end ms, group, phase entered) for every host phase
  struct Trace { double t0, t1; int g, phase; };
  std::vector<Trace> trace;
  const bool tracing = getenv("OAKB200_SEARCH_DEBUG") && atoi(getenv("OAKB200_SEARCH_DEBUG")) >= 2;
  const auto t_begin = std::chrono::steady_clock::now();
  auto ms_now = [&] { return (std::chrono::steady_clock::now() - t_begin).count() * 1e-6; };
  // scheduler: advance any group whose stream has drained; when none has, wait for one
  int live = n_groups;
  while (live > 0 && rc_out == OAKB200_OK) {
    bool progressed = false;
    for (int g = 0; g < n_groups && rc_out == OAKB200_OK; g++) {
      Group& G = groups[g];
      if (G.phase == 4) continue;
      if (g > 0 && G.wave == 0 && G.phase == 0 && !groups[g - 1].reached_leaves && groups[g - 1].phase != 4) continue;  // stagger
      const cudaError_t q = cudaStreamQuery(G.stream);
      if (q == cudaErrorNotReady) continue;
      if (fail_cuda(q, "search_run: stream")) break;
      const double t0 = tracing ? ms_now() : 0;
      const int ph = G.phase;
      advance(G, g);
      if (tracing) trace.push_back({t0, ms_now(), g, ph});
      progressed = true;
      if (G.phase == 4) live--;
    }
    if (!progressed && live > 0) {
      const auto t0 = std::chrono::steady_clock::now();
      std::this_thread::yield();
      idle_ns += (std::chrono::steady_clock::now() - t0).count();
    }
  }
  ctx->stream = saved_stream;
  ctx->ticket_slot = saved_slot;
  ctx->rollout_cta_cap = saved_cap;
  s->last_host_ms = host_ns * 1e-6;
  s->last_idle_ms = idle_ns * 1e-6;
  s->last_total_ms = (std::chrono::steady_clock::now() - t_begin).count() * 1e-6;
  s->last_waves = waves;
  if (getenv("OAKB200_SEARCH_DEBUG"))
    fprintf(stderr, "[oakb200] search_run: %u waves x %d groups, host phases %.1f ms, idle (every group on the device) %.1f ms\n", waves,
            n_groups, host_ns * 1e-6, idle_ns * 1e-6);
  if (tracing) {
    const double total = ms_now();
    fprintf(stderr, "[oakb200] timeline of the last wave (ms from run start; total %.2f):\n", total);
    size_t from = trace.size() > (size_t)(14 * n_groups) ? trace.size() - 14 * n_groups : 0;
    for (size_t k = from; k < trace.size(); k++)
      fprintf(stderr, "  g%d phase %d  %8.2f -> %8.2f  (%.2f)\n", trace[k].g, trace[k].phase, trace[k].t0, trace[k].t1, trace[k].t1 - trace[k].t0);
  }
  if (n_groups > 1)
    for (auto& G : groups) {
      cudaStreamSynchronize(G.stream);
      cudaStreamDestroy(G.stream);
    }
  return rc_out;
}

int oakb200_search_output(oakb200_search* s, uint8_t* k, uint8_t* choices, uint32_t* visit_matrix, double* value_matrix,
                          uint64_t* iterations, double* empirical_value, double* empirical, uint64_t* nodes) {
  if (!s) return OAKB200_EINVAL;
  for (size_t r = 0; r < s->n_roots; r++) {
    const Root& R = s->roots[r];
    if (k) { k[2 * r] = R.k1; k[2 * r + 1] = R.k2; }
    if (choices) memcpy(choices + 18 * r, R.choices, 18);
    if (visit_matrix) memcpy(visit_matrix + 81 * r, R.visit, sizeof(R.visit));
    if (value_matrix) memcpy(value_matrix + 81 * r, R.value, sizeof(R.value));
    if (iterations) iterations[r] = R.iterations;
    if (nodes) nodes[r] = R.nodes.size();
    // process_output (mcts.h:623-661) without the Nash solve, which stays with the caller (LRSNash)
    double total = 0, e1[9] = {0}, e2[9] = {0};
    for (int i = 0; i < R.k1; i++)
      for (int j = 0; j < R.k2; j++) {
        total += R.value[i * 9 + j];
        e1[i] += R.visit[i * 9 + j];
        e2[j] += R.visit[i * 9 + j];
      }
    const double it = R.iterations ? (double)R.iterations : 1.0;
    if (empirical_value) empirical_value[r] = total / it;
    if (empirical)
      for (int i = 0; i < 9; i++) {
        empirical[18 * r + i] = e1[i] / (float)it;
        empirical[18 * r + 9 + i] = e2[i] / (float)it;
      }
  }
  return OAKB200_OK;
}

// Matrix-game solve (the job of LRSNash::solve_fast, cpp/include/search/mcts.h:643-649, cpp/src/pyoak.cc:394-425).
// LRSNash (github.com/lab-oak/lrsnash, an empty submodule in the reference tree) walks the vertices of the game's
// polytopes in exact rational arithmetic; any equilibrium of the same game has the same value, the strategies are not
// unique. Here: the classical LP — shift the payoffs positive, maximise sum(y) subject to A y <= 1 — by a dense-tableau
// simplex in doubles with Bland's rule (no cycling); the row player's strategy is read from the duals. Data are
// integers <= denom on at most 9 x 9, so doubles are far from their limits.
int oakb200_solve_matrix(int rows, int cols, const int* payoff, int denom, float* p1, float* p2, float* value) {
  if (rows < 1 || cols < 1 || rows > 9 || cols > 9 || !payoff || denom <= 0) return OAKB200_EINVAL;
  const int m = rows, n = cols, W = n + m + 1;
  double T[10][20];  // m constraint rows + the objective row; columns: y (n), slack (m), rhs
  int lo = payoff[0];
  for (int i = 0; i < m * n; i++) lo = payoff[i] < lo ? payoff[i] : lo;
  const double shift = 1.0 - lo;  // every entry >= 1
  for (int i = 0; i < m; i++) {
    for (int j = 0; j < n; j++) T[i][j] = payoff[i * n + j] + shift;
    for (int j = 0; j < m; j++) T[i][n + j] = i == j;
    T[i][W - 1] = 1.0;
  }
  for (int j = 0; j < W; j++) T[m][j] = j < n ? -1.0 : 0.0;
  int basis[9];
  for (int i = 0; i < m; i++) basis[i] = n + i;
  for (int iter = 0; iter < 10000; iter++) {
    int enter = -1;
    for (int j = 0; j < n + m; j++)
      if (T[m][j] < -1e-12) { enter = j; break; }
    if (enter < 0) break;
    int leave = -1;
    double best = 0;
    for (int i = 0; i < m; i++) {
      if (T[i][enter] <= 1e-12) continue;
      const double ratio = T[i][W - 1] / T[i][enter];
      if (leave < 0 || ratio < best - 1e-15 || (ratio <= best + 1e-15 && basis[i] < basis[leave])) { leave = i; best = ratio; }
    }
    if (leave < 0) return OAKB200_EINVAL;  // unbounded: impossible for positive A
    const double piv = T[leave][enter];
    for (int j = 0; j < W; j++) T[leave][j] /= piv;
    for (int i = 0; i <= m; i++) {
      if (i == leave) continue;
      const double f = T[i][enter];
      if (f == 0) continue;
      for (int j = 0; j < W; j++) T[i][j] -= f * T[leave][j];
    }
    basis[leave] = enter;
  }
  const double total = T[m][W - 1];  // sum(y) = 1 / value of the shifted game
  if (!(total > 0)) return OAKB200_EINVAL;
  double y[9] = {0}, x[9];
  for (int i = 0; i < m; i++)
    if (basis[i] < n) y[basis[i]] = T[i][W - 1];
  for (int i = 0; i < m; i++) x[i] = T[m][n + i];  // duals
  double sx = 0, sy = 0;
  for (int i = 0; i < m; i++) { x[i] = x[i] < 0 ? 0 : x[i]; sx += x[i]; }
  for (int j = 0; j < n; j++) { y[j] = y[j] < 0 ? 0 : y[j]; sy += y[j]; }
  if (p1) for (int i = 0; i < m; i++) p1[i] = (float)(x[i] / sx);
  if (p2) for (int j = 0; j < n; j++) p2[j] = (float)(y[j] / sy);
  if (value) *value = (float)((1.0 / total - shift) / denom);
  return OAKB200_OK;
}

// process_output's Nash half (mcts.h:628-656): average values discretised to 1/256, solved, per root
int oakb200_search_nash(oakb200_search* s, float* nash, float* nash_value) {
  if (!s) return OAKB200_EINVAL;
  for (size_t r = 0; r < s->n_roots; r++) {
    const Root& R = s->roots[r];
    int M[81];
    for (int i = 0; i < R.k1; i++)
      for (int j = 0; j < R.k2; j++) {
        uint32_t n = R.visit[i * 9 + j];
        n += !n;
        M[R.k2 * i + j] = (int)(R.value[i * 9 + j] / n * 256);
      }
    float a[9] = {0}, b[9] = {0}, v = 0.5f;
    if (R.k1 >= 1 && R.k2 >= 1 && oakb200_solve_matrix(R.k1, R.k2, M, 256, a, b, &v) != OAKB200_OK) return OAKB200_EINVAL;
    if (nash) {
      for (int i = 0; i < 9; i++) {
        nash[18 * r + i] = i < R.k1 ? a[i] : 0.0f;
        nash[18 * r + 9 + i] = i < R.k2 ? b[i] : 0.0f;
      }
    }
    if (nash_value) nash_value[r] = v;
  }
  return OAKB200_OK;
}

int oakb200_search_set_leaf(oakb200_search* s, int mode) {
  if (!s) return OAKB200_EINVAL;
  oakb200_ctx* ctx = s->ctx;
  if (mode != 0 && mode != 1) return fail(ctx, OAKB200_EINVAL, "search_set_leaf: 0 = MonteCarlo playouts, 1 = network value");
  if (mode == 1) {
    if (!ctx->net) return fail(ctx, OAKB200_EINVAL, "search_set_leaf: no network loaded (oakb200_net_load)");
    if (!s->d_value) {
      const size_t B = s->roots_max * (size_t)s->P;
      CU(ctx, cudaSetDevice(ctx->device));
      CU(ctx, cudaMalloc(&s->d_value, B * 4));
      CU(ctx, cudaMalloc(&s->d_logits, B * 18 * 4));
      CU(ctx, cudaMalloc(&s->d_ecnt, B * 2));
      CU(ctx, cudaMallocHost(&s->h_value, B * 4));
      CU(ctx, cudaMallocHost(&s->h_logits, B * 18 * 4));
    }
  }
  s->leaf_mode = mode;
  return OAKB200_OK;
}

int oakb200_search_stats(oakb200_search* s, uint64_t* waves, uint64_t* levels) {
  if (!s) return OAKB200_EINVAL;
  if (waves) *waves = s->waves;
  if (levels) *levels = s->levels;
  return OAKB200_OK;
}

}  // extern "C"
