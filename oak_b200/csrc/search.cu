// Batched-leaf tree search: host trees, GPU transitions and leaves (BASELINE configs[3]).
//
// Reference algorithm: MCTS::Search::{run, run_root_iteration, run_iteration, init_stats_and_rollout}
// (cpp/include/search/mcts.h:154-496) with the joint UCB bandit (cpp/include/search/bandit/ucb.h:15-71,
// cpp/include/search/joint.h) on a Node heap keyed by (p1 index, p2 index, 16-byte chance observation)
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
// splitmix64(s ^ K * (2n + 2)), K = 0xD1B54A32D192ED03; the playout then picks as oakb200_rollout does.
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

// UCB::Bandit (ucb.h:15-66)
struct Bandit {
  float scores[9];
  uint32_t visits[9];
  uint8_t k = 0;
  void init(int n) {
    k = (uint8_t)n;
    for (int i = 0; i < n; i++) {
      scores[i] = 0.5f;
      visits[i] = 1;
    }
  }
  int select(float c) const {
    if (k == 1) return 0;
    uint64_t N = 0;
    for (int i = 0; i < k; i++) N += visits[i];
    const float sqrtN = (float)std::sqrt((double)N);
    float max = 0;
    int index = 0;
    for (int i = 0; i < k; i++) {
      const float e = c * sqrtN / k;
      const float a = (e + scores[i]) / visits[i];
      if (a > max) {
        max = a;
        index = i;
      }
    }
    return index;
  }
};

struct Child {
  uint8_t i, j;
  uint8_t obs[16];
  int32_t node;
};
struct Node {
  Bandit p1, p2;
  std::vector<Child> children;
  bool is_init() const { return p1.k != 0; }
};
struct PathEntry {
  int32_t node;
  uint8_t i, j;
};

struct Root {
  std::vector<Node> nodes;  // nodes[0] = root
  uint8_t k1 = 0, k2 = 0;
  uint8_t choices[18] = {0};
  uint32_t visit[81] = {0};
  double value[81] = {0};
  uint64_t iterations = 0;
  uint64_t seed = 0;
  uint64_t total_depth = 0;
};

struct Particle {
  int32_t cur = 0;
  bool active = false;
  std::vector<PathEntry> path;
};

// persistent worker pool: parallel_for over roots
class Pool {
 public:
  explicit Pool(int threads) {
    if (threads < 1) threads = 1;
    for (int t = 1; t < threads; t++) workers_.emplace_back([this] { loop(); });
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
      next_.store(0);
      pending_ = (int)workers_.size();
      gen_++;
    }
    cv_.notify_all();
    work();
    std::unique_lock<std::mutex> l(mu_);
    done_.wait(l, [this] { return pending_ == 0; });
  }

 private:
  void work() {
    for (;;) {
      const size_t i = next_.fetch_add(1);
      if (i >= n_) break;
      (*fn_)(i);
    }
  }
  void loop() {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [&] { return gen_ != seen; });
        seen = gen_;
        if (stop_) return;
      }
      work();
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
  std::atomic<size_t> next_{0};
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
  float *d_value = nullptr, *d_logits = nullptr, *h_value = nullptr;
  uint8_t* d_ecnt = nullptr;
  // pinned host mirrors
  uint8_t *h_idx1 = nullptr, *h_idx2 = nullptr, *h_rolls = nullptr, *h_res = nullptr, *h_act = nullptr, *h_cnt = nullptr, *h_score = nullptr;
  uint64_t *h_rng = nullptr, *h_pseed = nullptr;
  Pool* pool = nullptr;
  int pool_threads = 0;
  uint64_t levels = 0, waves = 0;
};

namespace {

void search_free(oakb200_search* s) {
  if (!s) return;
  if (s->ctx) cudaSetDevice(s->ctx->device);
  void* dev[] = {s->d_root_b, s->d_root_d, s->d_root_r, s->d_b, s->d_d, s->d_r, s->d_idx1, s->d_idx2, s->d_rolls, s->d_act, s->d_cnt,
                 s->d_score, s->d_rng, s->d_pseed, s->d_value, s->d_logits, s->d_ecnt};
  for (void* p : dev)
    if (p) cudaFree(p);
  void* host[] = {s->h_idx1, s->h_idx2, s->h_rolls, s->h_res, s->h_act, s->h_cnt, s->h_score, s->h_rng, s->h_pseed, s->h_value};
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
  s->roots.assign(n, Root());
  for (size_t r = 0; r < n; r++) {
    if ((results[r] & 15u) != 0) return fail(ctx, OAKB200_EINVAL, "search_set_roots: a root battle is already over");
    Root& R = s->roots[r];
    R.seed = seeds[r];
    R.k1 = cnt[2 * r];
    R.k2 = cnt[2 * r + 1];
    memcpy(R.choices, ch.data() + 18 * r, 18);
    R.nodes.reserve(1024);
    R.nodes.emplace_back();
    R.nodes[0].p1.init(R.k1);
    R.nodes[0].p2.init(R.k2);
  }
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
  const uint32_t waves = (iterations + P - 1) / P;
  if (waves == 0) return OAKB200_OK;

  // The roots are cut into groups that take turns: while the device runs one group's kernels (a tree level, or the
  // playouts, which are most of a wave's device time) the host threads select / expand / back-propagate in another
  // group's trees. Each group has its own stream and walks the same per-wave phases; a group's next host phase starts
  // by waiting for that group's stream only. Trees of different roots never interact and every random stream is a
  // function of (root seed, iteration), so the grouping does not change any result.
  struct Group {
    size_t r0, r1;           // roots [r0, r1)
    cudaStream_t stream;
    uint32_t wave = 0, depth = 0;
    int phase = 0;           // 0 spawn, 1 select + step, 2 expand, 3 back-propagate, 4 finished
  };
  // Measured on B200, 4 096 roots x 64 (tools/bench_search.py): 1 group 8.14 M iterations/s, 2 groups 8.82 M, 3 groups 7.8 M,
  // 4 groups 7.1 M. The gain is small because a wave's device time is almost all playouts, a persistent kernel that
  // holds every SM until it ends: another group's tree levels cannot run beside it, only its host phases can.
  int n_groups = 2;
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
  int rc_out = OAKB200_OK;
  auto fail_cuda = [&](cudaError_t e, const char* what) {
    if (e != cudaSuccess && rc_out == OAKB200_OK) rc_out = fail(ctx, OAKB200_ECUDA, what, e);
    return e != cudaSuccess;
  };

  long long wait_ns = 0, host_ns = 0;  // diagnostics: OAKB200_SEARCH_DEBUG=1 prints them
  // one host phase of group G (after its previous device work has finished) + the device work that follows it
  auto advance = [&](Group& G, int gi) {
    const size_t r0 = G.r0, nr = G.r1 - G.r0, b0 = r0 * P, nb = nr * P;
    ctx->stream = G.stream;
    ctx->ticket_slot = 8 + gi;
    const auto t_wait = std::chrono::steady_clock::now();
    if (fail_cuda(cudaStreamSynchronize(G.stream), "search_run: stream")) return;
    const auto t_host = std::chrono::steady_clock::now();
    wait_ns += (t_host - t_wait).count();
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
        for (uint32_t p = 0; p < P; p++) {
          const size_t b = r * P + p;
          const uint64_t nth = root.iterations + p;
          s->h_rng[b] = splitmix64(root.seed ^ (STREAM_K * (2 * nth + 1)));
          s->h_pseed[b] = splitmix64(root.seed ^ (STREAM_K * (2 * nth + 2)));
          Particle& q = s->particles[b];
          q.cur = 0;
          q.active = true;
          q.path.clear();
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
        for (uint32_t p = 0; p < P; p++) {
          const size_t b = r * P + p;
          Particle& q = s->particles[b];
          if (!q.active) continue;
          const PathEntry& pe = q.path.back();
          const uint8_t* obs = s->h_act + b * 16;
          int32_t child = -1;
          for (const Child& c : root.nodes[pe.node].children)
            if (c.i == pe.i && c.j == pe.j && memcmp(c.obs, obs, 16) == 0) {
              child = c.node;
              break;
            }
          if (child < 0) {
            child = (int32_t)root.nodes.size();
            root.nodes.emplace_back();
            Child c;
            c.i = pe.i;
            c.j = pe.j;
            memcpy(c.obs, obs, 16);
            c.node = child;
            root.nodes[pe.node].children.push_back(c);
          }
          q.cur = child;
          Node& cn = root.nodes[child];
          const bool over = (s->h_res[b] & 15u) != 0;
          if (cn.is_init() && !over) continue;  // keep descending (mcts.h:328)
          if (!over) {                           // new leaf: init its bandits, then a playout (mcts.h:399, :463-465)
            cn.p1.init(s->h_cnt[2 * b]);
            cn.p2.init(s->h_cnt[2 * b + 1]);
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
        size_t cnt = 0;
        for (uint32_t p = 0; p < P; p++) {
          const size_t b = r * P + p;
          Particle& q = s->particles[b];
          if (!q.active || depth >= MAX_DEPTH) {
            q.active = false;
            s->h_idx1[b] = s->h_idx2[b] = 255;
            continue;
          }
          Node& node = root.nodes[q.cur];
          const int i = node.p1.select(ucb_c), j = node.p2.select(ucb_c);
          node.p1.visits[i]++;  // the visit is counted now (virtual visit), the value at back-propagation
          node.p2.visits[j]++;
          q.path.push_back({q.cur, (uint8_t)i, (uint8_t)j});
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
        // the network's value of every leaf (mcts.h:409-424: UCB is not a contextual bandit, so value_inference) with
        // the durations tracked down the path; a finished battle's score replaces it at back-propagation
        const int rc = oakb200_eval_device(ctx, nb, s->d_b + b0 * 384, s->d_d + b0 * 8, s->d_r + b0, s->d_value + b0, s->d_logits + b0 * 18,
                                           s->d_ecnt + b0 * 2);
        if (rc) { rc_out = rc; return; }
        if (fail_cuda(cudaMemcpyAsync(s->h_value + b0, s->d_value + b0, nb * 4, cudaMemcpyDeviceToHost, G.stream), "search_run: copy")) return;
        if (fail_cuda(cudaMemcpyAsync(s->h_res + b0, s->d_r + b0, nb, cudaMemcpyDeviceToHost, G.stream), "search_run: copy")) return;
        G.phase = 3;
        return;
      }
      // playouts (a finished battle returns its score at once)
      const int rc = oakb200_rollout_device(ctx, nb, s->d_b + b0 * 384, s->d_d + b0 * 8, s->d_r + b0, s->d_pseed + b0, 100000, 0,
                                            s->d_score + b0, nullptr, nullptr);
      if (rc) { rc_out = rc; return; }
      if (fail_cuda(cudaMemcpyAsync(s->h_score + b0, s->d_score + b0, nb, cudaMemcpyDeviceToHost, G.stream), "search_run: copy")) return;
      G.phase = 3;
      return;
    }
    if (G.phase == 3) {
      // ---- back-propagation (mcts.h:363-384)
      s->pool->parallel_for(nr, [&](size_t k) {
        const size_t r = r0 + k;
        Root& root = s->roots[r];
        for (uint32_t p = 0; p < P; p++) {
          const size_t b = r * P + p;
          float v1;
          if (s->leaf_mode == 1) {
            const uint32_t t = s->h_res[b] & 15u;
            v1 = t == 0 ? s->h_value[b] : (t == 1 ? 1.0f : (t == 2 ? 0.0f : 0.5f));
          } else {
            v1 = (float)s->h_score[b] * 0.5f;
          }
          const float v2 = 1.0f - v1;
          const Particle& q = s->particles[b];
          for (const PathEntry& pe : q.path) {
            Node& node = root.nodes[pe.node];
            node.p1.scores[pe.i] += v1;
            node.p2.scores[pe.j] += v2;
          }
          if (!q.path.empty()) {
            root.visit[q.path[0].i * 9 + q.path[0].j]++;
            root.value[q.path[0].i * 9 + q.path[0].j] += v1;
          }
          root.iterations++;
        }
      });
      if (gi == 0) s->waves++;
      G.wave++;
      G.phase = G.wave < waves ? 0 : 4;
    }
  };

  for (;;) {
    bool any = false;
    for (int g = 0; g < n_groups && rc_out == OAKB200_OK; g++) {
      if (groups[g].phase == 4) continue;
      any = true;
      advance(groups[g], g);
    }
    if (!any || rc_out != OAKB200_OK) break;
  }
  ctx->stream = saved_stream;
  ctx->ticket_slot = saved_slot;
  if (getenv("OAKB200_SEARCH_DEBUG"))
    fprintf(stderr, "[oakb200] search_run: %u waves x %d groups, host phases %.1f ms, waiting for the device %.1f ms\n", waves, n_groups,
            host_ns * 1e-6, wait_ns * 1e-6);
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
