"""Self-play data generation on the GPU: many concurrent games per device, network policy, frames to .battle.data.

Mirrors the battle loop of the reference's `generate` program (cpp/src/generate.cc:218-359) for the part that is
data-parallel: build the battle and send out the leads, then per turn {legal choices, network value + policy,
RuntimePolicy temperature / floor / sample (cpp/include/util/policy.h), record the update, pkmn_gen1_battle_update},
and at game end serialise the game (cpp/include/train/battle/compressed-frame.h) into files flushed every
`buffer_mb` MiB (generate.cc:169-188,351-353). Every step of that is a kernel of liboakb200; this module only
sequences launches, refills finished slots and writes files. The search that the reference runs per turn
(RuntimeSearch::run) is the host control plane and is not part of this loop: the recorded "empirical" and "nash"
policies are the network's own softmax and both values the network's value, iterations 0 (a budget-0 agent).

Multi-GPU: one process per GPU, games sharded by rank, no collective on the data path; network weights are
broadcast from rank 0 and finished frames are gathered to rank 0's writer (oak_b200.sharding).
"""
import os

import numpy as np

from .engine import BattleEngine, OakB200Error

MASK64 = (1 << 64) - 1


def _splitmix64(x):
    z = (x + 0x9E3779B97F4A7C15) & MASK64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
    return z ^ (z >> 31)


def _splitmix64_np(x):
    """_splitmix64 on a uint64 array (wrapping arithmetic)."""
    with np.errstate(over="ignore"):
        z = x + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


class FrameWriter:
    """N.battle.data files of at least `buffer_mb` MiB each (generate.cc:169-188), plain concatenations of game records."""

    def __init__(self, out_dir, buffer_mb=8, first_index=0):
        self.out_dir, self.limit, self.index = out_dir, int(buffer_mb) << 20, first_index
        self.chunks, self.size, self.games, self.files = [], 0, 0, []
        if out_dir:
            os.makedirs(out_dir, exist_ok=True)

    def add(self, stream, games):
        if len(stream) == 0:
            return
        self.chunks.append(bytes(stream))
        self.size += len(stream)
        self.games += int(games)
        if self.size >= self.limit:
            self.flush()

    def flush(self):
        if self.size == 0 or not self.out_dir:
            return
        path = os.path.join(self.out_dir, f"{self.index}.battle.data")
        with open(path, "wb") as f:
            for c in self.chunks:
                f.write(c)
        self.files.append(path)
        self.index += 1
        self.chunks, self.size = [], 0


class SelfPlay:
    def __init__(self, games, net_blob, team_pool, seed=0, device=0, max_updates=1024, temperature=1.0, min_prob=0.0,
                 first_game=0, eval_mode=0):
        """team_pool[P,6,5] u8 = {species, move x4} sets to draw both sides from; game g of this shard has global id
        first_game + g-th started game and battle seed splitmix64(seed ^ id)."""
        import torch
        self.torch = torch
        self.n = int(games)
        self.dev = torch.device("cuda", device)
        self.eng = BattleEngine(self.n, device=device)
        self.eng.set_stream(torch.cuda.current_stream(self.dev).cuda_stream)
        if net_blob is not None:
            self.eng.load_network(net_blob)
            self.eng.set_eval_mode(eval_mode)
        self.eng.frames_create(self.n, max_updates)
        self.pool = np.ascontiguousarray(team_pool, np.uint8)
        self.seed, self.temperature, self.min_prob = int(seed), float(temperature), float(min_prob)
        self.next_game = int(first_game)
        self.rng = np.random.default_rng(self.seed ^ 0x5EED)
        n, dev = self.n, self.dev
        u8 = torch.uint8
        self.battles = torch.zeros((n, 384), dtype=u8, device=dev)
        self.durations = torch.zeros((n, 8), dtype=u8, device=dev)
        self.results = torch.full((n,), 1, dtype=u8, device=dev)  # "ended": the first respawn fills every slot
        self.choices = torch.zeros((n, 2, 9), dtype=u8, device=dev)
        self.counts = torch.zeros((n, 2), dtype=u8, device=dev)
        self.value = torch.zeros(n, dtype=torch.float32, device=dev)
        self.logits = torch.zeros((n, 2, 9), dtype=torch.float32, device=dev)
        self.ecounts = torch.zeros((n, 2), dtype=u8, device=dev)
        self.probs = torch.zeros((n, 2, 9), dtype=torch.float32, device=dev)
        self.c1 = torch.zeros(n, dtype=u8, device=dev)
        self.c2 = torch.zeros(n, dtype=u8, device=dev)
        self.before = torch.zeros(n, dtype=u8, device=dev)
        self.teams = torch.zeros((n, 2, 6, 5), dtype=u8, device=dev)
        self.seeds = torch.zeros(n, dtype=torch.int64, device=dev)
        self.step_index = 0
        self.updates = 0
        self.started = 0
        self._refill(np.ones(n, bool))

    def close(self):
        self.eng.close()

    def _refill(self, ended):
        """New battles into the slots whose game has ended (their frames are already finished or drained)."""
        idx = np.nonzero(ended)[0]
        if idx.size == 0:
            return
        torch = self.torch
        ids = self.next_game + np.arange(idx.size)
        self.next_game += idx.size
        self.started += idx.size
        pick = self.rng.integers(0, self.pool.shape[0], (idx.size, 2))
        teams = self.pool[pick]  # [k,2,6,5]
        seeds = _splitmix64_np(np.uint64(self.seed) ^ ids.astype(np.uint64))
        di = torch.from_numpy(idx).to(self.dev)
        self.teams[di] = torch.from_numpy(teams).to(self.dev)
        self.seeds[di] = torch.from_numpy(seeds.view(np.int64)).to(self.dev)
        self.eng.respawn_device(self.battles, self.durations, self.results, self.teams, self.seeds)
        self.eng.frames_begin_device(self.battles, self.results)

    def step(self):
        """One lockstep turn of every running game: one library call (oakb200_selfplay_step_device = choices, evaluation,
        policy sample, frame record, update, finish)."""
        self.eng.selfplay_step_device(self.battles, self.durations, self.results, self.seeds, self.step_index, self.temperature,
                                      self.min_prob, self.choices, self.counts, self.value, self.logits, self.ecounts, self.probs,
                                      self.c1, self.c2, self.before)
        self.step_index += 1

    def collect(self, device=False):
        """Drain finished games -> (stream, games) and restart their slots. device=True leaves the stream in device memory
        (a torch uint8 tensor, valid until the next collect) for sharding.gather_streams."""
        if device:
            buf = getattr(self, "_drain_dev", None)
            if buf is None:
                buf = self._drain_dev = self.torch.empty(64 << 20, dtype=self.torch.uint8, device=self.dev)
            try:
                stream, games = self.eng.frames_drain(out=buf)
            except OakB200Error:
                # too small (a failed drain consumes nothing): grow and retry
                buf = self._drain_dev = self.torch.empty(4 * buf.numel(), dtype=self.torch.uint8, device=self.dev)
                stream, games = self.eng.frames_drain(out=buf)
        else:
            stream, games = self.eng.frames_drain()
        ended = (self.results.cpu().numpy() & 15) != 0
        self.updates += int((~ended).sum())
        self._refill(ended)
        return stream, games

    def run(self, steps, writer=None, collect_every=8):
        """-> total games finished. Frames go to `writer` (FrameWriter) if given."""
        done = 0
        for s in range(steps):
            self.step()
            if (s + 1) % collect_every == 0 or s + 1 == steps:
                stream, games = self.collect()
                done += games
                if writer is not None:
                    writer.add(stream, games)
        return done


class PipelinedSelfPlay:
    """`lanes` SelfPlay generators (default two) on their own CUDA streams, software-pipelined: while the host drains one
    lane (the drain's size read-back, the ended-slot scan, new teams and seeds for the refill — all of which wait for that
    lane's kernels and then keep the host busy) the other lanes' next steps are already queued on the device, so neither the
    GPU nor the host waits for the other. Games are independent, so this is the same generation as one SelfPlay of
    lanes * games_per_lane games; lane l plays game ids first_game + l * 2^28 + ..."""

    def __init__(self, games_per_lane, net_blob, team_pool, seed=0, device=0, lanes=2, first_game=0, **kw):
        import torch
        self.torch = torch
        self.dev = torch.device("cuda", device)
        self.streams = [torch.cuda.Stream(self.dev) for _ in range(lanes)]
        self.lanes = []
        for l, st in enumerate(self.streams):
            st.wait_stream(torch.cuda.current_stream(self.dev))
            with torch.cuda.stream(st):
                self.lanes.append(SelfPlay(games_per_lane, net_blob, team_pool, seed=seed + 7919 * l, device=device,
                                           first_game=first_game + (l << 28), **kw))
        self.queued = [0] * lanes  # steps queued on the lane since its last collect

    @property
    def launch_count(self):
        return sum(sp.eng.launch_count for sp in self.lanes)

    def run(self, steps, on_drain, collect_every=8):
        """`steps` lockstep turns of EVERY lane; after each `collect_every` steps of a lane its finished games are drained and
        on_drain(stream, games) is called inside that lane's stream context with the device stream (valid until the lane's
        next drain). -> games finished."""
        torch, done = self.torch, 0
        assert steps % collect_every == 0
        for _ in range(steps // collect_every):
            for l, sp in enumerate(self.lanes):
                with torch.cuda.stream(self.streams[l]):
                    # the lane that was queued one block ago is drained only now: its kernels ran under the other lanes' host work
                    if self.queued[l]:
                        done += self._collect(l, on_drain)
                    for _ in range(collect_every):
                        sp.step()
                    self.queued[l] = collect_every
        for l in range(len(self.lanes)):
            with torch.cuda.stream(self.streams[l]):
                if self.queued[l]:
                    done += self._collect(l, on_drain)
        return done

    def _collect(self, l, on_drain):
        stream, games = self.lanes[l].collect(device=True)
        self.queued[l] = 0
        on_drain(stream, games)
        return games

    def close(self):
        for sp in self.lanes:
            sp.close()


def runtime_policy(empirical, nash, mode="e", temp=1.0, min_prob=0.0):
    """RuntimePolicy::get_policy (reference cpp/include/util/policy.h:22-98) for rows of 9 probabilities: `mode` is a
    '-'-joined list of <letter>[weight] with e = empirical, n = nash, x = argmax of the empirical policy; then temperature,
    floor and renormalisation in double. (p = the network prior and b = beta are not produced by this search.)"""
    emp, nsh = np.asarray(empirical, np.float64), np.asarray(nash, np.float64)
    pol = np.zeros_like(emp)
    for word in mode.split("-"):
        w = float(word[1:]) if len(word) > 1 else 1.0
        if word[0] == "e":
            pol += w * emp
        elif word[0] == "n":
            pol += w * nsh
        elif word[0] == "x":
            pol[np.arange(pol.shape[0]), emp.argmax(-1)] += w
        else:
            raise ValueError(f"RuntimePolicy: invalid mode char: {word[0]}")
    if temp != 1:
        pol = np.power(pol, temp)
        pol /= pol.sum(-1, keepdims=True)
    pol[pol < min_prob] = 0
    s = pol.sum(-1, keepdims=True)
    if (s == 0).any():
        raise RuntimeError(f"RuntimePolicy: zero policy, mode: {mode}")
    return pol / s


class SearchSelfPlay(SelfPlay):
    """The generate loop with its search (reference cpp/src/generate.cc:286-333): every turn of every running game gets
    a tree search (oakb200_search_*: `iterations` MonteCarlo iterations, joint UCB, `particles` descents in flight per
    root), the move is drawn from RuntimePolicy::get_policy of the search output, and the recorded update carries the
    search's iteration count, empirical and Nash policies and values (CompressedFrames::Update(output, c1, c2),
    compressed-frame.h:47-56). Leaves are MonteCarlo playouts without a network (`net_blob` None) and the network's value
    with one."""

    def __init__(self, games, team_pool, iterations, particles=16, policy_mode="e", ucb_c=1.0, rolls=(3, 1), net_blob=None,
                 net_leaves=None, **kw):
        from .search import TreeSearch
        self._net = net_blob is not None
        super().__init__(games, net_blob, team_pool, **kw)
        self.iterations, self.policy_mode, self.ucb_c, self.rolls = int(iterations), policy_mode, float(ucb_c), rolls
        self.search = TreeSearch(self.eng, self.n, particles=particles)
        if net_leaves if net_leaves is not None else self._net:  # a network, when given, evaluates the leaves (mcts.h:409-424)
            self.search.set_leaf(1)
        torch = self.torch
        self.iters_t = torch.zeros(self.n, dtype=torch.int32, device=self.dev)
        self.nash_value = torch.zeros(self.n, dtype=torch.float32, device=self.dev)
        self.nash_probs = torch.zeros((self.n, 2, 9), dtype=torch.float32, device=self.dev)
        self.searched = 0

    def close(self):
        self.search.close()
        super().close()

    def step(self):
        torch, e, n = self.torch, self.eng, self.n
        e.choices_device(self.battles, self.results, self.choices, self.counts)
        b, d, r = (x.cpu().numpy() for x in (self.battles, self.durations, self.results))
        seeds = self.seeds.cpu().numpy().view(np.uint64)
        live = np.nonzero((r & 15) == 0)[0]
        emp = np.zeros((n, 2, 9), np.float32)
        nsh = np.zeros((n, 2, 9), np.float32)
        ev, nv = np.zeros(n, np.float32), np.zeros(n, np.float32)
        it = np.zeros(n, np.int32)
        c1, c2 = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        if live.size:
            K = 0xD1B54A32D192ED03
            sseeds = np.array([_splitmix64(int(s) ^ ((K * (2 * self.step_index + 1)) & MASK64)) for s in seeds[live]], np.uint64)
            self.search.set_roots(b[live], d[live], r[live], sseeds)
            self.search.run(self.iterations, c=self.ucb_c, root_rolls=self.rolls[0], other_rolls=self.rolls[1])
            out = self.search.output()
            np_, nv_ = self.search.nash()
            emp[live], nsh[live] = out["empirical"].astype(np.float32), np_
            ev[live], nv[live] = out["empirical_value"].astype(np.float32), nv_
            it[live] = out["iterations"].astype(np.int32)
            self.searched += int(out["iterations"].sum())
            for p, dst in ((0, c1), (1, c2)):
                pol = runtime_policy(out["empirical"][:, p], np_[:, p], self.policy_mode, self.temperature, self.min_prob)
                u = np.array([(_splitmix64(int(s) ^ ((K * (2 * self.step_index + 2)) & MASK64) ^ p) >> 11) * 2.0 ** -53 for s in seeds[live]])
                # mt19937::sample_pdf (util/random.h:41-50): first index whose cumulative probability reaches u
                pick = np.minimum((np.cumsum(pol, -1) < u[:, None]).sum(-1), out["k"][:, p] - 1)
                pick = np.where(pol[np.arange(live.size), pick] > 0, pick, pol.argmax(-1))
                dst[live] = out["choices"][np.arange(live.size), p, pick]
        dev = self.dev
        self.c1.copy_(torch.from_numpy(c1).to(dev))
        self.c2.copy_(torch.from_numpy(c2).to(dev))
        self.probs.copy_(torch.from_numpy(emp).to(dev))
        self.nash_probs.copy_(torch.from_numpy(nsh).to(dev))
        self.value.copy_(torch.from_numpy(ev).to(dev))
        self.nash_value.copy_(torch.from_numpy(nv).to(dev))
        self.iters_t.copy_(torch.from_numpy(it).to(dev))
        self.before.copy_(self.results)
        e.frames_record_device(self.before, self.c1, self.c2, self.counts, self.value, self.nash_value, self.probs, self.nash_probs,
                               iterations=self.iters_t)
        e.step_device(self.battles, self.durations, self.results, self.c1, self.c2)
        e.frames_finish_device(self.results)
        self.step_index += 1
