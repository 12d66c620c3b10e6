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

from .engine import BattleEngine

MASK64 = (1 << 64) - 1


def _splitmix64(x):
    z = (x + 0x9E3779B97F4A7C15) & MASK64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
    return z ^ (z >> 31)


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
        seeds = np.array([_splitmix64(self.seed ^ int(g)) for g in ids], np.uint64)
        di = torch.from_numpy(idx).to(self.dev)
        self.teams[di] = torch.from_numpy(teams).to(self.dev)
        self.seeds[di] = torch.from_numpy(seeds.view(np.int64)).to(self.dev)
        self.eng.respawn_device(self.battles, self.durations, self.results, self.teams, self.seeds)
        self.eng.frames_begin_device(self.battles, self.results)

    def step(self):
        """One lockstep turn of every running game."""
        e = self.eng
        e.choices_device(self.battles, self.results, self.choices, self.counts)
        e.evaluate_device(self.battles, self.durations, self.results, self.value, self.logits, self.ecounts)
        e.policy_sample_device(self.choices, self.counts, self.logits, self.seeds, self.step_index, self.probs, self.c1, self.c2,
                               self.temperature, self.min_prob)
        self.before.copy_(self.results)
        e.frames_record_device(self.before, self.c1, self.c2, self.counts, self.value, self.value, self.probs, self.probs)
        e.step_device(self.battles, self.durations, self.results, self.c1, self.c2)
        e.frames_finish_device(self.results)
        self.step_index += 1

    def collect(self):
        """Drain finished games -> (stream, games) and restart their slots."""
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
