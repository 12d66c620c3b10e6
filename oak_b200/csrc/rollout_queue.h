#pragma once
#include <stdint.h>

// The shared playout queue of the tree search (search.cu; kernel in battle_kernels.cuh). The search runs many small root
// groups, each arriving at its leaves at its own time with a batch of a few thousand playouts. A launch per batch would run
// at the pace of the batch's longest playout with most lanes idle (playout lengths differ by an order of magnitude), so the
// batches share the lanes of all workers: a group's batch is its particle range [first, first + count) and one counter
// `next[g]` (the group publishes a batch by zeroing it: one 4-byte memset on its stream, ordered after its tree levels);
// a worker warp with idle lanes reads all counters in one coalesced load, picks a group that has playouts left and takes as
// many as it has idle lanes with ONE atomicAdd — an add that overshoots the count costs nothing, so there is no
// compare-and-swap to retry and nothing to wait for. Lanes refill the moment their playout ends, whichever group the next
// one belongs to. Every batch launches enough CTAs for itself; the CTAs count themselves in `live` and one that finds the
// SMs' share already taken (max_live) leaves at once, so together the workers hold a fixed part of every SM and the tree
// levels' kernels fit beside them. A CTA stays while any of its warps has work (its idle warps keep looking) and takes
// itself out of `live` BEFORE its last look at the counters: a batch published after that look finds room for its own
// worker, which is launched after it — that is what guarantees every playout is run, and no kernel ever waits for something
// the host does later. The last lane to finish a group's batch bumps the group's word in page-locked host memory, which
// the group's host thread polls.
struct oakb200_rollout_queue {
  uint32_t* next;             // [32] playouts of the group's current batch handed out so far (>= count: none left)
  uint32_t* done;             // [32] playouts of the group's current batch finished so far
  const uint32_t* group_first;  // [n_groups + 1] first particle index of each group
  uint32_t n_groups;          // <= 32: one lane looks at one group
  uint32_t* live;             // worker CTAs running now
  uint32_t max_live;
  uint32_t poll_mask;         // a warp with idle lanes looks at the counters every (poll_mask + 1)-th update
  uint32_t idle_ns;           // pause of a warp with nothing to do between looks
  uint32_t linger;            // looks an idle CTA makes before it leaves (the gaps between batches are short)
  volatile uint32_t* host_batches;  // page-locked, mapped: [32] batches completed
};
