"""Host-side shard plan for multi-GPU batches.

Rollouts are independent given their inputs and their offset in the ONE model stream the
reference's batched inference consumes (monte-carlo-model.hh:58-67), so a batch is cut into
contiguous rollout-index ranges, one per rank, with no data-path collective.  Rank r starts
its copy of the model engine `words_per_rollout * first` draws further on, and the only
exchange is a sum of the eight u64 counters.  This module is pure host logic (no CUDA): the
engine jump itself is `Engine.discard` (pcu_mt_discard, polynomial jump-ahead).
"""
from __future__ import annotations

from dataclasses import dataclass

WORDS_PER_STEP = 4  # two random_int per step, two 32-bit draws each (random.hh:26,28)


@dataclass(frozen=True)
class Shard:
    rank: int
    world: int
    first: int          # first global rollout index of this rank
    count: int          # rollouts of this rank
    engine_offset: int  # model-stream draws consumed by ranks before this one


def plan(n_total: int, world: int, rank: int, depth: int) -> Shard:
    """Contiguous split of n_total uniform-depth rollouts: the first n_total % world ranks
    take one extra rollout.  depth = depth_bound (RandomTree) or max_depth (MoldState)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, extra = divmod(n_total, world)
    first = rank * base + min(rank, extra)
    count = base + (1 if rank < extra else 0)
    return Shard(rank, world, first, count, WORDS_PER_STEP * depth * first)


def plan_ragged(depths, world: int, rank: int) -> Shard:
    """Same for per-rollout depths (a sequence of ints): the engine offset is the number of
    draws the rollouts before `first` consume, 4 * sum(depth[:first])."""
    n_total = len(depths)
    base, extra = divmod(n_total, world)
    first = rank * base + min(rank, extra)
    count = base + (1 if rank < extra else 0)
    return Shard(rank, world, first, count, WORDS_PER_STEP * int(sum(int(d) for d in depths[:first])))


def reduce_sums(sums, dist=None):
    """Sum the per-rank counters [wins, draws, losses, steps, hash sum, 0, 0, 0] over the
    process group (torch.distributed: NCCL on GPUs, gloo in the CPU tests).  `sums` is a
    torch int64 tensor on the backend's device; returns it reduced in place."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    return sums
