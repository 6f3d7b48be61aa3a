"""CPU-testable mirror of the multi-GPU split / combine arithmetic.

On GPUs the composition lives inside liblxyb200 (lxy_b200/csrc/comm.cu: lxy_comm_*,
lxy_scaff_score_sharded*, lxy_ga_set_sharding - NCCL on the context's own stream).  This module restates
the same arithmetic over torch.distributed so that the world-size-2 tests can run it on the gloo backend
without a GPU (tests/test_sharding_gloo.py; tests/test_abi.py checks shard_range against lxy_comm_slice).
No fitness is computed here - these helpers only split the work and combine what a per-rank scorer produced.

  population-sharded (BASELINE.json configs[2]): links replicated, candidates split across ranks,
      all-gather of the score slices; nothing else moves.
  link-sharded (configs[3]): records split across ranks, every rank scores the whole population
      against its shard, the partial sums are combined: exact int64 partials by all-reduce(SUM)
      (order-independent, bit-identical to the unsharded sum), fp64 partials by all-gather and a
      rank-ordered sum (deterministic run to run).
"""
import torch
import torch.distributed as dist


def shard_range(total, world, rank):
    """Contiguous, balanced split of range(total): the first total % world ranks get one more."""
    base, extra = divmod(int(total), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_counts(total, world):
    return [shard_range(total, world, r)[1] - shard_range(total, world, r)[0] for r in range(world)]


def gather_scores(local, total, group=None):
    """All-gather the per-rank score vectors (possibly of unequal length) into one [total] vector."""
    world = dist.get_world_size(group)
    counts = shard_counts(total, world)
    if len(set(counts)) == 1:
        out = torch.empty(total, dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    width = max(counts)
    padded = torch.zeros(width, dtype=local.dtype, device=local.device)
    padded[:local.numel()] = local
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded, group=group)
    return torch.cat([p[:c] for p, c in zip(parts, counts)])


def combine_link_partials_exact(x100_partial, group=None):
    """Sum of the per-shard exact integers sum(100*T*w*m): int64 all-reduce."""
    out = x100_partial.clone()
    dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out


def combine_link_partials_f64(partial, group=None):
    """fp64 partial fitness: all-gather then add in rank order (fixed order => reproducible)."""
    world = dist.get_world_size(group)
    parts = [torch.empty_like(partial) for _ in range(world)]
    dist.all_gather(parts, partial.contiguous(), group=group)
    total = torch.zeros_like(partial)
    for p in parts:
        total = total + p
    return total


def fitness_from_x100(x100):
    return -(x100.to(torch.float64) / 100.0)


def broadcast_elite(elite, src=0, group=None):
    """Elite individuals travel by broadcast (north_star): `elite` is an int32 [E][n] tensor."""
    dist.broadcast(elite, src=src, group=group)
    return elite


# ---- GA generation with sharded scoring ---------------------------------------------------------
# Every rank runs the same deterministic plan (include/lxy_ga_plan.h) and therefore holds the same
# children; rank r scores the slice ga_child_slice(C, world, r) and the fitness vector is
# all-gathered in place into the engine's score array (capacity >= world * width).
def ga_child_slice(n_children, world, rank):
    """(first, count, width): equal-width slices of ceil(C / world) children, the last ones short."""
    width = (int(n_children) + world - 1) // world
    lo = min(rank * width, n_children)
    hi = min((rank + 1) * width, n_children)
    return lo, hi - lo, width


def ga_gather_child_scores(scores, n_children, group=None):
    """`scores`: this rank's view of the children's score array (length >= world * width) with its own
    slice filled in; afterwards entries [0, C) hold every rank's scores."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    _, _, width = ga_child_slice(n_children, world, rank)
    if width == 0:
        return scores
    mine = scores[rank * width:(rank + 1) * width].clone()
    dist.all_gather_into_tensor(scores[:world * width], mine, group=group)
    return scores
