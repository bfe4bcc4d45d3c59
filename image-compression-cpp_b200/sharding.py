"""Data-parallel plumbing of the reward path: contiguous image shards per rank, and the path's only exchange --
an all-gather of the per-image rewards (the reference is single-GPU,
include/ema_baseline.hpp:13 takes the mean of the whole batch).  Once every rank holds the whole reward vector in image
order, Rule E runs on it locally: one collective per step, and the EMA value is bit-identical to the single-GPU one
because the sum runs over the same values in the same order.  `exchange` keeps the all-reduce of {sum, count} for
callers that only want the statistics.
torch.distributed is the transport (NCCL over NVLink on the GPU box, gloo in the CPU tests); nothing here computes."""
import torch
import torch.distributed as dist


def shard_range(B, world, rank):
    """Contiguous block of images owned by `rank`; the first B % world ranks take one extra image."""
    base, extra = divmod(B, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_rewards(rewards_local, B_total, group=None):
    """rewards_local f64 [b_r] -> rewards of the whole batch f64 [B_total] in image order; ONE all-gather.
    Ragged shards are padded to the largest shard."""
    world = dist.get_world_size(group)
    sizes = [shard_range(B_total, world, r)[1] - shard_range(B_total, world, r)[0] for r in range(world)]
    mx = max(sizes)
    pad = rewards_local
    if rewards_local.numel() < mx:
        pad = torch.zeros(mx, dtype=rewards_local.dtype, device=rewards_local.device)
        pad[:rewards_local.numel()] = rewards_local
    gathered = torch.empty(mx * world, dtype=rewards_local.dtype, device=rewards_local.device)
    dist.all_gather_into_tensor(gathered, pad.contiguous(), group=group)
    if all(s == mx for s in sizes):
        return gathered
    return torch.cat([gathered[r * mx:r * mx + sizes[r]] for r in range(world)])


def exchange(rewards_local, sum_count_local, B_total, group=None):
    """rewards_local f64 [b_r] (+ its {sum, count} f64 [2]) -> (rewards of the whole batch f64 [B_total] in image order,
    global {sum, count}).  Ragged shards are padded to the largest shard for the all-gather."""
    world = dist.get_world_size(group)
    sizes = [shard_range(B_total, world, r)[1] - shard_range(B_total, world, r)[0] for r in range(world)]
    mx = max(sizes)
    pad = rewards_local
    if rewards_local.numel() < mx:
        pad = torch.zeros(mx, dtype=rewards_local.dtype, device=rewards_local.device)
        pad[:rewards_local.numel()] = rewards_local
    gathered = torch.empty(mx * world, dtype=rewards_local.dtype, device=rewards_local.device)
    dist.all_gather_into_tensor(gathered, pad.contiguous(), group=group)
    sc = sum_count_local.clone()
    dist.all_reduce(sc, op=dist.ReduceOp.SUM, group=group)
    if all(s == mx for s in sizes):
        return gathered, sc
    return torch.cat([gathered[r * mx:r * mx + sizes[r]] for r in range(world)]), sc
