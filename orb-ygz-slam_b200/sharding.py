"""Host-side sharding of the two partitioned workloads: frames of a batched extraction
and query rows of a brute-force kNN are independent, so each rank takes a contiguous range and the only
communication is the gather of the results (NCCL on GPUs, gloo in the CPU tests).  No data-path collective.
"""
import torch
import torch.distributed as dist


def shard_range(n, rank, world):
    """Contiguous [begin, end) of rank `rank` when n units are split over `world` ranks (sizes differ by <= 1)."""
    base, rem = divmod(n, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_sizes(n, world):
    return [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]


def gather_rows(local, n_total, dst=0):
    """Gathers per-unit result rows (first dim = this rank's units, in shard order) to `dst`.
    Ragged shards are padded to the largest one for the collective.  Returns the (n_total, ...) tensor on dst,
    None elsewhere."""
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = shard_sizes(n_total, world)
    mx = max(sizes)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst)
    if rank != dst:
        return None
    return torch.cat([b[:s] for b, s in zip(bufs, sizes)], dim=0)


def max_over_ranks(value, device):
    t = torch.tensor([float(value)], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
