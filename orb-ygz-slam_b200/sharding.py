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


def gather_ragged(counts_local, payload_local, n_total, cap, dst=0):
    """The layout of ygzorb_extract_batch_sharded's gather, over torch.distributed (gloo in the CPU tests): counts first, then
    the packed ragged payload.  counts_local: (n_local,) int32; payload_local: (sum(counts_local), ...) rows of this rank's
    frames, back to back.  On `dst` returns (counts (n_total,), frame_offset (n_total,) int64, payload (n_total * cap, ...)):
    the rows of rank r's frames lie back to back from row r_begin * cap on; None elsewhere."""
    world, rank = dist.get_world_size(), dist.get_rank()
    ranges = [shard_range(n_total, r, world) for r in range(world)]
    totals = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(totals, torch.tensor([int(counts_local.sum())], dtype=torch.int64))      # sizes first: both sides of a transfer know it
    totals = [int(t.item()) for t in totals]
    if rank != dst:
        if len(counts_local):
            dist.send(counts_local.contiguous(), dst)
        if totals[rank]:
            dist.send(payload_local.contiguous(), dst)
        return None
    counts = torch.zeros(n_total, dtype=torch.int32)
    payload = torch.zeros((n_total * cap,) + tuple(payload_local.shape[1:]), dtype=payload_local.dtype)
    for r, (b, e) in enumerate(ranges):
        if r == dst:
            counts[b:e] = counts_local
            payload[b * cap:b * cap + totals[r]] = payload_local
            continue
        if e > b:
            buf = torch.zeros(e - b, dtype=torch.int32)
            dist.recv(buf, r)
            counts[b:e] = buf
        if totals[r]:
            buf = torch.zeros((totals[r],) + tuple(payload_local.shape[1:]), dtype=payload_local.dtype)
            dist.recv(buf, r)
            payload[b * cap:b * cap + totals[r]] = buf
    off = torch.zeros(n_total, dtype=torch.int64)
    for b, e in ranges:
        o = b * cap
        for f in range(b, e):
            off[f] = o
            o += int(counts[f])
    return counts, off, payload


def max_over_ranks(value, device):
    t = torch.tensor([float(value)], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
