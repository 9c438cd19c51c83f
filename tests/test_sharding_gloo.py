"""World-size-2 (and 3) gloo runs on CPU of the multi-GPU host logic: contiguous shards, ragged gather in
shard order, max-over-ranks timing reduction (orb-ygz-slam_b200/sharding.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ygz import PKG  # noqa: F401  (puts the package dir on sys.path)
import sharding


def test_shard_ranges_cover_and_are_contiguous():
    for n in (0, 1, 7, 4096, 1000003):
        for world in (1, 2, 3, 8):
            r = [sharding.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1 and sizes == sharding.shard_sizes(n, world)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, n):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b, e = sharding.shard_range(n, rank, world)
        # per-query result rows (idx, d1, d2) of a fake kNN: value encodes the global query index
        q = torch.arange(b, e, dtype=torch.int32)
        local = torch.stack([q * 3, q % 257, (q * 7) % 257], dim=1)
        full = sharding.gather_rows(local, n, dst=0)
        t = sharding.max_over_ranks(10.0 + rank, torch.device("cpu"))
        assert t == 10.0 + world - 1
        if rank == 0:
            g = torch.arange(n, dtype=torch.int32)
            assert full.shape == (n, 3)
            assert torch.equal(full, torch.stack([g * 3, g % 257, (g * 7) % 257], dim=1))
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 1001), (3, 10)])
def test_gather_rows_gloo(world, n):
    mp.spawn(_worker, args=(world, _free_port(), n), nprocs=world, join=True)


def _ragged_worker(rank, world, port, n, cap):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ygz import Y
        b, e = sharding.shard_range(n, rank, world)
        assert (b, e) == Y.shard_range(n, rank, world)            # the library's rule (ygzorb_shard_range) == the host mirror
        rng = np.random.default_rng(1234)                          # same stream on every rank: the global truth
        counts_all = rng.integers(0, cap + 1, size=n).astype(np.int32)
        counts_all[::5] = 0
        rows_all = [np.stack([np.full(c, f, np.int32), np.arange(c, dtype=np.int32)], 1) for f, c in enumerate(counts_all)]
        local_rows = [rows_all[f] for f in range(b, e)]
        payload = torch.from_numpy(np.concatenate(local_rows) if local_rows else np.zeros((0, 2), np.int32))
        res = sharding.gather_ragged(torch.from_numpy(counts_all[b:e].copy()), payload, n, cap)
        if rank == 0:
            counts, off, rows = res
            assert np.array_equal(counts.numpy(), counts_all)
            for r in range(world):
                rb, re = sharding.shard_range(n, r, world)
                if re > rb:
                    assert int(off[rb]) == rb * cap                 # a rank's region starts at its first frame's fixed-stride slot
            for f in range(n):
                o, c = int(off[f]), int(counts_all[f])
                assert np.array_equal(rows[o:o + c].numpy(), rows_all[f])
        else:
            assert res is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,cap", [(2, 37, 6), (3, 4, 3)])
def test_counts_then_payload_gather_layout_gloo(world, n, cap):
    """The gather layout of ygzorb_extract_batch_sharded (counts first, then the packed ragged payload; per-rank regions),
    world size 2 and 3 on CPU."""
    mp.spawn(_ragged_worker, args=(world, _free_port(), n, cap), nprocs=world, join=True)
