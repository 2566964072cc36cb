"""Host-side multi-GPU logic on CPU: the sample partition and the world_size-2 gloo reduce."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raytracing_rust_b200 import distributed as D


def test_sample_partition_is_exact_cover():
    for ns in (0, 1, 7, 64, 100, 10_000):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for rank in range(world):
                b, s, c = D.shard_samples(ns, rank, world)
                mine = list(range(b, ns, s))
                assert len(mine) == c
                seen += mine
            assert sorted(seen) == list(range(ns))
    with pytest.raises(ValueError):
        D.shard_samples(8, 2, 2)


def test_row_band_partition_is_exact_cover():
    for ny in (1, 3, 4, 5, 64, 117, 2160):
        for world in (1, 2, 3, 8):
            seen = []
            for rank in range(world):
                rows = D.shard_bands(ny, rank, world)
                assert all((r // 4) % world == rank for r in rows)
                seen += rows
            assert sorted(seen) == list(range(ny))


def test_resolve_host_matches_reference_quantisation():
    acc = np.array([[[0.0, 100.0, 25.0], [400.0, np.nan, -1.0]]])  # sums over ns = 100
    out = D.resolve_host(acc, 100)
    exp = [[0, int(255.99 * 1.0), int(255.99 * 0.5)], [255, 0, 0]]
    np.testing.assert_array_equal(out[0], exp)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, _, w = D.init_from_env(backend="gloo")
    assert (r, w) == (rank, world)
    ns, npix = 10, 6
    # stand-in for the per-rank render: sample s of pixel p contributes (s + 1) * (p + 1)
    b, st, c = D.shard_samples(ns, rank, world)
    acc = torch.zeros(npix, 3)
    for s in range(b, ns, st):
        acc += (s + 1) * torch.arange(1, npix + 1, dtype=torch.float32)[:, None]
    D.reduce_accumulators(acc, dst=0)
    if rank == 0:
        torch.save(acc, out_path)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_reduce_equals_single_rank_sum(tmp_path):
    out = str(tmp_path / "acc.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    acc = torch.load(out)
    full = sum(range(1, 11)) * torch.arange(1, 7, dtype=torch.float32)[:, None].expand(6, 3)
    torch.testing.assert_close(acc, full)
