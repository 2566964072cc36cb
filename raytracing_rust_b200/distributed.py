"""Multi-GPU plumbing: one process per GPU, samples sharded across ranks, one sum-reduce.

The reference's pixel loop has no inter-sample state (tests/test.rs:64-70), so GPU g of G renders
samples s in {g, g+G, ...} of every pixel into a private float3[W*H] accumulation buffer
(Philox counters make the sample set independent of G) and the buffers are combined by ONE
exchange step at the end: reduce(sum) to rank 0 over NCCL/NVLink (gloo on CPU in the tests),
followed by the resolve kernel (/ns, sqrt, clamp, quantise; tests/test.rs:71-78).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def shard_samples(ns: int, rank: int, world: int):
    """(sample_begin, sample_stride, my_count) of rank's slice {rank, rank+world, ...} < ns."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    count = (ns - rank + world - 1) // world if rank < ns else 0
    return rank, world, count


def init_from_env(backend: str | None = None):
    """Initialise torch.distributed from torchrun's environment; returns (rank, local_rank, world)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, local, world


def reduce_accumulators(acc: torch.Tensor, dst: int = 0) -> torch.Tensor:
    """Sum the per-rank radiance accumulation buffers into rank `dst` (in place)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM)
    return acc


def render_sharded(scene, cam, nx, ny, ns, opts, acc: torch.Tensor, rank: int, world: int, stream: int = 0):
    """Adds this rank's share of the ns samples into `acc` (CUDA float32 [ny,nx,3], caller-zeroed)
    on the given CUDA stream, without synchronising.  The scene must be uploaded."""
    begin, stride, _ = shard_samples(ns, rank, world)
    opts.sample_begin, opts.sample_stride = begin, stride
    scene.render_device(cam, nx, ny, ns, opts, acc.data_ptr(), stream)
    return acc


def resolve_host(acc_sum, ns):
    """tests/test.rs:71-78 on a host array (numpy or CPU tensor): /ns, sqrt, clamp, 255.99 truncation."""
    import numpy as np

    a = np.asarray(acc_sum, dtype=np.float64) / float(ns)
    c = np.clip(np.sqrt(np.maximum(a, 0.0)), 0.0, 1.0)
    return np.nan_to_num(255.99 * c, nan=0.0).astype(np.int32).astype(np.uint8)
