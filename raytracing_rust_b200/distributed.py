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


class PeerReducer:
    """The exchange step as ONE kernel over peer memory instead of an NCCL reduce + resolve.

    Every rank allocates its accumulation buffer through librtb200 (`rtb.DeviceBuffer`) and
    exports it (CUDA IPC); the root maps the peers' buffers once.  `reduce_resolve` then
      1. waits until every rank's render kernels finished (stream sync + barrier),
      2. on the root, launches k_reduce_resolve (csrc/peer.cu): P2P loads of all G partial sums
         over NVLink, added in rank order, quantised straight to rgb8 (tests/test.rs:71-78),
      3. releases the peers (barrier) so they may clear / reuse their buffers.
    Works on one node (NVLink / NVSwitch); `dist` only carries the 64-byte handles and barriers."""

    def __init__(self, nx: int, ny: int, local_device: int, root: int = 0):
        import raytracing_rust_b200 as rtb

        self.rtb, self.nx, self.ny, self.device, self.root = rtb, nx, ny, local_device, root
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        if self.world > rtb.MAX_PEER_BUFFERS:
            raise ValueError(f"PeerReducer supports up to {rtb.MAX_PEER_BUFFERS} ranks")
        self.buf = rtb.DeviceBuffer((ny, nx, 3), "float32", local_device)
        self.acc = torch.as_tensor(self.buf, device=torch.device("cuda", local_device))  # zero-copy view
        handles = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(handles, self.buf.peer_handle())
        self.ptrs, self._opened = [], []
        if self.rank == root:
            for r in range(self.world):
                if r == self.rank:
                    self.ptrs.append(self.buf.ptr)
                else:
                    p = rtb.peer_open(handles[r], local_device)
                    self._opened.append(p)
                    self.ptrs.append(p)
        if self.world > 1:
            dist.barrier()

    def reduce_resolve(self, ns: int, rgb8: torch.Tensor, total: torch.Tensor | None = None, stream: int = 0):
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier()
        if self.rank == self.root:
            self.rtb.reduce_resolve_device(self.ptrs, self.nx, self.ny, ns, rgb8.data_ptr(),
                                           total.data_ptr() if total is not None else 0, stream)
            torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier()

    def close(self):
        for p in self._opened:
            self.rtb.peer_close(p)
        self._opened = []
        del self.acc
        self.buf.free()


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
