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


class ShardedExchange:
    """The exchange step SHARDED over the ranks: nobody pulls whole images.

    Sample sharding (`reduce_resolve`): rank g owns the value slice [g*per, (g+1)*per) of the image.
      * mode "p2p":  every rank maps every rank's accumulation buffer and the root's rgb8 image (CUDA IPC);
        `rtb_reduce_resolve_range_device` on rank g then pulls its slice of the G partial sums over NVLink
        (7 x 12.4 MB inbound per GPU at 4K on 8 GPUs instead of 7 x 99.5 MB into the root), adds them in rank order and
        stores the quantised slice straight into the root's image.  Two host barriers fence the buffers.
      * mode "nccl": reduce_scatter(sum) -> k_resolve on the slice -> gather of the rgb8 slices (3.1 MB each) to the root.
    Row-band sharding (`gather_bands`; rtb_render_opts.band_begin/stride = rank/world): the accumulators are disjoint, so
    there is no arithmetic — "p2p": each rank quantises its bands into the root's image; "nccl": packed bands are
    gathered and the root scatters the rows.
    The accumulation buffer (`self.acc`, float32 [ny,nx,3]) is owned here; the root's image is `self.rgb8`."""

    def __init__(self, nx: int, ny: int, local_device: int, mode: str = "p2p", root: int = 0):
        import raytracing_rust_b200 as rtb

        if mode not in ("p2p", "nccl"):
            raise ValueError("mode must be 'p2p' or 'nccl'")
        self.rtb, self.nx, self.ny, self.device, self.root, self.mode = rtb, nx, ny, local_device, root, mode
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        if self.world > rtb.MAX_PEER_BUFFERS:
            raise ValueError(f"ShardedExchange supports up to {rtb.MAX_PEER_BUFFERS} ranks")
        dev = torch.device("cuda", local_device)
        self.n = 3 * nx * ny
        self.per = (((self.n + self.world - 1) // self.world) + 3) & ~3  # slice length: a multiple of 4 values
        self.first = min(self.n, self.rank * self.per)
        self.count = min(self.n - self.first, self.per)
        self.buf = rtb.DeviceBuffer((ny, nx, 3), "float32", local_device)
        self.acc = torch.as_tensor(self.buf, device=dev)  # zero-copy view
        self._opened, self.ptrs, self.root_rgb8_ptr = [], [], None
        if mode == "p2p":
            self.rgb8_buf = rtb.DeviceBuffer((ny, nx, 3), "uint8", local_device)
            self.rgb8 = torch.as_tensor(self.rgb8_buf, device=dev)
            handles = [None] * self.world
            if self.world > 1:
                dist.all_gather_object(handles, (self.buf.peer_handle(), self.rgb8_buf.peer_handle()))
            else:
                handles = [(None, None)]
            err = None
            try:
                for r in range(self.world):
                    if r == self.rank:
                        self.ptrs.append(self.buf.ptr)
                    else:
                        p = rtb.peer_open(handles[r][0], local_device)
                        self._opened.append(p)
                        self.ptrs.append(p)
                if self.rank == root:
                    self.root_rgb8_ptr = self.rgb8_buf.ptr
                else:
                    self.root_rgb8_ptr = rtb.peer_open(handles[root][1], local_device)
                    self._opened.append(self.root_rgb8_ptr)
            except Exception as e:  # e.g. no CUDA IPC / no peer path between two of the devices
                err = e
            if self.world > 1:
                # the ranks agree on the outcome, so either all of them own a working exchange or all of them raise here
                # (a rank that gave up alone would leave the others waiting in the next collective)
                ok = torch.tensor([0 if err is not None else 1], device=dev)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN)
                if int(ok) == 0 and err is None:
                    err = RuntimeError("peer mapping failed on another rank")
            if err is not None:
                self.close()
                raise RuntimeError(f"ShardedExchange(p2p): {err}")
        else:
            self.rgb8_buf = None
            self.rgb8 = torch.empty(ny, nx, 3, dtype=torch.uint8, device=dev)
            self._slice = torch.empty(self.per, dtype=torch.float32, device=dev)
            self._slice8 = torch.empty(self.per, dtype=torch.uint8, device=dev)
            self._pad = torch.zeros(self.per * self.world, dtype=torch.float32, device=dev) if self.per * self.world != self.n else None
            self._gather8 = torch.empty(self.world * self.per, dtype=torch.uint8, device=dev) if self.rank == root else None
        if self.world > 1:
            dist.barrier()

    # -- sample sharding: sum of the G partial images, quantised; the image is valid on the root afterwards
    def reduce_resolve(self, ns: int, stream: int = 0):
        if self.mode == "p2p":
            torch.cuda.synchronize(self.device)  # my render kernels are done ...
            if self.world > 1:
                dist.barrier()                   # ... and everybody else's
            self.rtb.reduce_resolve_range_device(self.ptrs, self.first, self.count, ns, self.root_rgb8_ptr, 0, stream)
            torch.cuda.synchronize(self.device)
            if self.world > 1:
                dist.barrier()                   # every slice has landed in the root's image; buffers may be reused
            return self.rgb8
        flat = self.acc.view(-1)
        if self.world > 1:
            src = flat
            if self._pad is not None:
                self._pad[: self.n].copy_(flat)
                src = self._pad
            dist.reduce_scatter_tensor(self._slice, src, op=dist.ReduceOp.SUM)
        else:
            self._slice[: self.n].copy_(flat)
        # quantise my slice: the fused kernel over a single buffer (same quantisation code as everywhere else)
        self.rtb.reduce_resolve_range_device([self._slice.data_ptr()], 0, self.per, ns, self._slice8.data_ptr(), 0, stream)
        if self.world > 1:
            dist.gather(self._slice8, list(self._gather8.view(self.world, self.per).unbind(0)) if self.rank == self.root else None, dst=self.root)
            if self.rank == self.root:
                self.rgb8.view(-1).copy_(self._gather8[: self.n])
        else:
            self.rgb8.view(-1).copy_(self._slice8[: self.n])
        return self.rgb8

    # -- row-band sharding: every rank rendered ALL samples of its bands into self.acc; no arithmetic in the exchange
    def gather_bands(self, ns: int, stream: int = 0):
        rtb = self.rtb
        if self.mode == "p2p":
            rtb.resolve_bands_device(self.buf.ptr, self.nx, self.ny, ns, self.rank, self.world, self.root_rgb8_ptr, False, stream)
            torch.cuda.synchronize(self.device)
            if self.world > 1:
                dist.barrier()
            return self.rgb8
        bands_all = (self.ny + rtb.BAND_ROWS - 1) // rtb.BAND_ROWS
        max_bands = (bands_all + self.world - 1) // self.world
        row = self.nx * 3
        packed = torch.zeros(max_bands * rtb.BAND_ROWS * row, dtype=torch.uint8, device=self.acc.device)
        rtb.resolve_bands_device(self.acc.data_ptr(), self.nx, self.ny, ns, self.rank, self.world, packed.data_ptr(), True, stream)
        if self.world > 1:
            parts = [torch.empty_like(packed) for _ in range(self.world)] if self.rank == self.root else None
            dist.gather(packed, parts, dst=self.root)
        else:
            parts = [packed]
        if self.rank == self.root:
            out = self.rgb8.view(self.ny, row)
            for r, part in enumerate(parts):
                rows = torch.cat([torch.arange(b * rtb.BAND_ROWS, min(self.ny, (b + 1) * rtb.BAND_ROWS)) for b in range(r, bands_all, self.world)]) \
                    if r < bands_all else torch.empty(0, dtype=torch.long)
                if rows.numel():
                    out[rows.to(out.device)] = part.view(-1, row)[: rows.numel()]
        return self.rgb8

    def close(self):
        for p in self._opened:
            try:
                self.rtb.peer_close(p)
            except Exception:
                pass
        self._opened = []
        self.acc = None
        self.buf.free()
        if self.rgb8_buf is not None:
            self.rgb8 = None
            self.rgb8_buf.free()


def render_sharded(scene, cam, nx, ny, ns, opts, acc: torch.Tensor, rank: int, world: int, stream: int = 0):
    """Adds this rank's share of the ns samples into `acc` (CUDA float32 [ny,nx,3], caller-zeroed)
    on the given CUDA stream, without synchronising.  The scene must be uploaded."""
    begin, stride, _ = shard_samples(ns, rank, world)
    opts.sample_begin, opts.sample_stride = begin, stride
    opts.band_begin, opts.band_stride = 0, 1
    scene.render_device(cam, nx, ny, ns, opts, acc.data_ptr(), stream)
    return acc


def shard_bands(ny: int, rank: int, world: int, band_rows: int = 4):
    """Rows of the image rank `rank` of `world` owns under row-band sharding: the 4-row bands rank, rank+world, ..."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    rows = []
    for b in range(rank, (ny + band_rows - 1) // band_rows, world):
        rows += list(range(b * band_rows, min(ny, (b + 1) * band_rows)))
    return rows


def render_band_sharded(scene, cam, nx, ny, ns, opts, acc: torch.Tensor, rank: int, world: int, stream: int = 0):
    """Adds ALL ns samples of this rank's 4-row bands into `acc` (full-size, caller-zeroed): the 'image tiles' split
    of the pixel loop (tests/test.rs:62-63).  Philox counters are per (pixel, sample), so the image equals the
    sample-sharded one."""
    opts.sample_begin, opts.sample_stride = 0, 1
    opts.band_begin, opts.band_stride = rank, world
    scene.render_device(cam, nx, ny, ns, opts, acc.data_ptr(), stream)
    return acc


def resolve_host(acc_sum, ns):
    """tests/test.rs:71-78 on a host array (numpy or CPU tensor): /ns, sqrt, clamp, 255.99 truncation."""
    import numpy as np

    a = np.asarray(acc_sum, dtype=np.float64) / float(ns)
    c = np.clip(np.sqrt(np.maximum(a, 0.0)), 0.0, 1.0)
    return np.nan_to_num(255.99 * c, nan=0.0).astype(np.int32).astype(np.uint8)
