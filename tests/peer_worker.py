"""torchrun worker for test_gpu_peer.py: 2 ranks render interleaved samples of cornell_box into
librtb200-owned buffers; rank 0 combines them with the fused P2P reduce + resolve kernel and checks
the image against its own single-GPU render of all samples and against the NCCL reduce path."""
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
from raytracing_rust_b200 import distributed as D

rank, local, world = D.init_from_env()
torch.cuda.set_device(local)
nx, ny, ns = 200, 160, 32
scene, w, cam = rtb.Scene.named("cornell_box", nx, ny)
scene.upload(w, local)
red = D.PeerReducer(nx, ny, local)
opts = rtb.default_opts(device=local)
D.render_sharded(scene, cam, nx, ny, ns, opts, red.acc, rank, world, torch.cuda.current_stream().cuda_stream)
rgb8 = torch.empty(ny, nx, 3, dtype=torch.uint8, device="cuda")
total = torch.empty(ny, nx, 3, dtype=torch.float32, device="cuda")
red.reduce_resolve(ns, rgb8, total)
# NCCL path on a copy of the partial sums
acc2 = red.acc.clone()
D.reduce_accumulators(acc2, dst=0)
if rank == 0:
    ref_sum, ref8, _ = scene.render(w, cam, nx, ny, ns, rtb.default_opts(device=local))
    t = total.cpu().numpy()
    np.testing.assert_allclose(t, ref_sum, rtol=2e-5, atol=1e-4)  # same samples, different f32 summation order
    np.testing.assert_allclose(t, acc2.cpu().numpy(), rtol=2e-6, atol=1e-5)
    diff = np.abs(rgb8.cpu().numpy().astype(int) - ref8.astype(int))
    assert diff.max() <= 1 and (diff > 0).mean() < 1e-3, (diff.max(), (diff > 0).mean())
    print("peer reduce ok", float(t.mean()), flush=True)
dist.barrier()
red.close()

# ---- the sharded exchange (every rank reduces 1/G of the image) and row-band sharding, P2P and NCCL forms ----------------
stream = torch.cuda.current_stream().cuda_stream
if rank == 0:
    ref8_t = torch.from_numpy(ref8).cuda()
for mode in ("p2p", "nccl"):
    ex = D.ShardedExchange(nx, ny, local, mode=mode)
    D.render_sharded(scene, cam, nx, ny, ns, rtb.default_opts(device=local), ex.acc, rank, world, stream)
    img = ex.reduce_resolve(ns, stream)
    torch.cuda.synchronize()
    if rank == 0:
        d = (img.int() - ref8_t.int()).abs()
        assert int(d.max()) <= 1 and float((d > 0).float().mean()) < 1e-3, (mode, int(d.max()))
    dist.barrier()
    ex.acc.zero_()
    D.render_band_sharded(scene, cam, nx, ny, ns, rtb.default_opts(device=local), ex.acc, rank, world, stream)
    img = ex.gather_bands(ns, stream)
    torch.cuda.synchronize()
    if rank == 0:
        d = (img.int() - ref8_t.int()).abs()
        assert int(d.max()) <= 1 and float((d > 0).float().mean()) < 1e-3, (mode, "bands", int(d.max()))
    dist.barrier()
    ex.close()
if rank == 0:
    print("sharded exchange ok", flush=True)
dist.barrier()
dist.destroy_process_group()
