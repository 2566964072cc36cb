"""Time-to-image of ONE final_scene 4K image at N GPUs for each way of splitting the work (run under torchrun):
  spp/p2p    samples sharded, sharded P2P exchange            (bench.py default)
  spp/nccl   samples sharded, reduce_scatter + resolve + gather
  spp/root   samples sharded, NCCL reduce of whole buffers to rank 0 + resolve (the r1 form)
  bands/p2p  4-row bands sharded (north_star's "image tiles"), each rank quantises its rows into rank 0's image
  bands/nccl bands sharded, packed rows gathered
at a low and a high sample count.  Scene resident; timed from a barrier to the image being complete on rank 0's device
(max over ranks, best of `reps`).  Prints a markdown table on rank 0.
usage: torchrun ... tools/shard_modes_bench.py [spp_low spp_high]"""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

import raytracing_rust_b200 as rtb
from raytracing_rust_b200 import distributed as D

rank, local, world = D.init_from_env()
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
NX, NY = 3840, 2160
spps = [int(x) for x in sys.argv[1:3]] if len(sys.argv) > 2 else [64, 10000]
scene, w, cam = rtb.Scene.named("final_scene", NX, NY)
scene.upload(w, local)
stream = torch.cuda.current_stream().cuda_stream
rows = []
exs = {"p2p": D.ShardedExchange(NX, NY, local, mode="p2p"), "nccl": D.ShardedExchange(NX, NY, local, mode="nccl")}
root_acc = torch.zeros(NY, NX, 3, dtype=torch.float32, device=dev)
root8 = torch.empty(NY, NX, 3, dtype=torch.uint8, device=dev)


def one_image(mode, ns):
    split, xch = mode.split("/")
    ex = exs.get(xch)
    acc = ex.acc if ex is not None else root_acc
    acc.zero_()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    o = rtb.default_opts(device=local)
    if split == "spp":
        D.render_sharded(scene, cam, NX, NY, ns, o, acc, rank, world, stream)
    else:
        D.render_band_sharded(scene, cam, NX, NY, ns, o, acc, rank, world, stream)
    torch.cuda.synchronize()
    t_render = time.perf_counter() - t0
    if ex is not None:
        (ex.reduce_resolve if split == "spp" else ex.gather_bands)(ns, stream)
    else:
        D.reduce_accumulators(acc, dst=0)
        rtb.resolve_device(acc.data_ptr(), NX, NY, ns, root8.data_ptr(), stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_all = time.perf_counter() - t0
    t = torch.tensor([t_all, t_render], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0]), float(t[1])


modes = ["spp/p2p", "spp/nccl", "spp/root", "bands/p2p", "bands/nccl"]
for ns in spps:
    for mode in modes:
        one_image(mode, min(ns, 8 * world))  # warm-up of this path
        best = min((one_image(mode, ns) for _ in range(3 if ns <= 1000 else 1)), key=lambda x: x[0])
        rows.append((ns, mode, best[0], best[1]))
if rank == 0:
    print(f"| GPUs | spp of the image | split / exchange | time to image ms | slowest rank's render ms | exchange + skew ms | Mpaths/s |")
    print("|---:|---:|---|---:|---:|---:|---:|")
    for ns, mode, t_all, t_r in rows:
        print(f"| {world} | {ns} | {mode} | {t_all * 1e3:.2f} | {t_r * 1e3:.2f} | {(t_all - t_r) * 1e3:.2f} | {NX * NY * ns / t_all / 1e6:.0f} |", flush=True)
for ex in exs.values():
    ex.close()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
