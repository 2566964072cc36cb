#!/usr/bin/env python
"""bench.py — headline benchmark: Mpaths/s on final_scene (BASELINE.json `metric`).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W     # the reference's CPU algorithm (oracle)

Workload: the reference's final_scene (tests/test.rs:419-523) as written, camera of
test_final_scene (tests/test.rs:819-830), at BASELINE.json's 3840x2160.  A *step* renders
SPP_PER_STEP camera samples of every pixel on every rank (the 10k-spp image is 156 such steps
per GPU at N=1); ranks take interleaved sample indices (or, --shard bands, 4-row image bands) and ONE
exchange step closes the timed region at N>1: sharded over the ranks through peer memory (each GPU
reduces + quantises 1/N of the image into rank 0's rgb8 buffer), NCCL reduce_scatter + gather as the
fallback (weak scaling: per-GPU work is fixed).  `e2e` at N>1 is ONE combined image on rank 0's pinned
host buffer per call, fresh scene handles on every rank, barrier to barrier.
A *path* is one top-level color(ray, world, 0) call (tests/test.rs:69).
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

SCENE = "final_scene"
NX, NY = 3840, 2160
SPP_PER_STEP = 64
METRIC = "Mpaths/s (final_scene)"
UNIT = "Mpaths/s"

# fixed FP32 lane-instruction costs per operation of the reference's formulas.
COST = {
    "node_test": 24.0,     # two AABB slab tests with precomputed 1/d (a9: 12 each)
    "prim_test": 20.0,     # sphere test (a11); rect 10, moving sphere 25, cube face scan more — 20 as the mean
    "medium_test": 12.0,   # interval clamps + free flight of one medium (a17: "2 x boundary + 12"); its two boundary queries are
                           # counted by the kernels as primitive tests (render.cu sample_media_const, trace.cuh sample_media) and
                           # priced there — r1 priced them here as well (52), a double count the r1 verdict caught
    "shade": [15.0, 25.0, 45.0, 2.0, 8.0],   # lambertian, metal, dielectric, diffuse_light, isotropic (a18)
    "tex": [0.0, 30.0, 645.0, 10.0],         # solid, checker, noise, image (a19/a20)
    "hit_record": 9.0,     # p / normal finalisation per surface segment
    "draws_per_segment": 5 * 8.0,            # one Philox block of five U01 per bounce (8 each)
    "raygen": 40.0 + 5 * 8.0,
    "accumulate": 3.0,
}


def algorithmic_lane_instr(st):
    """Σ(count_op × cost_op) for one render, from the device counters of an instrumented run."""
    w = st["node_tests"] * COST["node_test"] + st["prim_tests"] * COST["prim_test"]
    w += st["medium_tests"] * COST["medium_test"]
    w += sum(n * c for n, c in zip(st["shade"], COST["shade"]))
    w += sum(n * c for n, c in zip(st["tex_evals"], COST["tex"]))
    w += st["segments"] * (COST["hit_record"] + COST["draws_per_segment"])
    w += st["paths"] * (COST["raygen"] + COST["accumulate"])
    return w


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], threading.Event()
        self.th = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 3 + k and r[3 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    return json.loads(p.read_text()) if p.exists() else {}


def cpu_reference_sample(seconds=15.0, nthreads=None, seed=11, scene_name=None, nx=480, ny=270):
    """Times the CPU oracle (f64 restatement of the reference; the Rust original cannot be built
    here) on a bounded sample of the SAME workload: final_scene, same camera, 480x270, spp sized
    for ~`seconds` of wall time on `nthreads` host threads.  Returns (Mpaths/s, description, cores)."""
    from oracle import oracle_py as orc

    nthreads = nthreads or os.cpu_count() or 1
    scene_name = scene_name or SCENE
    if scene_name == "final_scene":
        # the headline scene comes from committed bytes (tests/golden/final_scene.rtscene.head + assets/earthmap…rgb8) and the
        # oracle's own camera struct: the reference arm loads liboracle.so only, nothing of the product
        o = orc.OracleScene(orc.final_scene_bytes(), orc.FINAL_SCENE_WORLD)
        cam = orc.final_scene_camera(nx, ny)
    else:  # --configs-table rows: the other scenes are built through the product's scene constructors
        import raytracing_rust_b200 as rtb
        s, w, cam = rtb.Scene.named(scene_name, nx, ny)
        o = orc.OracleScene(s.serialize(), w)
    t0 = time.perf_counter()
    o.render(cam, nx, ny, 1, seed=seed, nthreads=nthreads)  # calibration pass (also warms caches)
    per_spp = max(time.perf_counter() - t0, 1e-3)
    spp = max(1, min(4096, int(seconds / per_spp)))
    t0 = time.perf_counter()
    _, _, segs = o.render(cam, nx, ny, spp, seed=seed + 1, nthreads=nthreads)
    dt = time.perf_counter() - t0
    paths = nx * ny * spp
    desc = (f"{scene_name} {nx}x{ny} x {spp} spp = {paths / 1e6:.2f} Mpaths in {dt:.1f} s on {nthreads} threads "
            f"({segs / paths:.2f} segments/path); C++ f64 port of src/*.rs (oracle/oracle.cpp)")
    return paths / dt / 1e6, desc, nthreads, dt


def run_configs_table(args):
    """--configs-table: BASELINE.json's five configs rendered in FULL on one GPU (time-to-image measured, not
    extrapolated), each next to the CPU oracle timed on a bounded sample of the same scene and camera (the same
    cpu_baseline leg as the headline line).  Markdown on stdout (profiles/r1_configs.md)."""
    import raytracing_rust_b200 as rtb

    configs = [("C1", "two_spheres", 400, 225, 100), ("C2", "random_scene", 1920, 1080, 500), ("C3", "two_perlin_sphere", 1920, 1080, 1000),
               ("C3", "earth", 1920, 1080, 1000), ("C4", "cornell_box", 1024, 1024, 10000), ("C4", "cornell_smoke", 1024, 1024, 10000),
               ("C5", "final_scene", 3840, 2160, 10000),
               # the lit variant of C5 (light rect bounds ordered): radiance is non-zero, so the accumulation path (shared-memory
               # atomics, RED.global flush) is live in the timed launch — the as-written scene is black (F4) and never adds
               ("C5-lit", "final_scene_lit", 3840, 2160, 1000)]
    nth = os.cpu_count() or 1
    print(f"| cfg | scene (as written, black background; C5-lit = labelled lit variant) | image | spp | paths | GPU s | GPU Mpaths/s | GPU Mseg/s | seg/path | "
          f"CPU oracle Mpaths/s ({nth} thr) | CPU sample | speed-up |")
    print("|---|---|---|---:|---:|---:|---:|---:|---:|---:|---|---:|")
    for cfg, name, nx, ny, ns in configs:
        s, w, cam = rtb.Scene.named(name, nx, ny)
        s.render(w, cam, nx, ny, 2, want_sum=False, want_rgb8=False)
        t0 = time.perf_counter()
        _, _, st = s.render(w, cam, nx, ny, ns, want_sum=False, want_stats=True)
        wall = time.perf_counter() - t0
        cpu, desc, _, _ = cpu_reference_sample(seconds=min(args.cpu_seconds, 8.0), scene_name=name, nx=max(16, nx // 4), ny=max(16, ny // 4))
        paths = nx * ny * ns
        gpu = paths / (st["device_ms"] * 1e-3) / 1e6
        print(f"| {cfg} | {name} | {nx}x{ny} | {ns} | {paths / 1e9:.3f} G | {st['device_ms'] / 1e3:.2f} (wall {wall:.2f}) | {gpu:.0f} | "
              f"{st['segments'] / st['device_ms'] / 1e3:.0f} | {st['segments'] / st['paths']:.2f} | {cpu:.2f} | {desc.split(' on ')[0]} | "
              f"{gpu / cpu:.0f}x |", flush=True)


def run_reference(args):
    """--impl reference: the reference's CPU algorithm on the box's host cores (oracle port)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per_step = max(2.0, min(20.0, 60.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_reference_sample(seconds=per_step)
    vals, descs, t_tot = [], [], 0.0
    for _ in range(args.steps):
        v, d, cores, dt = cpu_reference_sample(seconds=per_step)
        vals.append(v)
        descs.append(d)
        t_tot += dt
    value = statistics.mean(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{SCENE} (tests/test.rs:419-523 as written), bounded sample per step: {descs[-1]}"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": descs[-1]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_gpu(args):
    # stdout carries exactly one JSON line: anything native libraries print on fd 1 while the job runs
    # (NCCL's version banner) is sent to stderr, and the JSON line is written to the saved stdout
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import numpy as np
    import torch
    import torch.distributed as dist

    import raytracing_rust_b200 as rtb
    from raytracing_rust_b200 import distributed as D

    rank, local, world = D.init_from_env()
    assert torch.cuda.is_available() and rtb.device_count() > 0, "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    pipe = {"auto": rtb.PIPE_AUTO, "mega": rtb.PIPE_MEGAKERNEL, "wavefront": rtb.PIPE_WAVEFRONT}[args.pipeline]
    spp = args.spp_per_step
    bands = args.shard == "bands"
    # weak scaling: every rank renders `spp` samples of every pixel per step (spp sharding), or — row-band sharding —
    # spp * world samples of its 1/world of the rows: the same number of paths per rank per step either way
    ns_total = spp * world  # samples per pixel rendered by one step across all ranks

    scene, w, cam = rtb.Scene.named(SCENE, NX, NY)
    scene.upload(w, local)
    # ---- the exchange step (N > 1).  Default: sharded over the ranks through peer memory; --reduce picks another form
    exchange_info = None
    mode = args.reduce
    ex = None
    legacy = None
    if world > 1:
        if mode in ("auto", "p2p"):
            try:
                ex = D.ShardedExchange(NX, NY, local, mode="p2p")
                mode = "p2p"
            except Exception as e:  # no CUDA IPC between these ranks: NCCL carries the sharded exchange
                if args.reduce == "p2p":
                    raise
                print(f"[bench] P2P exchange unavailable ({e}); using NCCL", file=sys.stderr)
                mode = "nccl"
            ok = torch.tensor([1 if ex is not None else 0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok) == 0 and mode == "p2p":
                if ex is not None:
                    ex.close()
                ex, mode = None, "nccl"
        if mode == "nccl":
            ex = D.ShardedExchange(NX, NY, local, mode="nccl")
        if mode == "p2p-root":
            legacy = D.PeerReducer(NX, NY, local)
    if ex is not None:
        acc, rgb8 = ex.acc, ex.rgb8
    elif legacy is not None:
        acc = legacy.acc
        rgb8 = torch.empty(NY, NX, 3, dtype=torch.uint8, device=dev)
    else:
        acc = torch.zeros(NY, NX, 3, dtype=torch.float32, device=dev)
        rgb8 = torch.empty(NY, NX, 3, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = torch.cuda.current_stream().cuda_stream

    def set_shard(opts, k):
        if bands:  # all ns_total samples of step k, on this rank's 4-row bands
            opts.sample_begin, opts.sample_stride = k * ns_total, 1
            opts.band_begin, opts.band_stride = rank, world
        else:      # step k renders samples [k*ns_total, (k+1)*ns_total) of the image; rank r takes r, r+world, ...
            opts.sample_begin, opts.sample_stride = k * ns_total + rank, world
            opts.band_begin, opts.band_stride = 0, 1

    def step(k, opts):
        set_shard(opts, k)
        return scene.render_device(cam, NX, NY, (k + 1) * ns_total, opts, acc.data_ptr(), stream)

    def exchange(ns_done):
        """Closes the image on rank 0: the path's one exchange step + quantisation.  Returns kernels launched by this rank."""
        if ex is not None:
            (ex.gather_bands if bands else ex.reduce_resolve)(ns_done, stream)
            return 1
        if legacy is not None:
            legacy.reduce_resolve(ns_done, rgb8, stream=stream)
            return 1 if rank == 0 else 0
        if world > 1:  # "nccl-root": sum-reduce of whole buffers to rank 0, then resolve there (the r1 form)
            D.reduce_accumulators(acc, dst=0)
        rtb.resolve_device(acc.data_ptr(), NX, NY, ns_done, rgb8.data_ptr(), stream)
        return 1

    opts = rtb.default_opts(pipeline=pipe, device=local)
    for k in range(args.warmup):
        step(k, opts)
    torch.cuda.synchronize()

    # instrumented (untimed) pass of one step: operation counts for the work model + launch count
    so = rtb.default_opts(pipeline=pipe, device=local, collect_stats=1)
    set_shard(so, 0)
    st = scene.render_device(cam, NX, NY, ns_total, so, acc.data_ptr(), stream, want_stats=True)
    launches_per_step = st["kernel_launches"]
    work_per_step = algorithmic_lane_instr(st)
    acc.zero_()
    exchange(ns_total)  # warm-up of the exchange path (first-use costs: NCCL channels, peer mappings)
    acc.zero_()

    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps + 1)]
    with ClockSampler(local) as clocks:
        for k in range(args.steps):
            flush.fill_(k & 0xFF)  # L2 flush between timed iterations (untimed)
            ev[k][0].record()
            step(k, opts)
            ev[k][1].record()
        ev[-1][0].record()
        exchange_launches = exchange(max(1, args.steps * ns_total))
        ev[-1][1].record()  # recorded after exchange() returned, i.e. after its closing host barrier
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    step_ms = [a.elapsed_time(b) for a, b in ev[:-1]]
    tail_ms = ev[-1][0].elapsed_time(ev[-1][1])
    # (the P2P exchange contains host barriers: its closing event is recorded after exchange() returns, so the event pair
    # brackets them; on a rank that finished its steps early the tail includes waiting for the slowest rank)
    total_ms = torch.tensor([sum(step_ms) + tail_ms], dtype=torch.float64, device=dev)
    kern_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    xch_ms = torch.tensor([tail_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(kern_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(xch_ms, op=dist.ReduceOp.MAX)
    total_ms, kern_ms, xch_ms = float(total_ms), float(kern_ms), float(xch_ms)
    paths_per_step_rank = NX * NY * spp
    value = world * args.steps * paths_per_step_rank / (total_ms * 1e-3) / 1e6
    final8 = rgb8.clone() if rank == 0 else None

    # ---- exchange self-check (N > 1): the P2P and the NCCL forms of the sharded exchange, and the r1 root reduce, give the
    # same image from the same partial sums (<= 1 LSB: float addition order) — run inside the driver's own invocation
    if world > 1:
        exchange_info = {"mode": ("bands-" if bands else "") + mode, "ms": xch_ms}
        if not bands:
            try:
                ns_chk = max(1, args.steps * ns_total)
                keep = acc.clone()
                imgs = {}
                for m2 in ("p2p", "nccl"):
                    if ex is not None and m2 == mode:
                        imgs[m2] = final8
                        continue
                    try:
                        e2 = D.ShardedExchange(NX, NY, local, mode=m2)
                    except Exception as e:
                        e2 = None
                        print(f"[bench] exchange_check: {m2} unavailable: {e}", file=sys.stderr)
                    okf = torch.tensor([1 if e2 is not None else 0], device=dev)
                    dist.all_reduce(okf, op=dist.ReduceOp.MIN)
                    if int(okf) == 0:
                        if e2 is not None:
                            e2.close()
                        continue
                    e2.acc.copy_(keep)
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    e2.reduce_resolve(ns_chk, stream)
                    torch.cuda.synchronize()
                    exchange_info[f"{m2}_ms_checked"] = (time.perf_counter() - t0) * 1e3
                    imgs[m2] = e2.rgb8.clone() if rank == 0 else None
                    e2.close()
                root_acc = keep.clone()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                D.reduce_accumulators(root_acc, dst=0)
                root8 = torch.empty_like(rgb8)
                rtb.resolve_device(root_acc.data_ptr(), NX, NY, ns_chk, root8.data_ptr(), stream)
                torch.cuda.synchronize()
                exchange_info["nccl_root_ms_checked"] = (time.perf_counter() - t0) * 1e3
                if rank == 0:
                    worst = 0
                    for m2, im in imgs.items():
                        worst = max(worst, int((im.int() - root8.int()).abs().max()))
                    exchange_info["exchange_check"] = "ok" if worst <= 1 and len(imgs) >= 1 else f"MISMATCH (max diff {worst})"
                    exchange_info["forms_compared"] = sorted(imgs) + ["nccl-root"]
                    exchange_info["max_lsb_diff"] = worst
                del keep, root_acc
            except Exception as e:
                exchange_info["exchange_check"] = f"error: {e}"

    # ---- e2e: ONE image on rank 0's host through the public API with host buffers, everything inside the timed region
    blob = scene.serialize()
    n_e2e = max(1, min(args.steps, 3))
    e2e = None
    if world == 1:
        # the host-buffer C-ABI call (rtb_render): scene flatten + upload, render, D2H of the image
        t_e2e, h2d, d2h = 0.0, 0, 0
        for k in range(-1, n_e2e):  # call -1 is an untimed warm-up (first-use costs of the library's per-device workspace)
            t0 = time.perf_counter()
            s2 = rtb.Scene.deserialize(blob)  # fresh handle: flatten + BVH build + H2D inside the timed call
            o2 = rtb.default_opts(pipeline=pipe, device=local)
            # create_image returns the quantised image only (tests/test.rs:71-84): the result read back is rgb8
            g_sum, g8, _ = s2.render(w, cam, NX, NY, ns_total, o2, want_sum=False)
            dt_call = time.perf_counter() - t0
            print(f"[bench] e2e call {k}: {dt_call * 1e3:.1f} ms" + (" (warm-up, not counted)" if k < 0 else ""), file=sys.stderr)
            if k >= 0:
                t_e2e += dt_call
            h2d = s2.device_bytes() + 256  # scene tables + camera/params
            d2h = g8.nbytes
            del s2
        e2e = {"value": paths_per_step_rank / (t_e2e / n_e2e) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "note": "rtb_render() on a fresh scene handle with host buffers: deserialize + flatten + BVH build + H2D, render, resolve, D2H of the rgb8 image (what create_image returns); host wall clock, mean of the calls after one untimed warm-up call"}
    else:
        # every rank: fresh scene handle from the host bytes -> flatten + BVH + H2D -> its shard of ONE ns_total-spp image ->
        # the exchange -> rank 0 copies the combined rgb8 image to pinned host memory.  Timed from a barrier to a barrier.
        host8 = torch.empty(NY, NX, 3, dtype=torch.uint8).pin_memory() if rank == 0 else None
        t_e2e, h2d = 0.0, 0
        for k in range(-1, n_e2e):
            acc.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            s2 = rtb.Scene.deserialize(blob)
            s2.upload(w, local)
            o2 = rtb.default_opts(pipeline=pipe, device=local)
            set_shard(o2, 0)
            s2.render_device(cam, NX, NY, ns_total, o2, acc.data_ptr(), stream)
            exchange(ns_total)
            if rank == 0:
                host8.copy_(rgb8, non_blocking=True)
            torch.cuda.synchronize()
            dist.barrier()
            dt_call = time.perf_counter() - t0
            if rank == 0:
                print(f"[bench] e2e image {k}: {dt_call * 1e3:.1f} ms" + (" (warm-up, not counted)" if k < 0 else ""), file=sys.stderr)
            if k >= 0:
                t_e2e += dt_call
            h2d = s2.device_bytes() + 256
            del s2
        t = torch.tensor([t_e2e / n_e2e], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": world * paths_per_step_rank / float(t) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h2d) * world,
               "d2h_bytes_per_step": int(NX * NY * 3),
               "note": f"one {ns_total}-spp 4K image on rank 0's pinned host buffer per call: on every rank a fresh scene handle from host bytes "
                       "(deserialize + flatten + BVH build + H2D of the tables), its shard of the samples, the exchange step "
                       f"({exchange_info['mode']}), D2H of the combined rgb8 image on rank 0; host wall clock from a barrier to a barrier, "
                       "max over ranks, mean of the calls after one untimed warm-up call"}

    if rank != 0:
        if ex is not None:
            ex.close()
        if world > 1:
            dist.destroy_process_group()
        return
    ck = clocks.summary()
    peaks = measured_peaks()
    prop = torch.cuda.get_device_properties(local)
    sm_mhz = ck["sm_mhz"] or peaks.get("sm_max_mhz", 1965.0)
    nominal_peak = prop.multi_processor_count * 128 * sm_mhz * 1e6 / 1e12  # T lane-instr/s at the clock seen
    # calibration on this device (csrc/calib.cu): FFMA chains alone, and FFMA + LOP3 interleaved (FMA and ALU pipe both fed)
    fma_peak, mixed_peak, slab_peak = rtb.issue_peak(local)
    issue_peak = max(fma_peak, mixed_peak, slab_peak)
    achieved = work_per_step / (kern_ms / args.steps * 1e-3) / 1e12
    roofline = {
        "bound": "fp32_issue", "achieved": achieved, "peak": issue_peak, "unit": "Tlane-instr/s",
        "frac": achieved / issue_peak, "traffic": ncu_traffic_bytes(NX, NY, spp),
        "note": ("SURVEY.md §8(d): scene tables (<1 MB + 1.5 MB texture) stay in L1/L2, no dense contraction, so the bound is "
                 "SM instruction issue: peak = lane-instructions/s measured on this device by an FFMA issue microbenchmark "
                 "(rtb_issue_peak; nominal SMs x 128 lanes x sampled SM clock alongside); achieved = algorithmic "
                 "lane-instructions (fixed cost table x device op counters of one step; a medium's two boundary queries are "
                 "priced once, as primitive tests) / kernel time. "
                 "MEASURED_PEAKS.json has no FP32 entry. traffic = dram bytes read + written by one launch of this "
                 "workload (ncu --set full, profiles/r2_mega_hw_4k64.json): megabytes against the algorithmic image + scene tables, "
                 "i.e. HBM is idle. "
                 "hardware = the ncu view of the same launch shape (profiles/), not measured in this run."),
        "segments_per_path": st["segments"] / st["paths"], "node_tests_per_segment": st["node_tests"] / st["segments"],
        "prim_tests_per_segment": st["prim_tests"] / st["segments"],
        "lane_instr_per_segment_model": work_per_step / st["segments"],
        "Msegments_per_s": value * st["segments"] / st["paths"],
        "hbm_peak_gbs_measured": peaks.get("hbm_gbs"),
        "peak_nominal": nominal_peak, "peak_measured_fma_only": fma_peak, "peak_measured_fma_lop3": mixed_peak, "peak_measured_slab_mix": slab_peak,
        "hardware": ncu_hardware_view(NX, NY, spp),
    }
    cpu_v, cpu_desc, cores, _ = cpu_reference_sample(seconds=args.cpu_seconds) if args.cpu_seconds > 0 else (None, "skipped", 0, 0)
    # the reference's own threading model is ONE thread (tests/test.rs:28 `--test-threads=1`, Rc world src/bvh.rs:11-12)
    cpu1_v, cpu1_desc, _, _ = cpu_reference_sample(seconds=min(8.0, args.cpu_seconds), nthreads=1) if args.cpu_seconds > 0 else (None, "skipped", 0, 0)
    par = f"{'row-band' if bands else 'spp'}-shard x{world}"
    if world > 1:
        par += {"p2p": " + sharded P2P exchange (each GPU reduces 1/N of the image over NVLink, rgb8 slices land on rank 0)",
                "nccl": " + NCCL reduce_scatter(sum) + per-slice resolve + gather of rgb8",
                "p2p-root": " + fused P2P reduce+resolve on rank 0", "nccl-root": " + NCCL reduce(sum) to rank 0"}.get(mode, "")
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{SCENE} as written (tests/test.rs:419-523), {NX}x{NY}, {spp} spp per step per GPU, max_depth 50, "
                               f"black background; 10k-spp image = {10000 // spp} steps; L2 flushed (256 MB write) between timed steps",
                   "pipeline": args.pipeline, "seconds_per_4k_10kspp_image": NX * NY * 10000 / (value * 1e6),
                   "parallelism": par, "exchange_ms": xch_ms},
        "e2e": e2e, "gpu_launches": int(launches_per_step * args.steps + exchange_launches), "clocks": ck, "roofline": roofline,
        "cpu_baseline": {"value": cpu_v, "unit": UNIT, "cores": cores, "kind": "port", "sample": cpu_desc,
                         "single_thread": {"value": cpu1_v, "unit": UNIT, "cores": 1, "sample": cpu1_desc,
                                           "note": "the reference is single-threaded (tests/test.rs:28, src/bvh.rs:11-12): this is its own execution model"}},
    }
    if exchange_info:
        line["exchange"] = exchange_info
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    if ex is not None:
        ex.close()
    if world > 1:
        dist.destroy_process_group()


def ncu_hardware_view(nx, ny, spp):
    """The ncu hardware counters of one k_render_mega launch of exactly this shape, from the committed summary
    (profiles/r2_mega_hw_4k64.json, written by tools/ncu_summary.py from an `ncu --set full` capture); None otherwise."""
    f = ROOT / "profiles" / "r2_mega_hw_4k64.json"
    if (nx, ny, spp) != (3840, 2160, 64) or not f.exists():
        return None
    try:
        return json.loads(f.read_text())
    except Exception:
        return None


def ncu_traffic_bytes(nx, ny, spp):
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_render_mega launch, from the committed `ncu --set full` capture of
    exactly this launch shape (profiles/r2_mega_hw_4k64.json: 3840x2160, 64 spp); None for any other shape."""
    hw = ncu_hardware_view(nx, ny, spp)
    if not hw or hw.get("dram_bytes_read") is None or hw.get("dram_bytes_write") is None:
        return None
    return float(hw["dram_bytes_read"]) + float(hw["dram_bytes_write"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--reduce", default="auto", choices=["auto", "p2p", "nccl", "p2p-root", "nccl-root"],
                    help="N>1 exchange step: sharded over the ranks through peer memory (p2p; auto falls back to nccl), sharded over NCCL "
                         "(reduce_scatter + resolve + gather), or the r1 forms that pull everything to rank 0")
    ap.add_argument("--shard", default="spp", choices=["spp", "bands"], help="N>1 work split: samples per pixel (default) or 4-row image bands")
    ap.add_argument("--pipeline", default="auto", choices=["auto", "mega", "wavefront"])
    ap.add_argument("--spp-per-step", type=int, default=SPP_PER_STEP)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--configs-table", action="store_true", help="render all five BASELINE configs in full and print a markdown table")
    args = ap.parse_args()
    if args.configs_table:
        run_configs_table(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
