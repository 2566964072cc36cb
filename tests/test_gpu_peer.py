"""The exchange step (csrc/peer.cu): fused reduce + resolve over several accumulation buffers.
Single-GPU cases use several local buffers (same kernel, same C-ABI call); the 2-GPU case maps a
peer rank's buffer over CUDA IPC and is skipped on a box with one GPU."""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

import raytracing_rust_b200 as rtb
from raytracing_rust_b200 import distributed as D

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _host_quantise(total, ns):
    return D.resolve_host(total, ns)


@pytest.mark.parametrize("nx,ny,n_bufs", [(64, 48, 1), (64, 48, 3), (37, 19, 2), (5, 3, 8), (256, 128, 16)])
def test_reduce_resolve_matches_host(nx, ny, n_bufs):
    import torch

    rng = np.random.default_rng(nx * 1000 + n_bufs)
    ns = 16 * n_bufs
    parts = [rng.random((ny, nx, 3), dtype=np.float32) * (2.0 * ns / n_bufs) for _ in range(n_bufs)]
    parts[0][0, 0] = [np.nan, -3.0, 1e9]  # NaN -> 0, negative -> 0 (sqrt NaN), overflow -> 255
    bufs = [rtb.DeviceBuffer((ny, nx, 3)) for _ in range(n_bufs)]
    views = [torch.as_tensor(b, device="cuda") for b in bufs]
    for v, p in zip(views, parts):
        v.copy_(torch.from_numpy(p))
    rgb8 = torch.empty(ny, nx, 3, dtype=torch.uint8, device="cuda")
    total = torch.empty(ny, nx, 3, dtype=torch.float32, device="cuda")
    rtb.reduce_resolve_device([b.ptr for b in bufs], nx, ny, ns, rgb8.data_ptr(), total.data_ptr())
    torch.cuda.synchronize()
    exp_total = parts[0].copy()
    for p in parts[1:]:
        exp_total = exp_total + p  # same order, f32
    np.testing.assert_array_equal(total.cpu().numpy(), exp_total)  # bit-exact: fixed summation order
    # the device quantises in f32 (sqrtf(sum * (1/ns))), the host check in f64: at most 1 LSB apart, rarely
    dq = np.abs(rgb8.cpu().numpy().astype(int) - _host_quantise(exp_total, ns).astype(int))
    assert dq.max() <= 1 and (dq > 0).mean() < 1e-3
    # and it agrees with the two-step path (sum, then rtb_resolve_device)
    rgb8_b = torch.empty_like(rgb8)
    rtb.resolve_device(total.data_ptr(), nx, ny, ns, rgb8_b.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(rgb8, rgb8_b)
    del views
    for b in bufs:
        b.free()


def test_reduce_resolve_rejects_bad_arguments():
    b = rtb.DeviceBuffer((4, 4, 3))
    with pytest.raises(rtb.RtbError):
        rtb.reduce_resolve_device([], 4, 4, 1, b.ptr)
    with pytest.raises(rtb.RtbError):
        rtb.reduce_resolve_device([b.ptr] * 17, 4, 4, 1, b.ptr)
    b.free()


# ---- the sharded exchange and row-band sharding, on one GPU (several local buffers stand in for the ranks) ---------------
@pytest.mark.parametrize("nx,ny,n_bufs", [(64, 48, 3), (37, 19, 2), (5, 3, 8), (256, 128, 8)])
def test_reduce_resolve_range_slices_equal_the_whole_image(nx, ny, n_bufs):
    import torch

    rng = np.random.default_rng(nx + n_bufs)
    ns = 8 * n_bufs
    bufs = [rtb.DeviceBuffer((ny, nx, 3)) for _ in range(n_bufs)]
    views = [torch.as_tensor(b, device="cuda") for b in bufs]
    for v in views:
        v.copy_(torch.from_numpy(rng.random((ny, nx, 3), dtype=np.float32) * (2.0 * ns / n_bufs)))
    ptrs = [b.ptr for b in bufs]
    whole8 = torch.empty(ny, nx, 3, dtype=torch.uint8, device="cuda")
    whole = torch.empty(ny, nx, 3, dtype=torch.float32, device="cuda")
    rtb.reduce_resolve_device(ptrs, nx, ny, ns, whole8.data_ptr(), whole.data_ptr())
    n = 3 * nx * ny
    for world in (1, 2, 3, 8):
        per = (((n + world - 1) // world) + 3) & ~3
        out8 = torch.zeros_like(whole8)
        out = torch.zeros_like(whole)
        for g in range(world):  # what rank g of `world` does
            first = min(n, g * per)
            rtb.reduce_resolve_range_device(ptrs, first, min(n - first, per), ns, out8.data_ptr(), out.data_ptr())
        torch.cuda.synchronize()
        assert torch.equal(out8, whole8) and torch.equal(out, whole), world
    del views
    for b in bufs:
        b.free()


@pytest.mark.parametrize("name,bg,nx,ny", [("cornell_box", 0, 96, 64), ("final_scene_lit", 0, 203, 117), ("two_spheres", 1, 64, 5)])
def test_row_band_shards_tile_the_image(name, bg, nx, ny):
    # the "image tiles" split of the pixel loop (tests/test.rs:62-63): shard g renders every sample of the 4-row bands
    # g, g+G, ...; the shards are disjoint and their union is the single-GPU image (same per-pixel Philox counters)
    import torch

    ns = 6
    s, w, cam = rtb.Scene.named(name, nx, ny)
    full, full8, st_full = s.render(w, cam, nx, ny, ns, rtb.default_opts(seed=4, background=bg, collect_stats=1), want_stats=True)
    for world in (2, 3, 8):
        acc = np.zeros_like(full)
        seg = paths = 0
        img8 = torch.zeros(ny, nx, 3, dtype=torch.uint8, device="cuda")
        for g in range(world):
            part, _, st = s.render(w, cam, nx, ny, ns, rtb.default_opts(seed=4, background=bg, band_begin=g, band_stride=world, collect_stats=1),
                                   want_stats=True)
            rows = D.shard_bands(ny, g, world)
            other = np.setdiff1d(np.arange(ny), rows)
            assert not part[other].any()                    # nothing outside the shard's bands
            assert st["paths"] == len(rows) * nx * ns
            acc += part
            seg, paths = seg + st["segments"], paths + st["paths"]
            d_part = torch.from_numpy(part).cuda()
            rtb.resolve_bands_device(d_part.data_ptr(), nx, ny, ns, g, world, img8.data_ptr(), compact=False)
            packed = torch.zeros(len(rows) * nx * 3 + 16, dtype=torch.uint8, device="cuda")
            rtb.resolve_bands_device(d_part.data_ptr(), nx, ny, ns, g, world, packed.data_ptr(), compact=True)
            torch.cuda.synchronize()
            if rows:
                assert torch.equal(packed[: len(rows) * nx * 3].view(len(rows), nx, 3).cpu(), img8[rows].cpu())
        np.testing.assert_allclose(acc, full, rtol=2e-5, atol=1e-5)
        assert paths == st_full["paths"] and seg == st_full["segments"]
        d8 = np.abs(img8.cpu().numpy().astype(int) - full8.astype(int))
        assert d8.max() <= 1 and (d8 > 0).mean() < 1e-3
    with pytest.raises(rtb.RtbError):
        s.render(w, cam, nx, ny, ns, rtb.default_opts(band_begin=1, band_stride=2, pipeline=rtb.PIPE_WAVEFRONT))


@pytest.mark.parametrize("mode", ["p2p", "nccl"])
def test_sharded_exchange_single_rank_is_plain_resolve(mode):
    import torch

    nx, ny, ns = 75, 41, 9  # 3*nx*ny not a multiple of 4: the scalar tail path
    s, w, cam = rtb.Scene.named("cornell_box", nx, ny)
    s.upload(w, 0)
    ex = D.ShardedExchange(nx, ny, 0, mode=mode)
    D.render_sharded(s, cam, nx, ny, ns, rtb.default_opts(seed=2), ex.acc, 0, 1)
    img = ex.reduce_resolve(ns).clone()
    ref = torch.empty_like(img)
    rtb.resolve_device(ex.acc.data_ptr(), nx, ny, ns, ref.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(img, ref)
    ex.acc.zero_()
    D.render_band_sharded(s, cam, nx, ny, ns, rtb.default_opts(seed=2), ex.acc, 0, 1)
    img2 = ex.gather_bands(ns)
    torch.cuda.synchronize()
    d = (img2.int() - ref.int()).abs()
    assert int(d.max()) <= 1
    ex.close()


def test_two_gpu_peer_reduce_equals_single_gpu_image():
    if rtb.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29533", str(ROOT / "tests" / "peer_worker.py")], env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "peer reduce ok" in r.stdout and "sharded exchange ok" in r.stdout


# ---- rtb_render_multi: create_image on several GPUs from one host call (csrc/multi.cu) --------------------------
def test_render_multi_on_one_device_is_render():
    nx, ny, ns = 96, 64, 12
    s, w, cam = rtb.Scene.named("cornell_box", nx, ny)
    a_sum, a8, sa = s.render(w, cam, nx, ny, ns, rtb.default_opts(seed=5), want_stats=True)
    b_sum, b8, sb = s.render_multi(w, cam, nx, ny, ns, devices=[0], opts=rtb.default_opts(seed=5), want_stats=True)
    # the same paths; several warps add into one pixel of so small an image, in no fixed order (1-ulp differences)
    np.testing.assert_allclose(a_sum, b_sum, rtol=2e-6, atol=1e-6)
    d8 = np.abs(a8.astype(int) - b8.astype(int))
    assert d8.max() <= 1 and (d8 > 0).mean() < 1e-3
    assert sb["devices"] == 1 and sb["paths"] == sa["paths"] == nx * ny * ns and sb["segments"] == sa["segments"]
    assert sb["kernel_launches"] == sa["kernel_launches"] + 1
    c_sum, c8, sc = s.render_multi(w, cam, nx, ny, ns, devices=[0], opts=rtb.default_opts(seed=5, multi_shard=rtb.SHARD_BANDS), want_stats=True)
    np.testing.assert_allclose(a_sum, c_sum, rtol=2e-6, atol=1e-6)
    assert np.abs(a8.astype(int) - c8.astype(int)).max() <= 1 and sc["paths"] == sa["paths"]
    with pytest.raises(rtb.RtbError):
        s.render_multi(w, cam, nx, ny, ns, devices=[0, 0])
    with pytest.raises(rtb.RtbError):
        s.render_multi(w, cam, nx, ny, ns, devices=[0, 99])
    with pytest.raises(rtb.RtbError):
        s.render_multi(w, cam, nx, ny, ns, devices=[0], opts=rtb.default_opts(sample_begin=1, sample_stride=2))


def test_render_multi_all_devices_equals_single_gpu_image():
    n_dev = rtb.device_count()
    if n_dev < 2:
        pytest.skip("needs >= 2 GPUs")
    nx, ny, ns = 203, 117, 21  # ragged everything: ns not a multiple of the device count
    for name, bg in (("final_scene_lit", 0), ("random_scene", 1)):
        s, w, cam = rtb.Scene.named(name, nx, ny)
        o = rtb.default_opts(seed=11, background=bg)
        a_sum, a8, sa = s.render(w, cam, nx, ny, ns, o, want_stats=True)
        for devs, shard in ((None, rtb.SHARD_SAMPLES), (list(range(n_dev))[::-1], rtb.SHARD_SAMPLES), (2, rtb.SHARD_SAMPLES),
                            (None, rtb.SHARD_BANDS), (2, rtb.SHARD_BANDS)):
            o.multi_shard = shard
            b_sum, b8, sb = s.render_multi(w, cam, nx, ny, ns, devices=devs, opts=o, want_stats=True)
            # same sample set (counter-based draws), different float add order
            np.testing.assert_allclose(a_sum, b_sum, rtol=1e-4, atol=1e-4)
            assert np.abs(a8.astype(int) - b8.astype(int)).max() <= 1
            assert sb["paths"] == sa["paths"] and sb["segments"] == sa["segments"]
            assert sb["devices"] == (2 if devs == 2 else min(n_dev, ns if shard == rtb.SHARD_SAMPLES else (ny + 3) // 4))
    # more devices than samples: the surplus devices stay idle
    s, w, cam = rtb.Scene.named("two_spheres", 64, 36)
    one_sum, _, _ = s.render(w, cam, 64, 36, 1, rtb.default_opts(background=1))
    m_sum, _, st = s.render_multi(w, cam, 64, 36, 1, opts=rtb.default_opts(background=1), want_stats=True)
    np.testing.assert_array_equal(one_sum, m_sum)
    assert st["devices"] == 1
