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


def test_two_gpu_peer_reduce_equals_single_gpu_image():
    if rtb.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29533", str(ROOT / "tests" / "peer_worker.py")], env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "peer reduce ok" in r.stdout


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
        for devs in (None, list(range(n_dev))[::-1], 2):
            b_sum, b8, sb = s.render_multi(w, cam, nx, ny, ns, devices=devs, opts=o, want_stats=True)
            # same sample set (counter-based draws), different float add order
            np.testing.assert_allclose(a_sum, b_sum, rtol=1e-4, atol=1e-4)
            assert np.abs(a8.astype(int) - b8.astype(int)).max() <= 1
            assert sb["paths"] == sa["paths"] and sb["segments"] == sa["segments"]
            assert sb["devices"] == (2 if devs == 2 else min(n_dev, ns))
    # more devices than samples: the surplus devices stay idle
    s, w, cam = rtb.Scene.named("two_spheres", 64, 36)
    one_sum, _, _ = s.render(w, cam, 64, 36, 1, rtb.default_opts(background=1))
    m_sum, _, st = s.render_multi(w, cam, 64, 36, 1, opts=rtb.default_opts(background=1), want_stats=True)
    np.testing.assert_array_equal(one_sum, m_sum)
    assert st["devices"] == 1
