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
