"""Helpers shared by the GPU parity tests and bench.py's parity report."""
import numpy as np


def gamma01(linear_mean):
    """tests/test.rs:73-75: sqrt then clamp to [0,1] (the value that gets quantised)."""
    return np.clip(np.sqrt(np.maximum(linear_mean, 0.0)), 0.0, 1.0)


def psnr(a, b, peak=1.0):
    mse = float(np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2))
    if mse == 0.0:
        return float("inf")
    return 10.0 * np.log10(peak * peak / mse)


def image_parity(gpu_mean, gpu_hi_mean, orc_a, orc_b):
    """Noise-floor-relative comparison of two Monte-Carlo estimates.

    gpu_mean / orc_a / orc_b: per-pixel linear means at the SAME spp (orc_a, orc_b: oracle with two
    seeds); gpu_hi_mean: GPU at >= 16x that spp.  Returns a dict with
      floor_db   PSNR(orc_a, orc_b)                 — what two unbiased estimates score
      gpu_db     PSNR(gpu, orc_a)                   — must be within 0.5 dB of floor_db
      bias_db    PSNR(gpu_hi, mean(orc_a, orc_b))   — ~ floor + 6 dB if the GPU is unbiased
      mean_ratio per-channel mean radiance ratio gpu_hi / oracle
    all on the gamma-encoded [0,1] image the reference writes."""
    ga, gb = gamma01(orc_a), gamma01(orc_b)
    gab = gamma01(0.5 * (orc_a + orc_b))
    out = {
        "floor_db": psnr(ga, gb),
        "gpu_db": psnr(gamma01(gpu_mean), ga),
        "bias_db": psnr(gamma01(gpu_hi_mean), gab),
    }
    o = 0.5 * (orc_a + orc_b)
    out["mean_ratio"] = [float(gpu_hi_mean[..., c].mean() / max(o[..., c].mean(), 1e-30)) for c in range(3)]
    return out
