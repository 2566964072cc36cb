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


def image_parity(gpu_mean, gpu_hi_mean, orc_a, orc_b, hi_factor=16.0):
    """Noise-floor-relative comparison of two Monte-Carlo estimates.

    gpu_mean / orc_a / orc_b: per-pixel linear means at the SAME spp (orc_a, orc_b: oracle with two
    seeds); gpu_hi_mean: GPU at >= 16x that spp.  Returns a dict with
      floor_db   PSNR(orc_a, orc_b)                 — what two unbiased estimates score
      gpu_db     PSNR(gpu, orc_a)                   — must be within 0.5 dB of floor_db
      bias_db    PSNR(gpu_hi, mean(orc_a, orc_b))   — ~ floor + 6 dB if the GPU is unbiased
      mean_ratio per-channel mean radiance ratio gpu_hi / oracle, and mean_ratio_sigma its standard error
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
    # 1-sigma of that ratio from the data: per pixel, (a - b)^2 / 2 estimates the variance of ONE oracle render; the
    # oracle mean o = (a + b)/2 carries half of it, the GPU render at hi_factor x the samples 1/hi_factor of it
    npix = orc_a.shape[0] * orc_a.shape[1]
    out["mean_ratio_sigma"] = []
    for c in range(3):
        var_one = float(np.sum((orc_a[..., c] - orc_b[..., c]) ** 2) / 2.0) / (npix * npix)
        out["mean_ratio_sigma"].append(float(np.sqrt(var_one * (0.5 + 1.0 / hi_factor)) / max(o[..., c].mean(), 1e-30)))
    return out
