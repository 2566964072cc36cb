"""Writes the parity table (profiles/r2_parity.md) from a run of the GPU parity tests ON THE GPU BOX:
    python tools/parity_report.py [pytest-log]  > gpurun_out/r2_parity.md
Without an argument it runs `pytest tests/test_gpu_parity.py tests/test_gpu_fields.py -m gpu -rP` itself.  Every figure is
what the tests printed (and asserted on): GPU f32 through the C ABI against the CPU f64 oracle, same scene bytes."""
import re
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
if len(sys.argv) > 1:
    log = Path(sys.argv[1]).read_text()
else:
    r = subprocess.run([sys.executable, "-m", "pytest", "tests/test_gpu_parity.py", "tests/test_gpu_fields.py", "-m", "gpu", "-q", "-rP"],
                       cwd=ROOT, capture_output=True, text=True)
    log = r.stdout + r.stderr
rows = {}
cases = {}
for line in log.splitlines():
    m = re.match(r"^([a-z_ ]+?) (\{.*\})\s*$", line)
    if not m:
        continue
    try:
        d = eval(m.group(2), {"np": np, "__builtins__": {}})
    except Exception:
        continue
    rows.setdefault(m.group(1), {}).update(d)
for m in re.finditer(r"test_image_matches_oracle_within_noise_floor\[([a-z_]+)-(\d)-(\d+)-(\d+)-(\d+)\]", log):
    cases[m.group(1)] = (int(m.group(3)), int(m.group(4)), int(m.group(5)), int(m.group(2)))
summary = [l for l in log.splitlines() if re.search(r"\d+ passed", l)]
print("# r2 parity table — GPU (f32, librtb200.so through the C ABI) vs CPU oracle (f64 restatement of src/*.rs)\n")
print("Written by `tools/parity_report.py` on the GPU box from the `-m gpu` tests' own output"
      + (f" (`{summary[-1].strip()}`)" if summary else "") + ".\n")
print("## Closest-hit primitive IDs (media off) and HitRecord fields\n")
print("primary = 2^20 jittered camera rays with lens and time samples; secondary = rays leaving the oracle's hit points in the four "
      "scatter families (normal + ball, mirror, through the surface, uniform ball), oracle with t_min = 0.001 "
      "(src/color.rs:7), GPU told which primitive the ray leaves (analytic self-hit exclusion).  `rehit self` = share of "
      "secondary rays whose reference answer is the primitive they left (far sphere root / another face).\n")
print("| scene | primary ID agree | secondary rays | secondary ID agree | rehit self | t rel err q99.9 (secondary) | p rel err q99.9 / max | normal err q99.9 | u / v err q99.9 |")
print("|---|---:|---:|---:|---:|---:|---:|---:|---:|")
for name, d in rows.items():
    if "secondary" not in d:
        continue
    print(f"| {name} | {100 * d.get('primary_agree', float('nan')):.4f} % | {d['secondary']} | {100 * d.get('secondary_agree', d['agree']):.4f} % | "
          f"{100 * d.get('rehit_self_ref', 0):.1f} % | {d.get('secondary_t_rel_q999', d.get('t_rel_q999', 0)):.1e} | {d.get('p_rel_q999', 0):.1e} / {d.get('p_rel_max', 0):.1e} | "
          f"{d.get('n_q999', 0):.1e} | {d.get('u_q999', 0):.1e} / {d.get('v_q999', 0):.1e} |")
print("\n## Images: PSNR on the gamma-encoded [0,1] image the reference writes (tests/test.rs:71-78)\n")
print("floor = oracle vs oracle (two seeds) at the same spp; GPU = GPU vs oracle at that spp; bias = GPU at 16x spp vs the mean of "
      "the two oracle renders (~ floor + 6 dB when unbiased); ratio = mean radiance GPU(16x) / oracle per channel, sigma its "
      "standard error.  Asserted: floor >= 40.5, GPU >= 40 (north_star), GPU >= floor - 0.5, bias >= floor + 4, |ratio - 1| <= 3 sigma + 5e-4.\n")
print("| scene | size | spp | background | floor dB | GPU dB | bias dB | mean ratio r/g/b | sigma |")
print("|---|---|---:|---|---:|---:|---:|---|---|")
for name, d in rows.items():
    if "floor_db" not in d:
        continue
    nx, ny, ns, bg = cases.get(name, (64, 64, 1024, 1))
    print(f"| {name} | {nx}x{ny} | {ns} | {'sky' if bg else 'black'} | {d['floor_db']:.2f} | {d['gpu_db']:.2f} | {d['bias_db']:.2f} | "
          + " / ".join(f"{x:.5f}" for x in d["mean_ratio"]) + " | " + " / ".join(f"{x:.5f}" for x in d["mean_ratio_sigma"]) + " |")
print("\n## Stage hooks (csrc/debug.cu)\n")
for name, d in rows.items():
    if "origin_err" in d:
        print(f"* camera `{name}`: origin err {d['origin_err']:.1e}, direction err {d['dir_err']:.1e} (relative), time err {d['time_err']:.1e} "
              "— device `camera_ray` vs the oracle's `get_ray` (src/camera.rs:53-67) on the same draws")
for key in ("ball", "disk"):
    for name, d in rows.items():
        if key in d and isinstance(d[key], dict):
            k = d[key]
            print(f"* sampler `{key}` (src/util.rs:4-24): KS vs analytic radial CDF {k['ks_vs_cdf']:.2e}, KS vs the oracle's rejection sampler "
                  f"{k['ks_vs_oracle']:.2e}, mean err {k['mean_err']:.1e}, second-moment err {k['second_moment_err']:.1e}")
for name, d in rows.items():
    if "p_scatter_ref" in d:
        extra = f", reflected fraction {d['reflect_frac_ref']:.4f} / {d['reflect_frac_dev']:.4f}" if "reflect_frac_ref" in d else ""
        print(f"* scatter `{name}` (last case printed): P(scatter) oracle {d['p_scatter_ref']:.4f} / device {d['p_scatter_dev']:.4f}{extra}")
