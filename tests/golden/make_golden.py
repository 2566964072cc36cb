"""Generates tests/golden/kat_ppm.json from the reference's two committed renders
(/root/reference/output/final_scene.ppm, cornell_smoke.ppm) — the only known-answer vectors the
reference ships.  Run in the build container, where /root/reference exists:

    python tests/golden/make_golden.py

The fixture stores, per file: byte length, sha256 of the whole P3 text, the header, and the
histogram of pixel lines, so tests can compare a freshly rendered P3 string byte-for-byte (via
the hash) without carrying 3.8 MB of "0 0 0" lines in the repo.
"""
import collections
import hashlib
import json
from pathlib import Path

REF = Path("/root/reference/output")
OUT = Path(__file__).resolve().parent / "kat_ppm.json"

# scene / camera / size / spp each file was rendered with: tests/test.rs:763-838
CONFIG = {
    "final_scene.ppm": {"scene": "final_scene", "nx": 800, "ny": 800, "ns": 100, "test": "tests/test.rs:802-838"},
    "cornell_smoke.ppm": {"scene": "cornell_smoke", "nx": 800, "ny": 800, "ns": 100, "test": "tests/test.rs:763-799"},
}


def main():
    out = {}
    for name, cfg in CONFIG.items():
        data = (REF / name).read_bytes()
        lines = data.split(b"\n")
        header = b"\n".join(lines[:3]).decode() + "\n"
        hist = collections.Counter(l.decode() for l in lines[3:] if l)
        out[name] = dict(cfg, bytes=len(data), sha256=hashlib.sha256(data).hexdigest(), header=header,
                         pixel_line_histogram=dict(hist))
    OUT.write_text(json.dumps(out, indent=1, sort_keys=True) + "\n")
    print(OUT.read_text())


if __name__ == "__main__":
    main()
