"""Writes tests/golden/jpeg_kat.json: small JPEG files (base64) with the SHA-256 of the RGB8 pixels
PIL (libjpeg-turbo) decodes them to — known answers for rtb_jpeg_decode that need no PIL at test time.
Also records the SHA-256 of the reference's texture/earthmap.jpg decode (== assets/earthmap_1024x512.rgb8).
Run here (needs PIL; /root/reference optional):  python tests/golden/make_jpeg_kat.py"""
import base64, hashlib, io, json
from pathlib import Path
import numpy as np
from PIL import Image

HERE = Path(__file__).resolve().parent
rng = np.random.default_rng(0xC0FFEE)
cases = []
for name, (w, h), ss, q, extra in [
    ("noise_444_q90", (24, 16), 0, 90, {}),
    ("noise_422_q75", (23, 17), 1, 75, {}),
    ("noise_420_q85_rst", (37, 29), 2, 85, {"restart_marker_blocks": 2}),
    ("ramp_420_q50", (40, 24), 2, 50, {}),
    ("grey_q80", (19, 11), None, 80, {}),
]:
    if name.startswith("ramp"):
        yy, xx = np.mgrid[0:h, 0:w]
        img = np.stack([(xx * 6) % 256, (yy * 10) % 256, (xx * yy) % 256], -1).astype(np.uint8)
    elif ss is None:
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    else:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    b = io.BytesIO()
    if ss is None:
        Image.fromarray(img, "L").save(b, "JPEG", quality=q)
    else:
        Image.fromarray(img).save(b, "JPEG", quality=q, subsampling=ss, **extra)
    data = b.getvalue()
    rgb = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
    cases.append({"name": name, "nx": w, "ny": h, "jpeg_b64": base64.b64encode(data).decode(),
                  "rgb_sha256": hashlib.sha256(rgb.tobytes()).hexdigest()})
out = {"decoder": "PIL %s (libjpeg-turbo)" % Image.__version__, "cases": cases}
ref = Path("/root/reference/texture/earthmap.jpg")
if ref.exists():
    rgb = np.asarray(Image.open(ref).convert("RGB"))
    out["earthmap"] = {"nx": rgb.shape[1], "ny": rgb.shape[0], "jpeg_sha256": hashlib.sha256(ref.read_bytes()).hexdigest(),
                       "rgb_sha256": hashlib.sha256(rgb.tobytes()).hexdigest()}
(HERE / "jpeg_kat.json").write_text(json.dumps(out, indent=1))
print("wrote", HERE / "jpeg_kat.json", sum(len(c["jpeg_b64"]) for c in cases), "b64 bytes")
