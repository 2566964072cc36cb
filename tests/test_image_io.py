"""Image formats either side of the hot path: the JPEG input of
tests/test.rs:201-205,489-493 (`image::open("texture/earthmap.jpg").to_rgb8()`) and PNG output.
CPU only.  The decoder is pinned (a) on committed known answers made with libjpeg-turbo through
PIL (tests/golden/jpeg_kat.json, generator beside it), (b) on the staged decode of the reference's
earthmap.jpg, (c) live against PIL over sizes x chroma sampling x quality x restart intervals."""
import base64, hashlib, io, json, zlib
from pathlib import Path

import numpy as np
import pytest

import raytracing_rust_b200 as rtb

GOLD = json.loads((Path(__file__).parent / "golden" / "jpeg_kat.json").read_text())
REF_JPG = Path("/root/reference/texture/earthmap.jpg")


@pytest.mark.parametrize("case", GOLD["cases"], ids=[c["name"] for c in GOLD["cases"]])
def test_jpeg_known_answers(case):
    rgb = rtb.decode_jpeg(base64.b64decode(case["jpeg_b64"]))
    assert rgb.shape == (case["ny"], case["nx"], 3)
    assert hashlib.sha256(rgb.tobytes()).hexdigest() == case["rgb_sha256"]


def test_staged_earthmap_is_the_known_decode():
    # assets/earthmap_1024x512.rgb8 (what the scenes use on the GPU box) has the recorded hash
    e = GOLD["earthmap"]
    a = rtb.earth_rgb8()
    assert a.shape == (e["ny"], e["nx"], 3)
    assert hashlib.sha256(a.tobytes()).hexdigest() == e["rgb_sha256"]


@pytest.mark.skipif(not REF_JPG.exists(), reason="reference tree not present on this machine")
def test_earthmap_jpg_decodes_to_the_staged_asset(tmp_path):
    assert hashlib.sha256(REF_JPG.read_bytes()).hexdigest() == GOLD["earthmap"]["jpeg_sha256"]
    assert np.array_equal(rtb.decode_jpeg(REF_JPG), rtb.earth_rgb8())
    # ImageTexture::new(image::open(..).to_rgb8().into_raw(), nx, ny) in one call gives the same scene bytes
    s1, s2 = rtb.Scene(), rtb.Scene()
    s1.image_file(REF_JPG)
    s2.image(rtb.earth_rgb8())
    assert s1.serialize() == s2.serialize()


def test_jpeg_matches_libjpeg_turbo_live():
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(11)
    n = 0
    for (w, h) in [(1, 1), (2, 3), (3, 5), (4, 4), (8, 8), (17, 13), (33, 31), (64, 48), (250, 129)]:
        yy, xx = np.mgrid[0:h, 0:w]
        ramp = np.stack([(xx * 7 + yy * 3) % 256, (xx * yy) % 256, (255 - xx * 2 - yy) % 256], -1).astype(np.uint8)
        for img in (ramp, rng.integers(0, 256, (h, w, 3), dtype=np.uint8)):
            for ss in (0, 1, 2):
                for q, extra in ((30, {}), (90, {"restart_marker_blocks": 3}), (100, {})):
                    b = io.BytesIO()
                    Image.fromarray(img).save(b, "JPEG", quality=q, subsampling=ss, **extra)
                    want = np.asarray(Image.open(io.BytesIO(b.getvalue())).convert("RGB"))
                    got = rtb.decode_jpeg(b.getvalue())
                    assert np.array_equal(got, want), (w, h, ss, q, extra)
                    n += 1
    assert n == 9 * 2 * 3 * 3


def test_jpeg_errors_are_reported_not_fatal(tmp_path):
    data = base64.b64decode(GOLD["cases"][2]["jpeg_b64"])
    for bad in (b"", b"\xff\xd8", b"not a jpeg at all", data[:40], data[: len(data) // 2][:-3] + b"\xff\xd9"[:0]):
        try:
            rtb.decode_jpeg(bad)  # a truncated scan may still decode (zero-padded), never crash
        except rtb.RtbError as e:
            assert e.code == -1
    with pytest.raises(rtb.RtbError) as ei:
        rtb.decode_jpeg(tmp_path / "missing.jpg")
    assert ei.value.code == -5  # RTB_ERR_IO
    # seeded mutations: error or image, never a crash
    rng = np.random.default_rng(5)
    for _ in range(300):
        d = bytearray(data)
        for _ in range(int(rng.integers(1, 5))):
            d[int(rng.integers(0, len(d)))] = int(rng.integers(0, 256))
        try:
            rtb.decode_jpeg(bytes(d))
        except rtb.RtbError:
            pass
    Image = pytest.importorskip("PIL.Image")
    b = io.BytesIO()
    Image.fromarray(np.zeros((16, 16, 3), np.uint8)).save(b, "JPEG", progressive=True)
    with pytest.raises(rtb.RtbError, match="progressive"):
        rtb.decode_jpeg(b.getvalue())


def _png_pixels(raw: bytes):
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, hdr = 8, b"", None
    while pos < len(raw):
        n = int.from_bytes(raw[pos:pos + 4], "big")
        typ, body = raw[pos + 4:pos + 8], raw[pos + 8:pos + 8 + n]
        assert zlib.crc32(typ + body) == int.from_bytes(raw[pos + 8 + n:pos + 12 + n], "big")
        if typ == b"IHDR":
            hdr = body
        if typ == b"IDAT":
            idat += body
        pos += 12 + n
    nx, ny = int.from_bytes(hdr[:4], "big"), int.from_bytes(hdr[4:8], "big")
    assert hdr[8:] == bytes([8, 2, 0, 0, 0])
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(ny, 1 + 3 * nx)
    assert (rows[:, 0] == 0).all()
    return rows[:, 1:].reshape(ny, nx, 3)


@pytest.mark.parametrize("shape", [(1, 1), (5, 7), (225, 400), (300, 257)])
def test_png_round_trip(tmp_path, shape):
    rng = np.random.default_rng(shape[0])
    img = rng.integers(0, 256, shape + (3,), dtype=np.uint8)
    p = tmp_path / "o.png"
    rtb.write_png(p, img)
    assert np.array_equal(_png_pixels(p.read_bytes()), img)  # chunk CRCs, zlib stream and Adler-32 check out
    try:
        from PIL import Image
    except ImportError:
        return
    assert np.array_equal(np.asarray(Image.open(p).convert("RGB")), img)
