"""Writes tests/golden/final_scene.rtscene.head: the serialized final_scene (tests/test.rs:419-523 as written, scene seed 0x5EED)
WITHOUT its trailing image pool, which is byte for byte assets/earthmap_1024x512.rgb8.  bench.py --impl reference
rebuilds the scene bytes from these two files and hands them to the oracle, so the reference arm never loads librtb200.so.
tests/test_oracle.py checks the fixture still equals what rtb_scene_named serializes.  usage: python tools/make_reference_fixture.py"""
import struct
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import raytracing_rust_b200 as rtb

s, w, cam = rtb.Scene.named("final_scene", 480, 270)
blob = s.serialize()
n_img = struct.unpack_from("<Q", blob, 32)[0]  # RtbSceneHeader.n_image_bytes
assert blob[-n_img:] == rtb.earth_rgb8().tobytes()
out = ROOT / "tests" / "golden" / "final_scene.rtscene.head"
out.write_bytes(blob[:-n_img])
print(out, len(blob) - n_img, "bytes; world handle", w)
