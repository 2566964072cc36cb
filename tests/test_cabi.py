"""C-ABI boundary: the library loads without a GPU, exports every symbol include/rtb200.h
declares, and mirrors the reference's error behaviour (None / panic -> error codes).  No compute
calls here."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    text = (ROOT / "include" / "rtb200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rtb_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(rtb):
    syms = declared_symbols()
    assert len(syms) >= 35
    missing = [s for s in syms if not hasattr(rtb.lib, s)]
    assert not missing, missing


def test_header_compiles_as_c(tmp_path):
    import subprocess
    src = tmp_path / "t.c"
    src.write_text('#include "rtb200.h"\nint main(void){ rtb_render_opts o; (void)o; return sizeof(RtbObjRec) == 112 ? 0 : 1; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", str(ROOT / "include"), "-c", str(src), "-o",
                    str(tmp_path / "t.o")], check=True)


def build_cpp_scene_test(tmp_path):
    """tests/cpp/test_scenes.cpp: the reference's #[test] bodies against include/rtb200.hpp, linked to librtb200.so"""
    import subprocess
    exe = tmp_path / "test_scenes"
    libdir = ROOT / "raytracing_rust_b200"
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-I", str(ROOT / "include"),
                    str(ROOT / "tests" / "cpp" / "test_scenes.cpp"), "-o", str(exe), "-L", str(libdir), "-lrtb200",
                    f"-Wl,-rpath,{libdir}"], check=True)
    return exe


def test_cpp_host_api_builds_and_fails_loudly_without_a_gpu(rtb, tmp_path):
    import subprocess
    exe = build_cpp_scene_test(tmp_path)
    if rtb.device_count() > 0:
        pytest.skip("a GPU is present (the run is tests/test_gpu_parity.py::test_cpp_host_api_scenes)")
    r = subprocess.run([str(exe), str(tmp_path), "8", "8", "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "no CUDA device" in r.stderr  # no CPU fallback behind create_image


def test_struct_layouts_match_ctypes(rtb):
    # sizes the Python mirror assumes (include/rtb200.h, rtb_scene_format.h)
    assert C.sizeof(rtb.RtbCamera) == 15 * 8
    assert C.sizeof(rtb.RenderOpts) == 96
    assert C.sizeof(rtb.Stats) == (5 + 5 + 4 + 1) * 8 + 8 + 8 * 8
    o = rtb.default_opts()
    assert (o.max_depth, o.t_min, o.background, o.sample_stride) == (50, 1e-3, 0, 1)  # src/color.rs:7,9


def test_bad_handles_are_errors_not_crashes(rtb):
    s = rtb.Scene()
    with pytest.raises(rtb.RtbError):
        s.lambertian(3)  # no such texture
    t = s.solid(0.5, 0.5, 0.5)
    with pytest.raises(rtb.RtbError):
        s.sphere((0, 0, 0), 1.0, 7)  # no such material
    m = s.lambertian(t)
    with pytest.raises(rtb.RtbError):
        s.flip_normals(0)  # no objects yet
    sp = s.sphere((0, 0, 0), 1.0, m)
    with pytest.raises(rtb.RtbError):
        s.list([sp, 99])
    with pytest.raises(rtb.RtbError):
        s.rect(5, 0, 0, 1, 1, 0, m)  # bad plane


def test_empty_bvh_is_the_reference_panic(rtb):
    # BVHNode::new on an empty slice panics (src/bvh.rs:43-51) -> RTB_ERR_NO_BBOX
    s = rtb.Scene()
    with pytest.raises(rtb.RtbError) as e:
        s.bvh([])
    assert e.value.code == -4


def test_serialize_roundtrip_is_identical(rtb):
    for name in ("cornell_smoke", "two_perlin_sphere", "random_scene"):
        s, w, _ = rtb.Scene.named(name, 32, 18)
        b = s.serialize()
        s2 = rtb.Scene.deserialize(b)
        assert s2.serialize() == b
        assert s2.prim_count(w) == s.prim_count(w)
    with pytest.raises(rtb.RtbError):
        rtb.Scene.deserialize(b[:100])
    bad = bytearray(b)
    bad[0] ^= 0xFF
    with pytest.raises(rtb.RtbError):
        rtb.Scene.deserialize(bytes(bad))


def test_named_scene_inventory(rtb):
    # primitive totals of tests/test.rs:89-523
    expect = {"two_spheres": 2, "two_perlin_sphere": 2, "earth": 1, "simple_light": 4, "cornell_box": 6 + 12,
              "cornell_smoke": 6 + 12, "final_scene": 2400 + 1 + 5 + 2 + 1 + 1000}
    for name, n in expect.items():
        s, w, cam = rtb.Scene.named(name, 64, 64)
        assert s.prim_count(w) == n, name
    s, w, cam = rtb.Scene.named("random_scene", 1280, 720)
    n = s.prim_count(w)
    assert 380 <= n <= 404  # 400 grid cells minus those near (4,0.2,0), plus ground + 3 (F10)
    assert abs(cam.aspect - 1280 / 720) < 1e-15 and cam.vertical_fov == 20.0 and cam.aperture == 0.1
    with pytest.raises(rtb.RtbError):
        rtb.Scene.named("no_such_scene", 8, 8)
    # seeds are reproducible and distinct
    a = rtb.Scene.named("random_scene", 8, 8, seed=7)[0].serialize()
    b = rtb.Scene.named("random_scene", 8, 8, seed=7)[0].serialize()
    c = rtb.Scene.named("random_scene", 8, 8, seed=8)[0].serialize()
    assert a == b and a != c


def test_ppm_p3_text_matches_reference_format(rtb):
    # tests/test.rs:59 header and :79 "{} {} {}\n" per pixel
    img = np.array([[[0, 1, 2], [10, 99, 100]], [[255, 254, 7], [0, 0, 0]]], np.uint8)
    assert rtb.ppm_p3(img) == b"P3\n2 2\n255\n0 1 2\n10 99 100\n255 254 7\n0 0 0\n"


def test_ppm_p6_and_file_writers(rtb, tmp_path):
    img = np.array([[[0, 1, 2], [10, 99, 100]], [[255, 254, 7], [0, 0, 0]]], np.uint8)
    assert rtb.ppm_p6(img) == b"P6\n2 2\n255\n" + img.tobytes()
    rtb.write_ppm(tmp_path / "a.ppm", img)
    assert (tmp_path / "a.ppm").read_bytes() == rtb.ppm_p3(img)
    rtb.write_ppm(tmp_path / "b.ppm", img, binary=True)
    assert (tmp_path / "b.ppm").read_bytes() == rtb.ppm_p6(img)
    with pytest.raises(rtb.RtbError):  # File::create(..).expect("REASON") upstream (tests/test.rs:536)
        rtb.write_ppm(tmp_path / "no_such_dir" / "c.ppm", img)
    # the reference's committed all-black 800x800 renders are exactly this P3 text of a zero image
    black = rtb.ppm_p3(np.zeros((800, 800, 3), np.uint8))
    assert len(black) == 3_840_015 and black.startswith(b"P3\n800 800\n255\n0 0 0\n")


def test_rtscene_file_roundtrip(rtb, tmp_path):
    s, w, _ = rtb.Scene.named("final_scene", 32, 32)
    s.save(tmp_path / "final.rtscene")
    assert (tmp_path / "final.rtscene").read_bytes() == s.serialize()
    s2 = rtb.Scene.load(tmp_path / "final.rtscene")
    assert s2.serialize() == s.serialize() and s2.prim_count(w) == s.prim_count(w)
    with pytest.raises(rtb.RtbError):
        rtb.Scene.load(tmp_path / "missing.rtscene")
    (tmp_path / "bad.rtscene").write_bytes(b"not a scene")
    with pytest.raises(rtb.RtbError):
        rtb.Scene.load(tmp_path / "bad.rtscene")


def test_no_device_is_a_loud_error(rtb):
    if rtb.device_count() > 0:
        pytest.skip("a GPU is present")
    s, w, cam = rtb.Scene.named("two_spheres", 8, 8)
    with pytest.raises(rtb.RtbError) as e:
        s.render(w, cam, 8, 8, 1)
    assert e.value.code == -2  # RTB_ERR_NO_DEVICE: no CPU fallback
    with pytest.raises(rtb.RtbError) as e:
        s.render_multi(w, cam, 8, 8, 1)
    assert e.value.code == -2


def test_primary_rays_are_deterministic_and_in_frustum(rtb):
    _, _, cam = rtb.Scene.named("cornell_box", 16, 16)
    a = rtb.primary_rays(cam, 16, 16, sample=3, seed=11)
    b = rtb.primary_rays(cam, 16, 16, sample=3, seed=11)
    c = rtb.primary_rays(cam, 16, 16, sample=4, seed=11)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert np.all((a[:, 6] >= 0) & (a[:, 6] < 1))  # time in [t0, t1)  (src/camera.rs:61)
    # lens offsets stay inside aperture/2 of the eye (src/camera.rs:57-59)
    assert np.all(np.linalg.norm(a[:, :3] - np.array([278, 278, -800.0]), axis=1) <= 0.05 + 1e-12)


def test_philox_known_answers(rtb):
    # Philox4x32-10 known-answer vectors published with Random123 (kat_vectors: zeros, all ones, digits of pi);
    # every draw of the renderer is one of these blocks cut into five 24-bit uniforms (csrc/philox.h)
    f = 0xFFFFFFFF
    kat = [((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
           ((f, f, f, f), (f, f), (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
           ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
            (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1))]
    for ctr, key, want in kat:
        assert rtb.philox4x32_10(ctr, key) == want


def test_deserialize_rejects_hostile_scene_files(rtb):
    """rtb_scene_deserialize / rtb_scene_load take arbitrary bytes (include/rtb_scene_format.h): every malformed record that
    could hang or fault a later stage is refused here — self-referencing or forward-referencing checkers (flatten's
    tex_needs_uv recursion, the device's texture_value loop), unknown texture / material types (the device falls through to
    the image branch), bad Rect plane / Rotate axis selectors, image extents outside the pool, counts that wrap the size sum."""
    import struct
    s = rtb.Scene()
    t_a, t_b = s.solid(0.1, 0.2, 0.3), s.solid(0.9, 0.9, 0.9)
    t_c = s.checker(t_a, t_b)
    t_i = s.image(np.zeros((4, 4, 3), np.uint8))
    m = s.lambertian(t_c)
    m2 = s.lambertian(t_i)
    r = s.rect(rtb.PLANE_XY, 0, 0, 1, 1, 0.5, m)
    w = s.list([s.rotate(rtb.AXIS_Y, r, 10.0), s.sphere((0, 0, 0), 1.0, m2)])
    good = s.serialize()
    assert rtb.Scene.deserialize(good) is not None
    HDR, TEX, MAT, OBJ = 40, 48, 16, 112
    n_tex, n_mat = struct.unpack_from("<II", good, 8)
    tex_off = lambda i: HDR + i * TEX
    mat_off = lambda i: HDR + n_tex * TEX + i * MAT
    obj_off = lambda i: HDR + n_tex * TEX + n_mat * MAT + i * OBJ

    def broken(off, fmt, *vals):
        b = bytearray(good)
        struct.pack_into(fmt, b, off, *vals)
        return bytes(b)

    cases = {
        "checker references itself": broken(tex_off(t_c) + 4, "<i", t_c),
        "checker references a later texture": broken(tex_off(t_c) + 8, "<i", t_i),
        "unknown texture type": broken(tex_off(t_a), "<i", 7),
        "negative texture type": broken(tex_off(t_a), "<i", -1),
        "image nx beyond the pool": broken(tex_off(t_i) + 4, "<i", 1 << 20),
        "image offset beyond the pool": broken(tex_off(t_i) + 16, "<Q", 1 << 60),
        "image nx*ny overflows": broken(tex_off(t_i) + 4, "<ii", 0x7FFFFFFF, 0x7FFFFFFF),
        "unknown material type": broken(mat_off(m), "<i", 9),
        "rect plane selector": broken(obj_off(r) + 12, "<i", 3),
        "rotate axis selector": broken(obj_off(r + 1) + 12, "<i", -1),
        "image byte count wraps the size sum": broken(32, "<Q", 0xFFFFFFFFFFFFFF00),
        "huge object count": broken(16, "<I", 0xFFFFFFFF),
    }
    for what, blob in cases.items():
        with pytest.raises(rtb.RtbError):
            rtb.Scene.deserialize(blob)
    # random byte flips never crash: each either loads (and then flattens / counts without fault) or is refused
    rng = np.random.default_rng(0)
    for _ in range(300):
        b = bytearray(good)
        for _ in range(rng.integers(1, 6)):
            b[rng.integers(8, len(b))] = rng.integers(0, 256)
        try:
            s2 = rtb.Scene.deserialize(bytes(b))
        except rtb.RtbError:
            continue
        s2.prim_count(w)
