"""CPU-side logic checks of the DEVICE code: csrc/trace.cuh compiled for the host (tools/host_emul, a developer tool —
not the product, not the oracle) against the f64 oracle, with the same field-level checks the GPU tests run
(tests/field_checks.py).  Catches traversal / self-hit / hit-record / scatter logic errors before a GPU is involved;
the numbers that count are the -m gpu ones."""
import numpy as np
import pytest

import field_checks as fc


@pytest.fixture(scope="module")
def emu():
    from tools.host_emul import emul_py as m
    return m


@pytest.mark.parametrize("name", fc.ALL_SCENES)
def test_emul_self_hit_and_records(rtb, orc, emu, name):
    nx = ny = 160
    s, w, cam = rtb.Scene.named(name, nx, ny)
    blob = s.serialize()
    o, e = orc.OracleScene(blob, w), emu.EmuScene(blob, w)
    rays = rtb.primary_rays(cam, nx, ny, sample=0, seed=20240607)
    r = fc.check_self_hit_ids(e, o, w, rays, np.random.default_rng(3), min_agree=0.9995)
    fc.check_hit_records(e, o, w, rays, tol_scale=2.0)
    if name in ("two_spheres", "cornell_box", "final_scene"):
        assert r["secondary"] > 1000


def test_emul_camera_and_samplers(rtb, orc, emu):
    for name in ("random_scene", "cornell_box", "final_scene"):
        _, _, cam = rtb.Scene.named(name, 64, 48)
        fc.check_camera_rays(emu, orc, rtb, cam, 64, 48)
    cam = rtb.camera((1, 2, 3), (0, 0, 0), (0, 1, 0), 35.0, 4 / 3, 0.0, 4.0, 0.2, 0.7)  # pinhole: the lens draw is skipped
    fc.check_camera_rays(emu, orc, rtb, cam, 64, 48)
    fc.check_samplers(emu, orc, n=100_000)


def test_emul_scatter_materials(rtb, orc, emu):
    s = rtb.Scene()
    tex = s.checker(s.solid(0.2, 0.3, 0.1), s.solid(0.9, 0.8, 0.7))
    mats = {"lambertian": s.lambertian(tex), "metal": s.metal(s.solid(0.8, 0.6, 0.2), 0.3), "metal0": s.metal(s.solid(0.7, 0.7, 0.7), 0.0),
            "metal_big": s.metal(s.solid(0.8, 0.8, 0.9), 10.0), "dielectric": s.dielectric(1.5), "isotropic": s.isotropic(s.solid(0.2, 0.4, 0.9)),
            "light": s.diffuse_light(s.solid(7, 7, 7))}
    w = s.list([s.sphere((0, 0, 0), 1.0, m) for m in mats.values()])
    blob = s.serialize()
    o, e = orc.OracleScene(blob, w), emu.EmuScene(blob, w)
    nrm = np.array([0.0, 1.0, 0.0])
    for key, kind in (("lambertian", "lambertian"), ("metal", "metal"), ("metal0", "metal"), ("metal_big", "metal"),
                      ("isotropic", "isotropic"), ("light", "light")):
        fc.check_scatter_material(e, o, w, mats[key], kind, (0.3, -1.0, 0.2), nrm, n=40_000)
    fc.check_scatter_material(e, o, w, mats["metal"], "metal", (1.0, -0.05, 0.0), nrm, n=40_000)  # grazing: part absorbed
    for d_in in ((0.3, -1.0, 0.2), (2.0, -0.3, 0.1), (0.2, 1.0, -0.1), (1.0, 0.6, 0.0), (1.0, 0.9, 0.0)):  # outside, grazing, inside, TIR
        fc.check_scatter_material(e, o, w, mats["dielectric"], "dielectric", d_in, nrm, n=40_000)


def test_flattener_knows_a_room_from_an_open_scene(rtb, emu):
    # FlatScene::enclosed picks the point up to which the megakernel scans primitives instead of walking a tree (render.cu)
    want = {"two_spheres": (2, False), "simple_light": (4, False), "cornell_box": (8, True), "cornell_smoke": (6, True),
            "random_scene": (None, False), "final_scene": (1407, False)}
    for name, (prims, enclosed) in want.items():
        s, w, _ = rtb.Scene.named(name, 16, 16)
        f = emu.EmuScene(s.serialize(), w).facts()
        assert f["enclosed"] == enclosed and (prims is None or f["prims"] == prims), (name, f)
    s = rtb.Scene()  # three walls of a room only: not enclosed
    m = s.lambertian(s.solid(0.5, 0.5, 0.5))
    w = s.list([s.rect(0, 0, 0, 10, 10, 0, m), s.rect(1, 0, 0, 10, 10, 0, m), s.rect(2, 0, 0, 10, 10, 10, m)])
    assert not emu.EmuScene(s.serialize(), w).facts()["enclosed"]
