"""Field-level GPU parity (run on the B200 box with -m gpu): each stage of the device path against the CPU oracle,
number against number, through the C ABI hooks of csrc/debug.cu.  The checks themselves live in tests/field_checks.py
(the same ones run on the CPU against the host-compiled device code, tests/test_emul_fields.py)."""
import os

import numpy as np
import pytest

import field_checks as fc

pytestmark = pytest.mark.gpu
NT = os.cpu_count() or 1


@pytest.fixture(scope="module", autouse=True)
def _need_gpu(rtb):
    assert rtb.device_count() > 0, "GPU tests need a CUDA device; the product has no CPU fallback"


# ---- the self-hit-exclusion path: every scattered ray of a render issues THIS query ------------------------------
@pytest.mark.parametrize("name", fc.ALL_SCENES)
def test_secondary_ray_ids_with_self_exclusion_agree_with_oracle(rtb, orc, name):
    nx = ny = 1024  # 2^20 primary rays -> up to 2^20 scattered rays leaving oracle hit points
    s, w, cam = rtb.Scene.named(name, nx, ny)
    o = orc.OracleScene(s.serialize(), w)
    rays = rtb.primary_rays(cam, nx, ny, sample=1, seed=20240608)
    r = fc.check_self_hit_ids(s, o, w, rays, np.random.default_rng(11), min_agree=0.9999)
    print(name, r)
    if name not in ("earth",):
        assert r["secondary"] >= 200_000, r  # not vacuous


def test_closest_hit_hooks_agree_on_both_trees(rtb, orc, monkeypatch):
    # the query kernels walk the upload-time choice of tree; RTB200_BVH_WIDTH forces either: same answers, ray by ray
    nx = ny = 512
    for name in ("final_scene", "random_scene", "cornell_box", "two_spheres"):
        s, w, cam = rtb.Scene.named(name, nx, ny)
        rays = rtb.primary_rays(cam, nx, ny, sample=4, seed=3)
        p_ref, _ = orc.OracleScene(s.serialize(), w).closest_hit(rays)
        got = {}
        for width in ("2", "4"):
            monkeypatch.setenv("RTB200_BVH_WIDTH", width)
            got[width], _ = s.closest_hit_ex(w, rays)
            assert float((got[width] == p_ref).mean()) >= 0.9999, (name, width)
        monkeypatch.delenv("RTB200_BVH_WIDTH")
        assert float((got["2"] == got["4"]).mean()) >= 0.99999


# ---- HitRecord fields -------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", fc.ALL_SCENES)
def test_hit_record_fields_match_oracle(rtb, orc, name):
    nx = ny = 512
    s, w, cam = rtb.Scene.named(name, nx, ny)
    o = orc.OracleScene(s.serialize(), w)
    rays = rtb.primary_rays(cam, nx, ny, sample=2, seed=99)
    print(name, fc.check_hit_records(s, o, w, rays))


# ---- geometry no named scene builds: Rotate X / Z, chains, transformed moving / textured spheres, Rects in a BVHNode ----
def test_zoo_rotations_chains_and_bvh_rects(rtb, orc):
    nx = ny = 1024
    s, w, cam = fc.build_zoo(rtb, nx, ny)
    o = orc.OracleScene(s.serialize(), w)
    for smp in (0, 1):
        rays = rtb.primary_rays(cam, nx, ny, sample=smp, seed=5)
        p, _ = o.closest_hit(rays)
        assert len(np.unique(p)) >= 25  # every object family is in view
        print(fc.check_self_hit_ids(s, o, w, rays, np.random.default_rng(3 + smp)))
        print(fc.check_hit_records(s, o, w, rays))
    # and as an image (uv-textured rotated sphere / rect, glass under a transform): noise-floor parity at 64x64
    from parity_utils import image_parity
    n, ns = 64, 1024
    s, w, cam = fc.build_zoo(rtb, n, n)
    a, _, _ = o.render(cam, n, n, ns, background=1, seed=101, nthreads=NT)
    b, _, _ = o.render(cam, n, n, ns, background=1, seed=202, nthreads=NT)
    g, _, _ = s.render(w, cam, n, n, ns, rtb.default_opts(background=1, seed=7), want_rgb8=False)
    h, _, _ = s.render(w, cam, n, n, 16 * ns, rtb.default_opts(background=1, seed=9), want_rgb8=False)
    r = image_parity(g / ns, h / (16 * ns), a, b)
    print("zoo image", r)
    assert r["gpu_db"] >= r["floor_db"] - 0.5 and r["bias_db"] >= r["floor_db"] + 4.0, r


def test_rect_in_bvh_is_true_geometry_not_the_reference_bbox_bug(rtb, orc):
    # Q8: Rect::bounding_box ignores `plane` (src/rect.rs:71-75), so the reference (and the oracle) CULL a ZX / YZ Rect
    # that sits directly in a BVHNode.  The GPU traverses its own tree over true bounds: it returns what the same Rects
    # give in a HittableList.  This pins the documented divergence (include/rtb200.h).
    def build(in_bvh):
        s = rtb.Scene()
        m = s.lambertian(s.solid(0.5, 0.5, 0.5))
        rects = [s.rect(rtb.PLANE_ZX, -2 + i, -2, -1.2 + i, 2, 0.3 * i, m) for i in range(4)] + \
                [s.rect(rtb.PLANE_YZ, -2, -2 + i, 2, -1.2 + i, 3.0 + 0.3 * i, m) for i in range(4)]
        return s, (s.bvh(rects, 0.0, 1.0, seed=1) if in_bvh else s.list(rects))
    cam = rtb.camera((6, 7, 9), (0, 0, 0), (0, 1, 0), 40.0, 1.0, 0.0, 10.0, 0.0, 1.0)
    rays = rtb.primary_rays(cam, 256, 256, sample=0, seed=2)
    s_b, w_b = build(True)
    s_l, w_l = build(False)
    p_gpu_bvh, _ = s_b.closest_hit_ex(w_b, rays)
    p_gpu_list, _ = s_l.closest_hit_ex(w_l, rays)
    p_orc_list, _ = orc.OracleScene(s_l.serialize(), w_l).closest_hit(rays)
    p_orc_bvh, _ = orc.OracleScene(s_b.serialize(), w_b).closest_hit(rays)
    assert (p_gpu_bvh == p_gpu_list).all() and (p_gpu_list == p_orc_list).mean() >= 0.9999
    assert (p_orc_list >= 0).mean() > 0.05
    assert (p_orc_bvh >= 0).sum() < (p_orc_list >= 0).sum()  # the reference loses hits to its own bounding boxes


# ---- a ConstantMedium far from the ray origin: the reference's +1e-4 restart in f32 (src/medium.rs:30-31) -----------------
@pytest.mark.parametrize("boundary", ["box", "xform"])
@pytest.mark.parametrize("distance", [60.0, 3000.0])
def test_far_constant_medium_transmittance(rtb, orc, boundary, distance):
    s, w, cam, T = fc.build_far_medium(rtb, distance=distance, boundary=boundary)
    n, ns = 64, 256
    g, _, _ = s.render(w, cam, n, n, ns, want_rgb8=False)
    a, _, _ = orc.OracleScene(s.serialize(), w).render(cam, n, n, 64, seed=1, nthreads=NT)
    sigma = np.sqrt(T * (1 - T) / (n * n * ns))
    assert abs(g[..., 0].mean() / ns - T) < 5 * sigma + 1e-4, (g[..., 0].mean() / ns, T)
    assert abs(a[..., 0].mean() - T) < 5 * np.sqrt(T * (1 - T) / (n * n * 64)) + 1e-4  # the f64 reference agrees with Beer-Lambert


# ---- camera, samplers, scatter ------------------------------------------------------------------------------------------------
def test_device_camera_rays_match_oracle_get_ray(rtb, orc):
    for name in ("random_scene", "simple_light", "cornell_box", "final_scene"):
        _, _, cam = rtb.Scene.named(name, 256, 144)
        print(name, fc.check_camera_rays(rtb, orc, rtb, cam, 256, 144))
    cam = rtb.camera((1, 2, 3), (0, 0, 0), (0, 1, 0), 35.0, 4 / 3, 0.0, 4.0, 0.2, 0.7)  # lens_radius == 0: no lens draw
    print("pinhole", fc.check_camera_rays(rtb, orc, rtb, cam, 200, 150))


def test_device_samplers_have_the_rejection_loops_distribution(rtb, orc):
    print(fc.check_samplers(rtb, orc, n=1_000_000))


def test_device_scatter_matches_oracle_per_material(rtb, orc):
    s = rtb.Scene()
    tex = s.checker(s.solid(0.2, 0.3, 0.1), s.solid(0.9, 0.8, 0.7))
    mats = {"lambertian": s.lambertian(tex), "metal": s.metal(s.solid(0.8, 0.6, 0.2), 0.3), "metal0": s.metal(s.solid(0.7, 0.7, 0.7), 0.0),
            "metal_big": s.metal(s.solid(0.8, 0.8, 0.9), 10.0), "dielectric": s.dielectric(1.5), "isotropic": s.isotropic(s.solid(0.2, 0.4, 0.9)),
            "light": s.diffuse_light(s.solid(7, 7, 7))}
    w = s.list([s.sphere((0, 0, 0), 1.0, m) for m in mats.values()])
    o = orc.OracleScene(s.serialize(), w)
    nrm = np.array([0.0, 1.0, 0.0])
    n = 400_000
    for key, kind in (("lambertian", "lambertian"), ("metal", "metal"), ("metal0", "metal"), ("metal_big", "metal"),
                      ("isotropic", "isotropic"), ("light", "light")):
        print(key, fc.check_scatter_material(s, o, w, mats[key], kind, (0.3, -1.0, 0.2), nrm, n=n))
    print("metal grazing", fc.check_scatter_material(s, o, w, mats["metal"], "metal", (1.0, -0.05, 0.0), nrm, n=n))
    # dielectric: from outside, grazing from outside, from inside below / near / beyond the critical angle (TIR)
    for d_in in ((0.3, -1.0, 0.2), (2.0, -0.3, 0.1), (0.2, 1.0, -0.1), (1.0, 0.6, 0.0), (1.0, 0.9, 0.0), (1.0, 0.2, 0.0)):
        print("dielectric", d_in, fc.check_scatter_material(s, o, w, mats["dielectric"], "dielectric", d_in, nrm, n=n))
