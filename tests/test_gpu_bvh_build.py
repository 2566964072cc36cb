"""The device-built BVH (csrc/bvh_build.cu): closest-hit answers do not depend on the tree, so the
LBVH built by CUDA kernels must give the oracle's primitive IDs, the host-SAH renderer's images and its path statistics."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["final_scene", "random_scene", "cornell_box", "two_spheres", "cornell_smoke_book"])
def test_device_lbvh_closest_hit_ids_agree_with_oracle(rtb, orc, name):
    nx = ny = 512
    s, w, cam = rtb.Scene.named(name, nx, ny)
    s.set_bvh_builder(rtb.BVH_DEVICE_LBVH)
    rays = rtb.primary_rays(cam, nx, ny, sample=0, seed=77)
    p_ref, t_ref = orc.OracleScene(s.serialize(), w).closest_hit(rays, media=False)
    p_gpu, t_gpu = s.closest_hit(w, rays)
    assert float((p_ref == p_gpu).mean()) >= 0.9999
    s.set_bvh_builder(rtb.BVH_HOST_SAH)  # the two builders answer identically (same primitive tests, same tie rule)
    p_sah, t_sah = s.closest_hit(w, rays)
    assert (p_sah == p_gpu).mean() >= 0.99999
    np.testing.assert_allclose(t_sah[p_sah == p_gpu], t_gpu[p_sah == p_gpu], rtol=1e-6)


@pytest.mark.parametrize("pipe", [1, 2])
def test_device_lbvh_renders_the_same_estimator(rtb, pipe):
    nx, ny, ns = 96, 64, 24
    for name in ("final_scene_lit", "cornell_box"):
        s, w, cam = rtb.Scene.named(name, nx, ny)
        a, _, sa = s.render(w, cam, nx, ny, ns, rtb.default_opts(seed=4, pipeline=pipe), want_rgb8=False, want_stats=True)
        s.set_bvh_builder(rtb.BVH_DEVICE_LBVH)
        b, _, sb = s.render(w, cam, nx, ny, ns, rtb.default_opts(seed=4, pipeline=pipe), want_rgb8=False, want_stats=True)
        # identical draws; a path changes only where two primitives tie within f32 rounding
        assert abs(sa["segments"] - sb["segments"]) <= 2e-4 * sa["segments"]
        close = np.isclose(a, b, rtol=1e-3, atol=1e-3).all(axis=-1)
        assert close.mean() >= 0.98 and abs(a.sum() / b.sum() - 1.0) < 2e-3


def test_device_lbvh_large_scene_and_edge_cases(rtb):
    # 60 000 random spheres + duplicates (identical Morton codes): device and host trees agree on every primary ray
    rng = np.random.default_rng(5)
    s = rtb.Scene()
    mat = s.lambertian(s.solid(0.5, 0.5, 0.5))
    objs = [s.sphere(c, r, mat) for c, r in zip(rng.uniform(-50, 50, (60000, 3)), rng.uniform(0.05, 0.6, 60000))]
    objs += [s.sphere((1.0, 2.0, 3.0), 0.5, mat) for _ in range(40)]  # coincident primitives
    w = s.list(objs)
    cam = rtb.camera((0, 0, -160), (0, 0, 0), (0, 1, 0), 40.0, 1.0, 0.0, 10.0, 0.0, 1.0)
    rays = rtb.primary_rays(cam, 256, 256, sample=0, seed=3)
    s.set_bvh_builder(rtb.BVH_DEVICE_LBVH)
    p_dev, t_dev = s.closest_hit(w, rays)
    s.set_bvh_builder(rtb.BVH_HOST_SAH)
    p_host, t_host = s.closest_hit(w, rays)
    assert (p_dev >= 0).mean() > 0.3
    same = p_dev == p_host
    assert same.mean() >= 0.9995  # (rays that hit the coincident spheres may pick a different twin)
    np.testing.assert_allclose(t_dev[same], t_host[same], rtol=1e-6)
    # two primitives (the smallest tree the device builder takes) and one (host path)
    for n in (2, 1):
        s2 = rtb.Scene()
        m2 = s2.lambertian(s2.solid(0.5, 0.5, 0.5))
        w2 = s2.list([s2.sphere((3.0 * k, 0, 0), 1.0, m2) for k in range(n)])
        s2.set_bvh_builder(rtb.BVH_DEVICE_LBVH)
        r = np.array([[3.0 * k, 0, -5, 0, 0, 1, 0] for k in range(n)] + [[100.0, 0, -5, 0, 0, 1, 0]], np.float64)
        p, t = s2.closest_hit(w2, r)
        assert list(p) == list(range(n)) + [-1] and np.allclose(t[:n], 4.0)
    with pytest.raises(rtb.RtbError):
        s2.set_bvh_builder(7)
