"""Field-level parity checks of the device path against the CPU oracle, written against a small backend interface
so the same checks run
  * on the GPU through the C ABI (tests/test_gpu_fields.py, -m gpu): backend = raytracing_rust_b200 itself;
  * on the CPU against the kernels' device functions compiled for the host (tests/test_emul_fields.py):
    backend = tools/host_emul — a logic check of trace.cuh that needs no GPU.
A backend `scene` offers closest_hit_ex(world, rays, self_id) and scatter(world, mat, inp, seed); a backend module
offers device_camera_rays(cam, nx, ny, sample, seed) and device_sample(kind, seed, n).
Reference lines cited per check."""
import numpy as np

ALL_SCENES = ("random_scene", "two_spheres", "two_perlin_sphere", "earth", "simple_light", "cornell_box",
              "cornell_smoke", "final_scene", "final_scene_lit", "cornell_smoke_book")


def _unit_ball(rng, n):
    v = rng.normal(size=(n, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    return v * rng.uniform(0, 1, (n, 1)) ** (1.0 / 3.0)


def scattered_rays(rays, prim, rec, rng):
    """Host-side stand-ins for Material::scatter (src/material.rs:49-168) at oracle hit points: the four direction
    families a render produces — normal + ball (Lambertian :50), mirror reflection (:9, Metal / Dielectric), a
    direction INTO the surface (refraction :13-23), uniform ball (Isotropic :166).  Returns (rays2 [m,7], self_id [m])."""
    hit = prim >= 0
    p, n = rec[hit, 3:6], rec[hit, 6:9]
    d_in = rays[hit, 3:6]
    m = p.shape[0]
    kind = rng.integers(0, 4, m)
    ball = _unit_ball(rng, m)
    du = d_in / np.linalg.norm(d_in, axis=1, keepdims=True)
    refl = du - 2.0 * np.sum(du * n, axis=1, keepdims=True) * n
    side = np.where(np.sum(d_in * n, axis=1, keepdims=True) > 0, 1.0, -1.0)  # continue through the surface
    through = side * n + 0.7 * ball
    d = np.where((kind == 0)[:, None], n + ball, np.where((kind == 1)[:, None], refl, np.where((kind == 2)[:, None], through, ball)))
    out = np.concatenate([p, d, rays[hit, 6:7]], axis=1)
    return out, prim[hit].astype(np.int32)


def check_self_hit_ids(backend_scene, oracle_scene, world, rays, rng, min_agree=0.9999):
    """Secondary rays leaving oracle hit points: oracle f64 with t_min = 0.001 (src/color.rs:7) vs the device query that is
    told which surface the ray leaves (trace.cuh intersect_prim `self`).  Returns a dict of figures; asserts agreement."""
    p1, _, rec1, _ = oracle_scene.closest_hit(rays, media=False, want_rec=True)
    rays2, self_id = scattered_rays(rays, p1, rec1, rng)
    out = {"primary_hits": int((p1 >= 0).sum()), "secondary": int(rays2.shape[0])}
    if rays2.shape[0] == 0:
        out["agree"] = 1.0
        return out
    p_ref, t_ref = oracle_scene.closest_hit(rays2, t_min=1e-3, media=False)
    p_dev, rec_dev = backend_scene.closest_hit_ex(world, rays2, self_id=self_id, t_min=1e-3)
    agree = float((p_ref == p_dev).mean())
    out["agree"] = out["secondary_agree"] = agree
    out["rehit_self_ref"] = float((p_ref == self_id).mean())  # how often the ray meets its own primitive again (far root / other face)
    both = (p_ref == p_dev) & (p_ref >= 0)
    if both.any():
        rel = np.abs(rec_dev[both, 0] - t_ref[both]) / np.maximum(np.abs(t_ref[both]), 1e-3)
        out["t_rel_q999"] = out["secondary_t_rel_q999"] = float(np.quantile(rel, 0.999))
    assert agree >= min_agree, out
    return out


def check_hit_records(backend_scene, oracle_scene, world, rays, tol_scale=1.0):
    """HitRecord fields {t,u,v,p,normal} (src/hittable.rs:9-16) of primary hits, number against number.
    Tolerances are f32 round-off at scene scale: |p| up to ~1e3 -> 1e-3 absolute; normals 1e-3; u,v 2e-3 away from the
    atan2 seam (src/sphere.rs:9-15, Q5) and the Rect/Cube uv of src/rect.rs:54-55."""
    p_ref, t_ref, rec_ref, _ = oracle_scene.closest_hit(rays, media=False, want_rec=True)
    p_dev, rec_dev = backend_scene.closest_hit_ex(world, rays, self_id=None)
    both = (p_ref == p_dev) & (p_ref >= 0)
    out = {"agree": float((p_ref == p_dev).mean()), "hits": int(both.sum())}
    if not both.any():
        return out
    a, b = rec_ref[both], rec_dev[both].astype(np.float64)
    # p = o + t*d: its round-off scales with the longer of |p| and the ray length t*|d|
    ray_len = np.abs(a[:, 0:1] * rays[both, 3:6]).max(axis=1)
    scale = np.maximum(1.0, np.maximum(np.abs(a[:, 3:6]).max(axis=1), ray_len))
    rel_t = np.abs(a[:, 0] - b[:, 0]) / np.maximum(np.abs(a[:, 0]), 1e-3)
    dp = np.abs(a[:, 3:6] - b[:, 3:6]).max(axis=1) / scale
    dn = np.abs(a[:, 6:9] - b[:, 6:9]).max(axis=1)
    du = np.abs(a[:, 1] - b[:, 1])
    du = np.minimum(du, 1.0 - du)  # the atan2 seam: u wraps
    dv = np.abs(a[:, 2] - b[:, 2])
    out.update(t_rel_q999=float(np.quantile(rel_t, 0.999)), p_rel_q999=float(np.quantile(dp, 0.999)), p_rel_max=float(dp.max()), n_q999=float(np.quantile(dn, 0.999)),
               n_max=float(dn.max()), u_q999=float(np.quantile(du, 0.999)), v_q999=float(np.quantile(dv, 0.999)))
    assert out["agree"] >= 0.9999, out
    assert out["t_rel_q999"] < 3e-4 * tol_scale, out
    # relative to |p|: a few f32 ulps; the tail is grazing hits on the r = 1000 ground spheres, where t is ill-conditioned
    assert out["p_rel_q999"] < 2e-4 * tol_scale and out["p_rel_max"] < 2e-3 * tol_scale, out
    assert out["n_q999"] < 2e-3 * tol_scale, out          # normal = (p - c)/r: f32 cancellation on the r = 1000 spheres
    assert out["u_q999"] < 2e-3 * tol_scale and out["v_q999"] < 2e-3 * tol_scale, out
    return out


def check_camera_rays(backend_mod, orc_mod, rtb_mod, cam, nx, ny, seed=77):
    """Camera::get_ray (src/camera.rs:53-67): device f32 rays vs the oracle's get_ray fed the same draws."""
    draws = rtb_mod.camera_draws(nx, ny, sample=3, seed=seed)
    ref = orc_mod.camera_rays(cam, draws)
    host = rtb_mod.primary_rays(cam, nx, ny, sample=3, seed=seed)
    dev = backend_mod.device_camera_rays(cam, nx, ny, sample=3, seed=seed).astype(np.float64)
    # (1) the library's f64 ray generator IS the oracle's get_ray on these draws
    np.testing.assert_allclose(host, ref, rtol=1e-12, atol=1e-12)
    # (2) the device's f32 evaluation of it
    o_scale = max(1.0, np.abs(ref[:, :3]).max())
    d_scale = np.abs(ref[:, 3:6]).max()
    out = {"origin_err": float(np.abs(dev[:, :3] - ref[:, :3]).max() / o_scale),
           "dir_err": float(np.abs(dev[:, 3:6] - ref[:, 3:6]).max() / d_scale),
           "time_err": float(np.abs(dev[:, 6] - ref[:, 6]).max())}
    assert out["origin_err"] < 2e-6 and out["dir_err"] < 2e-5 and out["time_err"] < 1e-6, out
    return out


def _ks_against_cdf(x, cdf):
    x = np.sort(x)
    n = x.size
    f = cdf(x)
    return float(max(np.abs(f - np.arange(1, n + 1) / n).max(), np.abs(f - np.arange(0, n) / n).max()))


def check_samplers(backend_mod, orc_mod, n=400_000):
    """random_in_unit_sphere / _disk (src/util.rs:4-24): the device's direct samplers have the rejection loops'
    distribution — radial CDF r^3 (ball) / r^2 (disk) by a KS test against the analytic CDF AND against the oracle's
    rejection sampler, isotropy through the first two moments."""
    out = {}
    crit = 1.95 / np.sqrt(n)  # KS critical value at alpha ~ 0.001
    for kind, power, name in ((0, 3.0, "ball"), (1, 2.0, "disk")):
        d = backend_mod.device_sample(kind, 1234 + kind, n).astype(np.float64)
        o = orc_mod.sample(kind, 99 + kind, n)
        rd, ro = np.linalg.norm(d, axis=1), np.linalg.norm(o, axis=1)
        assert rd.max() <= 1.0 + 1e-6
        ks_cdf = _ks_against_cdf(rd, lambda r: np.clip(r, 0, 1) ** power)
        # two-sample KS vs the oracle's rejection sampler
        allr = np.concatenate([rd, ro])
        order = np.argsort(allr, kind="mergesort")
        step = np.where(order < n, 1.0 / n, -1.0 / n)
        ks_two = float(np.abs(np.cumsum(step)).max())
        dims = 3 if kind == 0 else 2
        mean_err = float(np.abs(d[:, :dims].mean(axis=0)).max())
        second = (d[:, :dims] ** 2).mean(axis=0)
        want_second = 0.2 if kind == 0 else 0.25  # E[x^2] of the uniform ball / disk
        cross = float(np.abs(np.cov(d[:, :dims].T) - np.diag(np.diag(np.cov(d[:, :dims].T)))).max())
        out[name] = {"ks_vs_cdf": ks_cdf, "ks_vs_oracle": ks_two, "mean_err": mean_err,
                     "second_moment_err": float(np.abs(second - want_second).max()), "cross_cov": cross}
        assert ks_cdf < crit and ks_two < crit * np.sqrt(2.0), out
        assert mean_err < 5.0 / np.sqrt(n) and out[name]["second_moment_err"] < 5.0 / np.sqrt(n) and cross < 5.0 / np.sqrt(n), out
        if kind == 1:
            assert np.all(d[:, 2] == 0.0)
    return out


def scatter_inputs(n, d_in, normal, p=(1.0, 2.0, 3.0), uv=(0.3, 0.6)):
    """n copies of one (ray, hit) pair in the oracle's orc_scatter layout: {o, d, time, t, u, v, p, normal}."""
    d_in, normal, p = np.asarray(d_in, float), np.asarray(normal, float), np.asarray(p, float)
    row = np.concatenate([p - d_in, d_in, [0.25], [1.0], uv, p, normal])
    return np.tile(row, (n, 1))


def check_scatter_material(backend_scene, oracle_scene, world, mat, kind, d_in, normal, n=200_000):
    """Material::scatter for ONE fixed (ray, hit) and n independent draws, device vs oracle (src/material.rs:49-168).
    kind: 'lambertian' | 'metal' | 'dielectric' | 'isotropic' | 'light'.  Compares the scatter probability, the
    attenuation (exact: same texture code path), and the direction distribution (branch frequencies and per-branch
    direction for the two-outcome dielectric; first two moments for the ball-based ones)."""
    inp = scatter_inputs(n, d_in, normal)
    a = oracle_scene.scatter(mat, inp, seed=5)
    b = backend_scene.scatter(world, mat, inp, seed=6).astype(np.float64)
    out = {"p_scatter_ref": float(a[:, 0].mean()), "p_scatter_dev": float(b[:, 0].mean())}
    sig = 4.5 * np.sqrt(max(out["p_scatter_ref"] * (1 - out["p_scatter_ref"]), 1e-12) * 2.0 / n)
    assert abs(out["p_scatter_ref"] - out["p_scatter_dev"]) <= sig + 1e-12, out
    if kind == "light":
        assert out["p_scatter_dev"] == 0.0, out  # DiffuseLight never scatters (:144-146)
        return out
    oka, okb = a[:, 0] > 0, b[:, 0] > 0
    if okb.any() and oka.any():
        np.testing.assert_allclose(b[okb, 1:4], np.tile(inp[0, 10:13], (okb.sum(), 1)), rtol=1e-6)  # origin = hit.p
        att_a, att_b = a[oka, 7:10], b[okb, 7:10]
        assert np.abs(att_a.mean(axis=0) - att_b.mean(axis=0)).max() < 1e-6, (att_a.mean(axis=0), att_b.mean(axis=0))
        da, db = a[oka, 4:7], b[okb, 4:7]
        if kind == "dielectric":
            nrm = np.asarray(normal, float)
            # two outcomes: reflect(d, n) and the refracted direction; classify by the side of the surface
            side_a, side_b = np.sign(da @ nrm), np.sign(db @ nrm)
            incoming = np.sign(np.asarray(d_in, float) @ nrm)
            fa, fb = float((side_a != incoming).mean()), float((side_b != incoming).mean())  # reflected fraction
            out.update(reflect_frac_ref=fa, reflect_frac_dev=fb)
            assert abs(fa - fb) <= 4.5 * np.sqrt(max(fa * (1 - fa), 1e-12) * 2.0 / n) + 1e-12, out
            for sgn in (-1.0, 1.0):
                ma, mb = side_a == sgn, side_b == sgn
                if ma.any() and mb.any():
                    assert np.abs(da[ma].mean(axis=0) - db[mb].mean(axis=0)).max() < 2e-5 * max(1.0, np.abs(da).max()), out
                    assert db[mb].std(axis=0).max() < 1e-5 * max(1.0, np.abs(da).max()), out  # each branch is one direction
        else:
            tol = 5.0 / np.sqrt(min(oka.sum(), okb.sum()))
            out.update(dir_mean_err=float(np.abs(da.mean(axis=0) - db.mean(axis=0)).max()),
                       dir_cov_err=float(np.abs(np.cov(da.T) - np.cov(db.T)).max()))
            assert out["dir_mean_err"] < tol and out["dir_cov_err"] < tol, out
    return out


def build_zoo(rtb, nx=512, ny=512):
    """Geometry the eight named scenes never build (VERDICT r1 'untested geometry'): Rotate about X and Z
    (src/rotate.rs:13-19), Rotate∘Traslate∘Rotate chains, a MovingSphere and an image-textured Sphere under transforms,
    FlipNormals under a transform, and XY Rects directly inside a BVHNode (src/bvh.rs:17-66 with the one plane whose
    Rect::bounding_box is right, Q8).  Returns (scene, world, camera)."""
    s = rtb.Scene()
    grey = s.lambertian(s.solid(0.7, 0.7, 0.7))
    chk = s.lambertian(s.checker(s.solid(0.2, 0.3, 0.1), s.solid(0.9, 0.9, 0.9)))
    img = s.lambertian(s.image(rtb.earth_rgb8()[::8, ::8].copy()))
    glass = s.dielectric(1.5)
    objs = []
    # row 1: one cube per rotation axis
    for k, axis in enumerate((rtb.AXIS_X, rtb.AXIS_Y, rtb.AXIS_Z)):
        objs.append(s.translate(s.rotate(axis, s.cube((-1, -1, -1), (1, 1.5, 2), grey), 25.0 + 10 * k), (-6 + 6 * k, 4, 0)))
    # row 2: chains  T(Rz(T(Rx(cube))))  and  Ry(Rz(Rx(rect)))
    inner = s.translate(s.rotate(rtb.AXIS_X, s.cube((-1, -1, -1), (1, 1, 1), chk), -35.0), (0.5, 0.25, -0.5))
    objs.append(s.translate(s.rotate(rtb.AXIS_Z, inner, 50.0), (-6, 0, 0)))
    r = s.rect(rtb.PLANE_ZX, -1.5, -1.5, 1.5, 1.5, 0.0, img)
    objs.append(s.translate(s.rotate(rtb.AXIS_Y, s.rotate(rtb.AXIS_Z, s.rotate(rtb.AXIS_X, r, 20.0), 30.0), 40.0), (0, 0, 0)))
    objs.append(s.translate(s.rotate(rtb.AXIS_X, s.flip_normals(s.rect(rtb.PLANE_YZ, -1, -1, 1, 1, 0.0, grey)), 70.0), (6, 0, 0)))
    # row 3: spheres under transforms (image-textured: uv is taken in the sphere's own frame; moving; glass)
    objs.append(s.translate(s.rotate(rtb.AXIS_Z, s.sphere((0, 0, 0), 1.5, img), 65.0), (-6, -4, 0)))
    objs.append(s.translate(s.rotate(rtb.AXIS_X, s.moving_sphere((0, 0, 0), (0.5, 1.0, 0.0), 0.0, 1.0, 1.0, chk), 45.0), (0, -4, 0)))
    objs.append(s.rotate(rtb.AXIS_Z, s.translate(s.sphere((0, 0, 0), 1.2, glass), (6, -4, 0)), 5.0))
    # XY rects directly in a BVHNode, itself under a rotation about Z
    rects = [s.rect(rtb.PLANE_XY, -9 + 2.0 * i, -7.5, -8 + 2.0 * i, -6.5, 0.1 * i, grey) for i in range(9)]
    objs.append(s.rotate(rtb.AXIS_Z, s.bvh(rects, 0.0, 1.0, seed=3), 3.0))
    w = s.list(objs)
    cam = rtb.camera((0.5, 0.3, 22.0), (0, 0, 0), (0, 1, 0), 42.0, nx / ny, 0.3, 22.0, 0.0, 1.0)
    return s, w, cam


def build_far_medium(rtb, distance=3000.0, density=0.02, thickness=40.0, boundary="box"):
    """A ConstantMedium (src/medium.rs:28-56) whose boundary is crossed at t ~ `distance` in units of |d| ~ 1, where
    the reference's `hit1.t + 0.0001` (medium.rs:31) is below one f32 ulp; black absorber in front of a white light."""
    s = rtb.Scene()
    light = s.diffuse_light(s.solid(1, 1, 1))
    z0 = distance
    if boundary == "box":
        b = s.cube((-400, -400, z0), (400, 400, z0 + thickness), s.lambertian(s.solid(0, 0, 0)))
    else:  # a rotated + translated box: the generic (transform) boundary path
        b = s.translate(s.rotate(rtb.AXIS_Z, s.cube((-400, -400, 0), (400, 400, thickness), s.lambertian(s.solid(0, 0, 0))), 30.0),
                        (0, 0, z0))
    med = s.constant_medium(b, density, s.solid(0, 0, 0))
    w = s.list([med, s.rect(rtb.PLANE_XY, -5000, -5000, 5000, 5000, z0 + thickness + 50.0, light)])
    cam = rtb.camera((0, 0, 0), (0, 0, 1), (0, 1, 0), 2.0, 1.0, 0.0, 1.0, 0.0, 1.0)
    return s, w, cam, float(np.exp(-density * thickness))
