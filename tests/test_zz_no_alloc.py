"""rtb_closest_hit_noalloc — the closest-hit query that allocates nothing (static device scratch; include/rtb200.h) — against
rtb_closest_hit_ex, through the C ABI.  No counterpart in the reference.
Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_query(rtb, torch, s, rays, self_id):
    d_rays = torch.from_numpy(np.ascontiguousarray(rays, dtype=np.float32)).cuda()
    d_self = torch.from_numpy(np.ascontiguousarray(self_id, dtype=np.int32)).cuda() if self_id is not None else None
    n = rays.shape[0]
    d_prim = torch.full((n,), -7, dtype=torch.int32, device="cuda")
    d_rec = torch.full((n, 9), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    s.closest_hit_noalloc(d_rays.data_ptr(), n, d_prim.data_ptr(), d_rec.data_ptr(), d_self.data_ptr() if d_self is not None else 0,
                          stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return d_prim.cpu().numpy(), d_rec.cpu().numpy()


def test_noalloc_closest_hit_equals_closest_hit_ex_on_consecutive_calls(rtb):
    import torch

    assert rtb.device_count() > 0, "GPU tests need a CUDA device; the product has no CPU fallback"
    # two scenes in turn, the larger first: the shared static table keeps the previous call's contents, which must not leak
    for name, nx, ny in (("final_scene", 256, 144), ("cornell_box", 160, 160), ("final_scene", 256, 144)):
        s, w, cam = rtb.Scene.named(name, nx, ny)
        rays = rtb.primary_rays(cam, nx, ny, sample=0, seed=5)
        p1, rec1 = s.closest_hit_ex(w, rays)  # (uploads the scene: the set-up step that may allocate)
        hit = p1 >= 0
        # scattered rays that leave the surfaces just found: the self_id path, i.e. the one that uses the scratch table
        rng = np.random.default_rng(3)
        d2 = rec1[hit, 6:9] + 0.9 * rng.uniform(-0.57, 0.57, (int(hit.sum()), 3))
        rays2 = np.concatenate([rec1[hit, 3:6], d2, rays[hit, 6:7]], axis=1).astype(np.float32).astype(np.float64)
        sid = p1[hit].astype(np.int32)
        ref_p, ref_rec = s.closest_hit_ex(w, rays2, self_id=sid)
        for call in range(2):  # two calls in a row
            got_p, got_rec = _device_query(rtb, torch, s, rays2, sid)
            assert np.array_equal(got_p, ref_p), (name, call, float((got_p == ref_p).mean()))  # integer IDs: exact
            np.testing.assert_allclose(got_rec, ref_rec, rtol=1e-6, atol=0)                     # same kernel: bit-equal in practice
        # and without self_id (the table is not consulted)
        got_p, got_rec = _device_query(rtb, torch, s, rays.astype(np.float32).astype(np.float64), None)
        ref_p0, ref_rec0 = s.closest_hit_ex(w, rays.astype(np.float32).astype(np.float64))
        assert np.array_equal(got_p, ref_p0)
        np.testing.assert_allclose(got_rec, ref_rec0, rtol=1e-6, atol=0)
        assert (ref_p >= 0).mean() > 0.2  # not vacuous


def test_noalloc_rejects_what_it_cannot_serve(rtb):
    import torch

    out = torch.zeros(8, dtype=torch.int32, device="cuda")
    rays = torch.zeros(8, 7, dtype=torch.float32, device="cuda")
    rays[:, 5] = 1.0
    # not uploaded: uploading allocates, so it is the caller's step, never this call's
    s, w, cam = rtb.Scene.named("two_spheres", 8, 8)
    with pytest.raises(rtb.RtbError, match="not uploaded"):
        s.closest_hit_noalloc(rays.data_ptr(), 8, out.data_ptr())
    s.upload(w, 0)
    s.closest_hit_noalloc(rays.data_ptr(), 8, out.data_ptr())
    s.closest_hit_noalloc(rays.data_ptr(), 0, 0)  # empty batch
    # more primitive IDs than the static table holds (a Cube owns six): refused with a clear error
    big = rtb.Scene()
    m = big.lambertian(big.solid(0.5, 0.5, 0.5))
    n_cubes = rtb.NOALLOC_MAX_PRIM_IDS // 6 + 1
    objs = [big.cube((2.0 * i, 0, 0), (2.0 * i + 1, 1, 1), m) for i in range(n_cubes)]
    wb = big.list(objs)
    big.upload(wb, 0)
    with pytest.raises(rtb.RtbError, match="RTB_NOALLOC_MAX_PRIM_IDS"):
        big.closest_hit_noalloc(rays.data_ptr(), 8, out.data_ptr())
    torch.cuda.synchronize()
