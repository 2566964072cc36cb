"""ctypes binding of oracle/_build/liboracle.so — the CPU f64 restatement of the reference.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / `--impl reference` legs.  The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = _HERE / "_build" / "liboracle.so"


def build():
    subprocess.run(["make", "-C", str(_HERE)], check=True, capture_output=True)


if not _LIB.exists():
    build()
lib = C.CDLL(str(_LIB))

_vp, _i, _d, _u64, _sz = C.c_void_p, C.c_int, C.c_double, C.c_uint64, C.c_size_t
lib.orc_scene_load.restype = _vp
lib.orc_scene_load.argtypes = [_vp, _sz, _i, C.c_char_p, _sz]
lib.orc_scene_free.argtypes = [_vp]
lib.orc_scene_prim_count.argtypes = [_vp]
lib.orc_scene_bvh_nodes.restype = _u64
lib.orc_scene_bvh_nodes.argtypes = [_vp]
lib.orc_render.restype = _u64
lib.orc_render.argtypes = [_vp, _vp, _i, _i, _i, _i, _d, _i, _u64, _i, _vp, _vp]
lib.orc_closest_hit.argtypes = [_vp, _vp, _sz, _d, _i, _u64, _vp, _vp, _vp, _vp]
lib.orc_texture_value.argtypes = [_vp, _i, _vp, _sz, _vp]
lib.orc_scatter.argtypes = [_vp, _i, _vp, _sz, _u64, _vp]
lib.orc_camera_rays.argtypes = [_vp, _vp, _sz, _vp]
lib.orc_sample.argtypes = [_i, _u64, _sz, _vp]


class Camera(C.Structure):
    """Camera::new arguments, src/camera.rs:21-31 (RtbCamera of include/rtb_scene_format.h): the oracle's own mirror, so that
    a caller (bench.py --impl reference) needs nothing from the product package."""
    _fields_ = [("look_from", C.c_double * 3), ("look_at", C.c_double * 3), ("view_up", C.c_double * 3),
                ("vertical_fov", C.c_double), ("aspect", C.c_double), ("aperture", C.c_double),
                ("focus_dist", C.c_double), ("time0", C.c_double), ("time1", C.c_double)]


FINAL_SCENE_WORLD = 1415  # world handle inside tests/golden/final_scene.rtscene.head (tools/make_reference_fixture.py)


def final_scene_bytes() -> bytes:
    """final_scene as written (tests/test.rs:419-523) as .rtscene bytes, from committed files only: the table part
    (tests/golden/final_scene.rtscene.head) + the decoded earthmap the scene's ImageTexture uses (assets/)."""
    root = _HERE.parent
    return (root / "tests" / "golden" / "final_scene.rtscene.head").read_bytes() + (root / "assets" / "earthmap_1024x512.rgb8").read_bytes()


def final_scene_camera(nx: int, ny: int) -> Camera:
    """set_camera of test_final_scene, tests/test.rs:819-830: (478,278,-600) -> (278,278,0), vfov 40, aspect nx/ny,
    focus_dist 10, aperture 0.1, shutter [0,1]."""
    c = Camera()
    c.look_from[:] = (478.0, 278.0, -600.0)
    c.look_at[:] = (278.0, 278.0, 0.0)
    c.view_up[:] = (0.0, 1.0, 0.0)
    c.vertical_fov, c.aspect, c.aperture, c.focus_dist, c.time0, c.time1 = 40.0, nx / ny, 0.1, 10.0, 0.0, 1.0
    return c


class OracleScene:
    """A scene loaded from the .rtscene bytes the product serialises (same bytes, same scene)."""

    def __init__(self, scene_bytes: bytes, world: int):
        err = C.create_string_buffer(256)
        self._p = lib.orc_scene_load(scene_bytes, len(scene_bytes), world, err, 256)
        if not self._p:
            raise RuntimeError("oracle: " + err.value.decode())

    def __del__(self):
        p, self._p = getattr(self, "_p", None), None
        if p:
            lib.orc_scene_free(p)

    def prim_count(self) -> int:
        return lib.orc_scene_prim_count(self._p)

    def bvh_nodes(self) -> int:
        return lib.orc_scene_bvh_nodes(self._p)

    def render(self, cam, nx, ny, ns, max_depth=50, t_min=1e-3, background=0, seed=1, nthreads=1):
        """create_image (tests/test.rs:55-85).  Returns (mean_rgb float64 [ny,nx,3] linear,
        rgb8 uint8 [ny,nx,3], segments)."""
        mean = np.empty((ny, nx, 3), np.float64)
        rgb8 = np.empty((ny, nx, 3), np.uint8)
        segs = lib.orc_render(self._p, C.addressof(cam), nx, ny, ns, max_depth, t_min, background, seed, nthreads,
                              mean.ctypes.data, rgb8.ctypes.data)
        return mean, rgb8, int(segs)

    def closest_hit(self, rays: np.ndarray, t_min=1e-3, media=False, seed=1, want_rec=False):
        r = np.ascontiguousarray(rays, dtype=np.float64)
        n = r.shape[0]
        prim = np.empty(n, np.int32)
        t = np.empty(n, np.float64)
        rec = np.empty((n, 9), np.float64) if want_rec else None
        counters = np.zeros(2, np.uint64)
        lib.orc_closest_hit(self._p, r.ctypes.data, n, t_min, int(media), seed, prim.ctypes.data, t.ctypes.data,
                            rec.ctypes.data if want_rec else None, counters.ctypes.data)
        if want_rec:
            return prim, t, rec, counters
        return prim, t

    def texture_value(self, tex: int, uvp: np.ndarray) -> np.ndarray:
        q = np.ascontiguousarray(uvp, dtype=np.float64)
        out = np.empty((q.shape[0], 3), np.float64)
        if lib.orc_texture_value(self._p, tex, q.ctypes.data, q.shape[0], out.ctypes.data) != 0:
            raise RuntimeError("oracle: bad texture handle")
        return out

    def scatter(self, mat: int, inp: np.ndarray, seed=1) -> np.ndarray:
        q = np.ascontiguousarray(inp, dtype=np.float64)
        out = np.empty((q.shape[0], 10), np.float64)
        if lib.orc_scatter(self._p, mat, q.ctypes.data, q.shape[0], seed, out.ctypes.data) != 0:
            raise RuntimeError("oracle: bad material handle")
        return out


def camera_rays(cam, inp: np.ndarray) -> np.ndarray:
    """Camera::get_ray with explicit random inputs; inp = [n,5] {s,t,rd_x,rd_y,xi_time}."""
    q = np.ascontiguousarray(inp, dtype=np.float64)
    out = np.empty((q.shape[0], 7), np.float64)
    lib.orc_camera_rays(C.addressof(cam), q.ctypes.data, q.shape[0], out.ctypes.data)
    return out


def sample(kind: int, seed: int, n: int) -> np.ndarray:
    out = np.empty((n, 3), np.float64)
    lib.orc_sample(kind, seed, n, out.ctypes.data)
    return out
