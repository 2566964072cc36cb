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
