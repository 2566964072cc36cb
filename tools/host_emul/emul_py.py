"""ctypes binding of tools/host_emul/_build/libemul.so — DEVELOPER TOOL (see emul.cpp): the kernels' device functions
compiled for the host.  Used by tests/ (CPU-side logic checks of the device code against the oracle) and by the BVH
layout experiments; never imported by the product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = _HERE / "_build" / "libemul.so"


def build():
    subprocess.run(["make", "-C", str(_HERE)], check=True, capture_output=True)


build()
lib = C.CDLL(str(_LIB))
_vp, _i, _d, _u64, _sz = C.c_void_p, C.c_int, C.c_double, C.c_uint64, C.c_size_t
lib.emu_scene_load.restype = _vp
lib.emu_scene_load.argtypes = [_vp, _sz, _i]
lib.emu_scene_free.argtypes = [_vp]
lib.emu_scene_facts.argtypes = [_vp, _vp]
lib.emu_set_bvh_width.argtypes = [_vp, _i]
lib.emu_closest_hit_ex.argtypes = [_vp, _vp, _sz, _d, _vp, _vp, _vp, _vp]
lib.emu_scatter.argtypes = [_vp, _i, _vp, _sz, _u64, _vp]
lib.emu_sample.argtypes = [_i, _u64, _sz, _vp]
lib.emu_camera_rays.argtypes = [_vp, _i, _i, _i, _u64, _vp]
lib.emu_render.argtypes = [_vp, _vp, _i, _i, _i, _i, _d, _i, _u64, _i, _vp, _vp, _vp, _sz]
lib.emu_wide_stats.argtypes = [_vp, _vp, _sz, _i, _i, _vp]
lib.emu_wide_steps.argtypes = [_vp, _vp, _sz, _i, _i, _vp, _vp]


class EmuScene:
    """Same query methods as raytracing_rust_b200.Scene (closest_hit_ex, scatter, render), run on the host."""

    def __init__(self, scene_bytes: bytes, world: int):
        self._p = lib.emu_scene_load(scene_bytes, len(scene_bytes), world)
        if not self._p:
            raise RuntimeError("emul: scene load failed")
        self.world = world
        self.counters = np.zeros(2, np.uint64)

    def __del__(self):
        p, self._p = getattr(self, "_p", None), None
        if p:
            lib.emu_scene_free(p)

    def facts(self):
        """{'prims': flattened primitives, 'enclosed': FlatScene::enclosed (>= 4 faces of the world box are walls)}"""
        out = (C.c_int * 2)()
        lib.emu_scene_facts(self._p, out)
        return {"prims": out[0], "enclosed": bool(out[1])}

    def set_bvh_width(self, width):
        lib.emu_set_bvh_width(self._p, width)

    def closest_hit_ex(self, world, rays, self_id=None, t_min=1e-3, device=0, want_rec=True):
        r = np.ascontiguousarray(rays, dtype=np.float64)
        n = r.shape[0]
        sid = None if self_id is None else np.ascontiguousarray(self_id, dtype=np.int32)
        prim = np.empty(n, np.int32)
        rec = np.empty((n, 9), np.float32) if want_rec else None
        lib.emu_closest_hit_ex(self._p, r.ctypes.data, n, t_min, sid.ctypes.data if sid is not None else None, prim.ctypes.data,
                               rec.ctypes.data if want_rec else None, self.counters.ctypes.data)
        return prim, rec

    def scatter(self, world, mat, inp, seed=1, device=0):
        q = np.ascontiguousarray(inp, dtype=np.float64)
        out = np.empty((q.shape[0], 10), np.float32)
        lib.emu_scatter(self._p, mat, q.ctypes.data, q.shape[0], seed, out.ctypes.data)
        return out

    def render(self, cam, nx, ny, ns, max_depth=50, t_min=1e-3, background=0, seed=1, nthreads=None, dump_segments=0):
        out = np.empty((ny, nx, 3), np.float32)
        st = np.zeros(3, np.uint64)
        dump = np.zeros((dump_segments, 9), np.float32) if dump_segments else None
        lib.emu_render(self._p, C.addressof(cam), nx, ny, ns, max_depth, t_min, background, seed, nthreads or os.cpu_count() or 1,
                       out.ctypes.data, st.ctypes.data, dump.ctypes.data if dump_segments else None, dump_segments)
        d = {"segments": int(st[0]), "node_tests": int(st[1]), "prim_tests": int(st[2])}
        if dump_segments:
            d["rays"] = dump[: min(dump_segments, d["segments"])]
        return out, d

    def wide_stats(self, rays9, width, mode):
        r = np.ascontiguousarray(rays9, dtype=np.float32)
        out = np.zeros(16, np.uint64)
        lib.emu_wide_stats(self._p, r.ctypes.data, r.shape[0], width, mode, out.ctypes.data)
        return out

    def wide_steps(self, rays9, width, mode):
        """wide_stats plus the node steps every ray took."""
        r = np.ascontiguousarray(rays9, dtype=np.float32)
        out = np.zeros(16, np.uint64)
        steps = np.zeros(r.shape[0], np.uint16)
        lib.emu_wide_steps(self._p, r.ctypes.data, r.shape[0], width, mode, out.ctypes.data, steps.ctypes.data)
        return out, steps


def device_camera_rays(cam, nx, ny, sample=0, seed=1, device=0):
    out = np.empty((nx * ny, 7), np.float32)
    lib.emu_camera_rays(C.addressof(cam), nx, ny, sample, seed, out.ctypes.data)
    return out


def device_sample(kind, seed, n, device=0):
    out = np.empty((n, 3), np.float32)
    lib.emu_sample(kind, seed, n, out.ctypes.data)
    return out
