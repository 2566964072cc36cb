"""raytracing_rust_b200 — Python host mirror of the rtb200 C ABI (include/rtb200.h).

The product is ``librtb200.so`` (C++ scene host + hand-written sm_100a kernels).  This module is a
thin ctypes binding used by the tests and ``bench.py``; it mirrors the reference crate's
constructor names (``Sphere::new`` -> :meth:`Scene.sphere`, ... ; raytracing_rust
``src/*.rs``) and its driver ``create_image`` (``tests/test.rs:55-85``).

There is no CPU fallback: importing without the built library raises, and every render call
fails loudly when no CUDA device is usable.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
# RTB200_LIB: developer override used by tools/ to A/B experiment builds of the same library
_LIB_PATH = Path(os.environ["RTB200_LIB"]).resolve() if os.environ.get("RTB200_LIB") else _HERE / "librtb200.so"
ASSETS = _HERE.parent / "assets"

if not _LIB_PATH.exists():
    raise ImportError(
        f"{_LIB_PATH} is missing: build it with `make` (or `python -c 'import __graft_entry__ as g; g.build()'`). "
        "rtb200 has no CPU fallback."
    )
lib = C.CDLL(str(_LIB_PATH))

PLANE_YZ, PLANE_ZX, PLANE_XY = 0, 1, 2
AXIS_X, AXIS_Y, AXIS_Z = 0, 1, 2
BG_BLACK, BG_SKY = 0, 1
BVH_AUTO, BVH_HOST_SAH, BVH_DEVICE_LBVH = 0, 1, 2
PIPE_AUTO, PIPE_MEGAKERNEL, PIPE_WAVEFRONT, PIPE_POOL, PIPE_DUAL = 0, 1, 2, 3, 4
SHARD_SAMPLES, SHARD_BANDS = 0, 1
BAND_ROWS = 4
NOALLOC_MAX_PRIM_IDS = 65536

NAMED_SCENES = (
    "random_scene", "two_spheres", "two_perlin_sphere", "earth", "simple_light",
    "cornell_box", "cornell_smoke", "final_scene",
)
LIT_VARIANTS = ("final_scene_lit", "cornell_smoke_book")


class RtbCamera(C.Structure):
    _fields_ = [
        ("look_from", C.c_double * 3), ("look_at", C.c_double * 3), ("view_up", C.c_double * 3),
        ("vertical_fov", C.c_double), ("aspect", C.c_double), ("aperture", C.c_double),
        ("focus_dist", C.c_double), ("time0", C.c_double), ("time1", C.c_double),
    ]


class RenderOpts(C.Structure):
    _fields_ = [
        ("max_depth", C.c_int), ("t_min", C.c_double), ("background", C.c_int), ("seed", C.c_uint64),
        ("sample_begin", C.c_int), ("sample_stride", C.c_int), ("pipeline", C.c_int), ("device", C.c_int),
        ("collect_stats", C.c_int), ("band_begin", C.c_int), ("band_stride", C.c_int), ("multi_shard", C.c_int),
        ("reserved", C.c_int * 7),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("paths", C.c_uint64), ("segments", C.c_uint64), ("node_tests", C.c_uint64),
        ("prim_tests", C.c_uint64), ("medium_tests", C.c_uint64), ("shade", C.c_uint64 * 5),
        ("tex_evals", C.c_uint64 * 4), ("kernel_launches", C.c_uint64), ("device_ms", C.c_double),
        ("reserved", C.c_uint64 * 8),
    ]

    def as_dict(self):
        return {
            "paths": self.paths, "segments": self.segments, "node_tests": self.node_tests,
            "prim_tests": self.prim_tests, "medium_tests": self.medium_tests, "shade": list(self.shade),
            "tex_evals": list(self.tex_evals), "kernel_launches": self.kernel_launches,
            "device_ms": self.device_ms,
        }


def _sig(name, restype, *argtypes):
    f = getattr(lib, name)
    f.restype = restype
    f.argtypes = list(argtypes)
    return f


_vp, _i, _d, _u64, _sz = C.c_void_p, C.c_int, C.c_double, C.c_uint64, C.c_size_t
_dp, _u8p = C.POINTER(C.c_double), C.POINTER(C.c_uint8)

_sig("rtb_last_error", C.c_char_p)
_sig("rtb_version", C.c_char_p)
_sig("rtb_device_count", _i)
_sig("rtb_scene_create", _vp)
_sig("rtb_scene_destroy", None, _vp)
_sig("rtb_tex_solid", _i, _vp, _d, _d, _d)
_sig("rtb_tex_checker", _i, _vp, _i, _i)
_sig("rtb_tex_noise", _i, _vp, _d, _u64)
_sig("rtb_tex_noise_tables", _i, _vp, _d, _vp, _vp, _vp, _vp)
_sig("rtb_tex_image", _i, _vp, _vp, C.c_uint32, C.c_uint32)
_sig("rtb_mat_lambertian", _i, _vp, _i)
_sig("rtb_mat_metal", _i, _vp, _i, _d)
_sig("rtb_mat_dielectric", _i, _vp, _d)
_sig("rtb_mat_diffuse_light", _i, _vp, _i)
_sig("rtb_mat_isotropic", _i, _vp, _i)
_sig("rtb_obj_sphere", _i, _vp, _dp, _d, _i)
_sig("rtb_obj_moving_sphere", _i, _vp, _dp, _dp, _d, _d, _d, _i)
_sig("rtb_obj_rect", _i, _vp, _i, _d, _d, _d, _d, _d, _i)
_sig("rtb_obj_cube", _i, _vp, _dp, _dp, _i)
_sig("rtb_obj_flip_normals", _i, _vp, _i)
_sig("rtb_obj_rotate", _i, _vp, _i, _i, _d)
_sig("rtb_obj_translate", _i, _vp, _i, _dp)
_sig("rtb_obj_constant_medium", _i, _vp, _i, _d, _i)
_sig("rtb_obj_list", _i, _vp, C.POINTER(C.c_int), _i)
_sig("rtb_obj_bvh", _i, _vp, C.POINTER(C.c_int), _i, _d, _d, _u64)
_sig("rtb_scene_named", _i, _vp, C.c_char_p, _u64, _vp, C.c_uint32, C.c_uint32, _i, _i, C.POINTER(RtbCamera))
_sig("rtb_scene_serialized_size", _sz, _vp)
_sig("rtb_scene_serialize", _i, _vp, _vp, _sz)
_sig("rtb_scene_deserialize", _vp, _vp, _sz)
_sig("rtb_scene_prim_count", _i, _vp, _i)
_sig("rtb_scene_set_bvh_builder", _i, _vp, _i)
_sig("rtb_render_opts_default", None, C.POINTER(RenderOpts))
_sig("rtb_render", _i, _vp, _i, C.POINTER(RtbCamera), _i, _i, _i, C.POINTER(RenderOpts), _vp, _vp, C.POINTER(Stats))
_sig("rtb_render_multi", _i, _vp, _i, C.POINTER(RtbCamera), _i, _i, _i, C.POINTER(RenderOpts), C.POINTER(C.c_int), _i, _vp, _vp,
     C.POINTER(Stats))
_sig("rtb_philox4x32_10", _i, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32))
_sig("rtb_issue_peak", _i, _i, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double))
PROGRESS_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_uint8), C.c_double, C.c_double)
_sig("rtb_render_progressive", _i, _vp, _i, C.POINTER(RtbCamera), _i, _i, _i, C.POINTER(RenderOpts), _i, _i, PROGRESS_FN, _vp,
     _vp, _vp, C.POINTER(C.c_int))
_sig("rtb_scene_upload", _i, _vp, _i, _i)
_sig("rtb_scene_device_bytes", _sz, _vp)
_sig("rtb_render_device", _i, _vp, C.POINTER(RtbCamera), _i, _i, _i, C.POINTER(RenderOpts), _vp, _vp, C.POINTER(Stats))
_sig("rtb_resolve_device", _i, _vp, _i, _i, _i, _vp, _vp)
_sig("rtb_device_alloc", _i, _sz, _i, C.POINTER(C.c_void_p))
_sig("rtb_device_free", _i, _vp)
_sig("rtb_peer_export", _i, _vp, _vp)
_sig("rtb_peer_open", _i, _vp, _i, C.POINTER(C.c_void_p))
_sig("rtb_peer_close", _i, _vp)
_sig("rtb_reduce_resolve_device", _i, C.POINTER(C.c_void_p), _i, _i, _i, _i, _vp, _vp, _vp)
_sig("rtb_reduce_resolve_range_device", _i, C.POINTER(C.c_void_p), _i, _sz, _sz, _i, _vp, _vp, _vp)
_sig("rtb_resolve_bands_device", _i, _vp, _i, _i, _i, _i, _i, _vp, _i, _vp)
_sig("rtb_closest_hit", _i, _vp, _i, _vp, _sz, _d, _i, _vp, _vp)
_sig("rtb_debug_texture_value", _i, _vp, _i, _i, _vp, _sz, _i, _vp)
_sig("rtb_closest_hit_ex", _i, _vp, _i, _vp, _sz, _d, _vp, _i, _vp, _vp)
_sig("rtb_debug_camera_rays", _i, C.POINTER(RtbCamera), _i, _i, _i, _u64, _i, _vp)
_sig("rtb_debug_camera_draws", _i, _i, _i, _i, _u64, _vp)
_sig("rtb_debug_scatter", _i, _vp, _i, _i, _vp, _sz, _u64, _i, _vp)
_sig("rtb_debug_sample", _i, _i, _u64, _sz, _i, _vp)
_sig("rtb_closest_hit_noalloc", _i, _vp, _vp, _sz, _d, _vp, _vp, _vp, _vp)
_sig("rtb_debug_check_failures", _i, _i, C.POINTER(C.c_uint32))
_sig("rtb_primary_rays", _i, C.POINTER(RtbCamera), _i, _i, _i, _u64, _vp)
_sig("rtb_ppm_p3", _sz, _vp, _i, _i, _vp, _sz)
_sig("rtb_ppm_p6", _sz, _vp, _i, _i, _vp, _sz)
_sig("rtb_image_write_ppm", _i, C.c_char_p, _vp, _i, _i, _i)
_sig("rtb_release_workspace", _i)
_sig("rtb_image_write_png", _i, C.c_char_p, _vp, _i, _i)
_sig("rtb_jpeg_decode", _i, _vp, _sz, C.POINTER(C.c_void_p), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32))
_sig("rtb_jpeg_decode_file", _i, C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32))
_sig("rtb_image_free", None, _vp)
_sig("rtb_tex_image_file", _i, _vp, C.c_char_p)
_sig("rtb_scene_save", _i, _vp, C.c_char_p)
_sig("rtb_scene_load", _vp, C.c_char_p)


class RtbError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        super().__init__(f"{where}: [{code}] {lib.rtb_last_error().decode()}")


def _h(code, where):
    if code < 0:
        raise RtbError(code, where)
    return code


def _vec3(v):
    return (C.c_double * 3)(*[float(x) for x in v])


def philox4x32_10(counter, key):
    """One Philox4x32-10 block on the host (rtb_philox4x32_10): counter = 4 x u32, key = 2 x u32 -> 4 x u32."""
    c, k, o = (C.c_uint32 * 4)(*counter), (C.c_uint32 * 2)(*key), (C.c_uint32 * 4)()
    _h(lib.rtb_philox4x32_10(c, k, o), "philox4x32_10")
    return tuple(int(x) for x in o)


def issue_peak(device=0):
    """(FFMA only, FFMA+LOP3 1:1, FFMA+FMNMX 1:2) lane-instructions per second in 1e12 on `device` (rtb_issue_peak)."""
    a, b, c = C.c_double(0), C.c_double(0), C.c_double(0)
    _h(lib.rtb_issue_peak(device, C.byref(a), C.byref(b), C.byref(c)), "issue_peak")
    return a.value, b.value, c.value


def device_count() -> int:
    return lib.rtb_device_count()


def earth_rgb8() -> np.ndarray:
    """Decoded texture/earthmap.jpg (1024x512 RGB8), staged in assets/ (SURVEY F9)."""
    a = np.fromfile(ASSETS / "earthmap_1024x512.rgb8", dtype=np.uint8)
    return a.reshape(512, 1024, 3)


def default_opts(**kw) -> RenderOpts:
    o = RenderOpts()
    lib.rtb_render_opts_default(C.byref(o))
    for k, v in kw.items():
        setattr(o, k, v)
    return o


def camera(look_from, look_at, view_up, vertical_fov, aspect, aperture, focus_dist, time0, time1) -> RtbCamera:
    """Camera::new, src/camera.rs:21-31 (same argument order)."""
    c = RtbCamera()
    c.look_from[:] = look_from
    c.look_at[:] = look_at
    c.view_up[:] = view_up
    c.vertical_fov, c.aspect, c.aperture = vertical_fov, aspect, aperture
    c.focus_dist, c.time0, c.time1 = focus_dist, time0, time1
    return c


class Scene:
    """One rtb_scene.  Builder methods return handles and mirror the crate's constructors."""

    def __init__(self, _ptr=None):
        self._p = _ptr if _ptr is not None else lib.rtb_scene_create()
        if not self._p:
            raise RtbError(-1, "scene_create")
        self._keep = []

    def __del__(self):
        p, self._p = getattr(self, "_p", None), None
        if p and lib is not None:  # `lib` is already None when the interpreter tears the module down
            lib.rtb_scene_destroy(p)

    # -- textures (src/texture.rs)
    def solid(self, r, g, b):
        return _h(lib.rtb_tex_solid(self._p, r, g, b), "tex_solid")

    def checker(self, odd, even):
        return _h(lib.rtb_tex_checker(self._p, odd, even), "tex_checker")

    def noise(self, scale, seed=0x5EED):
        return _h(lib.rtb_tex_noise(self._p, scale, seed), "tex_noise")

    def image(self, rgb8: np.ndarray):
        a = np.ascontiguousarray(rgb8, dtype=np.uint8)
        ny, nx = a.shape[0], a.shape[1]
        return _h(lib.rtb_tex_image(self._p, a.ctypes.data, nx, ny), "tex_image")

    def image_file(self, path):
        """ImageTexture::new(image::open(path).to_rgb8().into_raw(), nx, ny), tests/test.rs:201-205"""
        return _h(lib.rtb_tex_image_file(self._p, str(path).encode()), "tex_image_file")

    # -- materials (src/material.rs)
    def lambertian(self, tex):
        return _h(lib.rtb_mat_lambertian(self._p, tex), "mat_lambertian")

    def metal(self, tex, fuzz):
        return _h(lib.rtb_mat_metal(self._p, tex, fuzz), "mat_metal")

    def dielectric(self, ref_idx):
        return _h(lib.rtb_mat_dielectric(self._p, ref_idx), "mat_dielectric")

    def diffuse_light(self, tex):
        return _h(lib.rtb_mat_diffuse_light(self._p, tex), "mat_diffuse_light")

    def isotropic(self, tex):
        return _h(lib.rtb_mat_isotropic(self._p, tex), "mat_isotropic")

    # -- hittables
    def sphere(self, center, radius, mat):
        return _h(lib.rtb_obj_sphere(self._p, _vec3(center), radius, mat), "obj_sphere")

    def moving_sphere(self, c0, c1, t0, t1, radius, mat):
        return _h(lib.rtb_obj_moving_sphere(self._p, _vec3(c0), _vec3(c1), t0, t1, radius, mat), "obj_moving_sphere")

    def rect(self, plane, x0, y0, x1, y1, k, mat):
        return _h(lib.rtb_obj_rect(self._p, plane, x0, y0, x1, y1, k, mat), "obj_rect")

    def cube(self, p_min, p_max, mat):
        return _h(lib.rtb_obj_cube(self._p, _vec3(p_min), _vec3(p_max), mat), "obj_cube")

    def flip_normals(self, obj):
        return _h(lib.rtb_obj_flip_normals(self._p, obj), "obj_flip_normals")

    def rotate(self, axis, obj, angle_deg):
        return _h(lib.rtb_obj_rotate(self._p, axis, obj, angle_deg), "obj_rotate")

    def translate(self, obj, offset):
        return _h(lib.rtb_obj_translate(self._p, obj, _vec3(offset)), "obj_translate")

    def constant_medium(self, boundary, density, tex):
        return _h(lib.rtb_obj_constant_medium(self._p, boundary, density, tex), "obj_constant_medium")

    def list(self, objs):
        arr = (C.c_int * len(objs))(*objs)
        return _h(lib.rtb_obj_list(self._p, arr, len(objs)), "obj_list")

    def bvh(self, objs, t0=0.0, t1=1.0, seed=0):
        arr = (C.c_int * len(objs))(*objs)
        return _h(lib.rtb_obj_bvh(self._p, arr, len(objs), t0, t1, seed), "obj_bvh")

    # -- named scenes (tests/test.rs:89-523)
    @classmethod
    def named(cls, name, nx, ny, seed=0x5EED):
        """Returns (scene, world_handle, camera) for one of NAMED_SCENES / LIT_VARIANTS."""
        s = cls()
        cam = RtbCamera()
        e = earth_rgb8() if name in ("earth", "final_scene", "final_scene_lit") else None
        ep = e.ctypes.data if e is not None else None
        w = _h(lib.rtb_scene_named(s._p, name.encode(), seed, ep, 1024 if e is not None else 0,
                                   512 if e is not None else 0, nx, ny, C.byref(cam)), f"scene_named({name})")
        return s, w, cam

    # -- serialisation
    def serialize(self) -> bytes:
        n = lib.rtb_scene_serialized_size(self._p)
        buf = C.create_string_buffer(n)
        _h(lib.rtb_scene_serialize(self._p, buf, n), "scene_serialize")
        return buf.raw

    @classmethod
    def deserialize(cls, data: bytes):
        p = lib.rtb_scene_deserialize(data, len(data))
        if not p:
            raise RtbError(-1, "scene_deserialize")
        return cls(_ptr=p)

    def save(self, path):
        """Writes the scene as a .rtscene file (include/rtb_scene_format.h)."""
        _h(lib.rtb_scene_save(self._p, str(path).encode()), "scene_save")

    @classmethod
    def load(cls, path):
        p = lib.rtb_scene_load(str(path).encode())
        if not p:
            raise RtbError(-1, "scene_load")
        return cls(_ptr=p)

    def set_bvh_builder(self, builder: int):
        """BVH_AUTO / BVH_HOST_SAH / BVH_DEVICE_LBVH; takes effect at the next upload."""
        _h(lib.rtb_scene_set_bvh_builder(self._p, builder), "scene_set_bvh_builder")

    def prim_count(self, world):
        return _h(lib.rtb_scene_prim_count(self._p, world), "scene_prim_count")

    # -- rendering
    def render(self, world, cam, nx, ny, ns, opts=None, want_sum=True, want_rgb8=True, want_stats=False):
        """create_image (tests/test.rs:55-85) with host buffers.  Returns (sum_rgb float32
        [ny,nx,3] or None, rgb8 uint8 [ny,nx,3] or None, stats dict or None)."""
        o = opts if opts is not None else default_opts()
        out_sum = np.empty((ny, nx, 3), np.float32) if want_sum else None
        out_rgb8 = np.empty((ny, nx, 3), np.uint8) if want_rgb8 else None
        st = Stats() if want_stats else None
        _h(lib.rtb_render(self._p, world, C.byref(cam), nx, ny, ns, C.byref(o),
                          out_sum.ctypes.data if want_sum else None, out_rgb8.ctypes.data if want_rgb8 else None,
                          C.byref(st) if want_stats else None), "render")
        return out_sum, out_rgb8, (st.as_dict() if want_stats else None)

    def render_multi(self, world, cam, nx, ny, ns, devices=None, opts=None, want_sum=True, want_rgb8=True, want_stats=False):
        """create_image on several GPUs of this node from one call (rtb_render_multi): samples sharded over
        `devices` (list of ordinals, an int = the first n, None = all), partial sums combined on devices[0]."""
        o = opts if opts is not None else default_opts()
        out_sum = np.empty((ny, nx, 3), np.float32) if want_sum else None
        out_rgb8 = np.empty((ny, nx, 3), np.uint8) if want_rgb8 else None
        st = Stats() if want_stats else None
        if devices is None or isinstance(devices, int):
            arr, n = None, (devices or 0)
        else:
            arr, n = (C.c_int * len(devices))(*devices), len(devices)
        _h(lib.rtb_render_multi(self._p, world, C.byref(cam), nx, ny, ns, C.byref(o), arr, n,
                                out_sum.ctypes.data if want_sum else None, out_rgb8.ctypes.data if want_rgb8 else None,
                                C.byref(st) if want_stats else None), "render_multi")
        d = st.as_dict() if want_stats else None
        if d is not None:
            d["devices"] = int(st.reserved[0])
        return out_sum, out_rgb8, d

    def render_progressive(self, world, cam, nx, ny, ns, chunk_spp, on_progress=None, preview_every=0, opts=None, want_sum=False):
        """create_image in chunks with real progress (rtb_render_progressive).  on_progress(done, total,
        preview rgb8 [ny,nx,3] or None, mpaths_per_s, eta_s) -> truthy to stop.  Returns (sum or None, rgb8, done)."""
        o = opts if opts is not None else default_opts()
        out_sum = np.empty((ny, nx, 3), np.float32) if want_sum else None
        out_rgb8 = np.empty((ny, nx, 3), np.uint8)
        done = C.c_int(0)

        def _cb(_user, d, total, img, rate, eta):
            if on_progress is None:
                return 0
            prev = np.ctypeslib.as_array(img, shape=(ny, nx, 3)).copy() if img else None
            return 1 if on_progress(d, total, prev, rate, eta) else 0

        cb = PROGRESS_FN(_cb)
        _h(lib.rtb_render_progressive(self._p, world, C.byref(cam), nx, ny, ns, C.byref(o), chunk_spp, preview_every, cb, None,
                                      out_sum.ctypes.data if want_sum else None, out_rgb8.ctypes.data, C.byref(done)),
           "render_progressive")
        return out_sum, out_rgb8, done.value

    def upload(self, world, device=0):
        _h(lib.rtb_scene_upload(self._p, world, device), "scene_upload")

    def device_bytes(self) -> int:
        return lib.rtb_scene_device_bytes(self._p)

    def render_device(self, cam, nx, ny, ns, opts, d_sum_ptr: int, stream: int = 0, want_stats=False):
        st = Stats() if want_stats else None
        _h(lib.rtb_render_device(self._p, C.byref(cam), nx, ny, ns, C.byref(opts), d_sum_ptr, stream,
                                 C.byref(st) if want_stats else None), "render_device")
        return st.as_dict() if want_stats else None

    def closest_hit(self, world, rays: np.ndarray, t_min=1e-3, device=0):
        """rays: [n,7] float64 {o, d, time}.  Returns (prim int32 [n], t float32 [n])."""
        r = np.ascontiguousarray(rays, dtype=np.float64)
        n = r.shape[0]
        prim = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        _h(lib.rtb_closest_hit(self._p, world, r.ctypes.data, n, t_min, device, prim.ctypes.data, t.ctypes.data),
           "closest_hit")
        return prim, t

    def closest_hit_ex(self, world, rays: np.ndarray, self_id=None, t_min=1e-3, device=0, want_rec=True):
        """rtb_closest_hit_ex: rays [n,7] float64; self_id [n] int32 primitive ID each ray leaves (-1 none) or None.
        Returns (prim int32 [n], rec float32 [n,9] {t,u,v,p,normal} or None)."""
        r = np.ascontiguousarray(rays, dtype=np.float64)
        n = r.shape[0]
        sid = None if self_id is None else np.ascontiguousarray(self_id, dtype=np.int32)
        prim = np.empty(n, np.int32)
        rec = np.empty((n, 9), np.float32) if want_rec else None
        _h(lib.rtb_closest_hit_ex(self._p, world, r.ctypes.data, n, t_min, sid.ctypes.data if sid is not None else None,
                                  device, prim.ctypes.data, rec.ctypes.data if want_rec else None), "closest_hit_ex")
        return prim, rec

    def closest_hit_noalloc(self, d_rays_ptr: int, n: int, d_out_prim_ptr: int, d_out_rec_ptr: int = 0, d_self_id_ptr: int = 0,
                            t_min=1e-3, stream: int = 0):
        """rtb_closest_hit_noalloc: device pointers in and out, no allocation inside the call; the scene must be uploaded."""
        _h(lib.rtb_closest_hit_noalloc(self._p, d_rays_ptr, n, t_min, d_self_id_ptr or None, d_out_prim_ptr, d_out_rec_ptr or None,
                                       stream or None), "closest_hit_noalloc")

    def scatter(self, world, mat, inp: np.ndarray, seed=1, device=0):
        """rtb_debug_scatter: inp [n,16] {ray o,d,time, hit t,u,v,p,normal}; returns [n,10] {ok, o, d, atten}."""
        q = np.ascontiguousarray(inp, dtype=np.float64)
        out = np.empty((q.shape[0], 10), np.float32)
        _h(lib.rtb_debug_scatter(self._p, world, mat, q.ctypes.data, q.shape[0], seed, device, out.ctypes.data), "debug_scatter")
        return out

    def texture_value(self, world, tex, uvp: np.ndarray, device=0):
        q = np.ascontiguousarray(uvp, dtype=np.float64)
        out = np.empty((q.shape[0], 3), np.float32)
        _h(lib.rtb_debug_texture_value(self._p, world, tex, q.ctypes.data, q.shape[0], device, out.ctypes.data),
           "texture_value")
        return out


def resolve_device(d_sum_ptr: int, nx, ny, ns, d_rgb8_ptr: int, stream: int = 0):
    _h(lib.rtb_resolve_device(d_sum_ptr, nx, ny, ns, d_rgb8_ptr, stream), "resolve_device")


IPC_HANDLE_BYTES = 64
MAX_PEER_BUFFERS = 16


class DeviceBuffer:
    """A zeroed device allocation owned by librtb200 (rtb_device_alloc): a whole CUDA allocation,
    so it can be exported to the other ranks (peer_handle) for the fused reduce + resolve.
    `__cuda_array_interface__` lets torch / cupy view it without a copy."""

    def __init__(self, shape, dtype=np.float32, device=0):
        self.shape, self.dtype, self.device = tuple(int(x) for x in shape), np.dtype(dtype), device
        self.nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        p = C.c_void_p()
        _h(lib.rtb_device_alloc(self.nbytes, device, C.byref(p)), "device_alloc")
        self.ptr = p.value

    @property
    def __cuda_array_interface__(self):
        return {"shape": self.shape, "typestr": self.dtype.str, "data": (self.ptr, False), "version": 3, "strides": None}

    def peer_handle(self) -> bytes:
        h = C.create_string_buffer(IPC_HANDLE_BYTES)
        _h(lib.rtb_peer_export(self.ptr, h), "peer_export")
        return h.raw

    def free(self):
        if self.ptr:
            _h(lib.rtb_device_free(self.ptr), "device_free")
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def peer_open(handle: bytes, device: int) -> int:
    """Maps a peer rank's DeviceBuffer into this process; returns the device pointer."""
    p = C.c_void_p()
    _h(lib.rtb_peer_open(C.create_string_buffer(handle, IPC_HANDLE_BYTES), device, C.byref(p)), "peer_open")
    return p.value


def peer_close(ptr: int):
    _h(lib.rtb_peer_close(ptr), "peer_close")


def reduce_resolve_device(sum_ptrs, nx, ny, ns, d_rgb8_ptr: int, d_total_ptr: int = 0, stream: int = 0):
    """One kernel: sum of the accumulation buffers in `sum_ptrs` (device pointers, own or
    peer-mapped, added in list order) -> quantised rgb8 (and the f32 total if d_total_ptr)."""
    arr = (C.c_void_p * len(sum_ptrs))(*sum_ptrs)
    _h(lib.rtb_reduce_resolve_device(arr, len(sum_ptrs), nx, ny, ns, d_total_ptr or None, d_rgb8_ptr, stream or None),
       "reduce_resolve_device")


def reduce_resolve_range_device(sum_ptrs, first_value, n_values, ns, d_rgb8_ptr: int, d_total_ptr: int = 0, stream: int = 0):
    """The sharded exchange: sum + quantise values [first_value, first_value + n_values) of the full-size buffers in
    `sum_ptrs` (base pointers) into the full-size rgb8 image at d_rgb8_ptr (base pointer, own or peer-mapped)."""
    arr = (C.c_void_p * len(sum_ptrs))(*sum_ptrs)
    _h(lib.rtb_reduce_resolve_range_device(arr, len(sum_ptrs), first_value, n_values, ns, d_total_ptr or None, d_rgb8_ptr,
                                           stream or None), "reduce_resolve_range_device")


def resolve_bands_device(d_sum_ptr: int, nx, ny, ns, band_begin, band_stride, d_rgb8_ptr: int, compact=False, stream: int = 0):
    """Quantise this shard's 4-row bands of a full-size accumulation buffer (row-band sharding)."""
    _h(lib.rtb_resolve_bands_device(d_sum_ptr, nx, ny, ns, band_begin, band_stride, d_rgb8_ptr, int(bool(compact)), stream or None),
       "resolve_bands_device")


def primary_rays(cam, nx, ny, sample=0, seed=1) -> np.ndarray:
    out = np.empty((nx * ny, 7), np.float64)
    _h(lib.rtb_primary_rays(C.byref(cam), nx, ny, sample, seed, out.ctypes.data), "primary_rays")
    return out


def device_camera_rays(cam, nx, ny, sample=0, seed=1, device=0) -> np.ndarray:
    """Camera::get_ray as the render kernels evaluate it (device f32): [nx*ny, 7] float32."""
    out = np.empty((nx * ny, 7), np.float32)
    _h(lib.rtb_debug_camera_rays(C.byref(cam), nx, ny, sample, seed, device, out.ctypes.data), "debug_camera_rays")
    return out


def camera_draws(nx, ny, sample=0, seed=1) -> np.ndarray:
    """The draws behind primary_rays / device_camera_rays as explicit get_ray inputs: [nx*ny, 5] {s,t,rd.x,rd.y,xi_time}."""
    out = np.empty((nx * ny, 5), np.float64)
    _h(lib.rtb_debug_camera_draws(nx, ny, sample, seed, out.ctypes.data), "debug_camera_draws")
    return out


def check_failures(device=0):
    """Self-check counters of an RTB_DEBUG_CHECKS build (read and cleared); raises RtbError on a release build."""
    out = (C.c_uint32 * 16)()
    _h(lib.rtb_debug_check_failures(device, out), "debug_check_failures")
    names = ["stack", "node", "prim", "mat", "tex", "texel", "tile", "pixel", "queue", "nan"]
    return {n: int(out[i]) for i, n in enumerate(names)}


def device_sample(kind, seed, n, device=0) -> np.ndarray:
    """Device samplers (0 = unit ball, 1 = unit disk): [n, 3] float32."""
    out = np.empty((n, 3), np.float32)
    _h(lib.rtb_debug_sample(kind, seed, n, device, out.ctypes.data), "debug_sample")
    return out


def ppm_p3(rgb8: np.ndarray) -> bytes:
    a = np.ascontiguousarray(rgb8, dtype=np.uint8)
    ny, nx = a.shape[0], a.shape[1]
    n = lib.rtb_ppm_p3(a.ctypes.data, nx, ny, None, 0)
    buf = C.create_string_buffer(n + 1)
    lib.rtb_ppm_p3(a.ctypes.data, nx, ny, buf, n + 1)
    return buf.raw[:n]


def ppm_p6(rgb8: np.ndarray) -> bytes:
    a = np.ascontiguousarray(rgb8, dtype=np.uint8)
    ny, nx = a.shape[0], a.shape[1]
    n = lib.rtb_ppm_p6(a.ctypes.data, nx, ny, None, 0)
    buf = C.create_string_buffer(n)
    lib.rtb_ppm_p6(a.ctypes.data, nx, ny, buf, n)
    return buf.raw[:n]


def write_ppm(path, rgb8: np.ndarray, binary=False):
    """File::create + write! of the reference's #[test] wrappers (tests/test.rs:536,557); P3 text or P6."""
    a = np.ascontiguousarray(rgb8, dtype=np.uint8)
    _h(lib.rtb_image_write_ppm(str(path).encode(), a.ctypes.data, a.shape[1], a.shape[0], int(bool(binary))), "image_write_ppm")


def write_png(path, rgb8: np.ndarray):
    a = np.ascontiguousarray(rgb8, dtype=np.uint8)
    _h(lib.rtb_image_write_png(str(path).encode(), a.ctypes.data, a.shape[1], a.shape[0]), "image_write_png")


def decode_jpeg(src) -> np.ndarray:
    """image::open(path).to_rgb8() (tests/test.rs:201-205): `src` is a path or the file's bytes; (ny, nx, 3) uint8."""
    out, nx, ny = C.c_void_p(), C.c_uint32(), C.c_uint32()
    if isinstance(src, (bytes, bytearray, memoryview)):
        b = bytes(src)
        _h(lib.rtb_jpeg_decode(b, len(b), C.byref(out), C.byref(nx), C.byref(ny)), "jpeg_decode")
    else:
        _h(lib.rtb_jpeg_decode_file(str(src).encode(), C.byref(out), C.byref(nx), C.byref(ny)), "jpeg_decode_file")
    try:
        a = np.ctypeslib.as_array(C.cast(out, C.POINTER(C.c_uint8)), shape=(ny.value, nx.value, 3)).copy()
    finally:
        lib.rtb_image_free(out)
    return a


def create_image(ny, nx, ns, cam, scene: Scene, world: int, opts=None) -> bytes:
    """create_image(ny, nx, ns, cam, world) -> String (tests/test.rs:55): same argument order,
    same P3 text; the pixel loop runs on the GPU."""
    _, rgb8, _ = scene.render(world, cam, nx, ny, ns, opts, want_sum=False)
    return ppm_p3(rgb8)
