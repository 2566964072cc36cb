//! rtb200-sys — mirrors include/rtb200.h (C ABI of librtb200.so).  Every call returns an `int`
//! status or handle (negative = error, text in `rtb_last_error()`); none of them unwinds.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_double, c_int, c_void};

#[repr(C)]
pub struct rtb_scene {
    _private: [u8; 0],
}

pub const RTB_OK: c_int = 0;
pub const RTB_ERR_INVALID: c_int = -1;
pub const RTB_ERR_NO_DEVICE: c_int = -2;

pub const RTB_PLANE_YZ: c_int = 0; // src/rect.rs:8-12
pub const RTB_PLANE_ZX: c_int = 1;
pub const RTB_PLANE_XY: c_int = 2;
pub const RTB_AXIS_X: c_int = 0; // src/rotate.rs:7-11
pub const RTB_AXIS_Y: c_int = 1;
pub const RTB_AXIS_Z: c_int = 2;
pub const RTB_BG_BLACK: c_int = 0; // src/color.rs:21 as written
pub const RTB_BG_SKY: c_int = 1; // the gradient commented out at src/color.rs:18-20

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct RtbCamera {
    // Camera::new arguments, src/camera.rs:21-31
    pub look_from: [f64; 3],
    pub look_at: [f64; 3],
    pub view_up: [f64; 3],
    pub vertical_fov: f64,
    pub aspect: f64,
    pub aperture: f64,
    pub focus_dist: f64,
    pub time0: f64,
    pub time1: f64,
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rtb_render_opts {
    pub max_depth: c_int, // 50, src/color.rs:9
    pub t_min: f64,       // 1e-3, src/color.rs:7
    pub background: c_int,
    pub seed: u64,
    pub sample_begin: c_int,
    pub sample_stride: c_int,
    pub pipeline: c_int,
    pub device: c_int,
    pub collect_stats: c_int,
    pub band_begin: c_int,  // row-band sharding: 4-row bands band_begin + i*band_stride (tests/test.rs:62-63 split)
    pub band_stride: c_int,
    pub multi_shard: c_int, // rtb_render_multi: RTB_SHARD_SAMPLES | RTB_SHARD_BANDS
    pub reserved: [c_int; 7],
}
pub const RTB_SHARD_SAMPLES: c_int = 0;
pub const RTB_SHARD_BANDS: c_int = 1;
pub const RTB_BAND_ROWS: c_int = 4;

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rtb_stats {
    pub paths: u64,
    pub segments: u64,
    pub node_tests: u64,
    pub prim_tests: u64,
    pub medium_tests: u64,
    pub shade: [u64; 5],
    pub tex_evals: [u64; 4],
    pub kernel_launches: u64,
    pub device_ms: f64,
    pub reserved: [u64; 8],
}

pub type rtb_progress_fn = Option<
    unsafe extern "C" fn(user: *mut c_void, samples_done: c_int, samples_total: c_int, rgb8: *const u8, mpaths_per_s: c_double, eta_seconds: c_double) -> c_int,
>;

extern "C" {
    pub fn rtb_last_error() -> *const c_char;
    pub fn rtb_version() -> *const c_char;
    pub fn rtb_device_count() -> c_int;
    pub fn rtb_scene_create() -> *mut rtb_scene;
    pub fn rtb_scene_destroy(s: *mut rtb_scene);
    // textures, src/texture.rs
    pub fn rtb_tex_solid(s: *mut rtb_scene, r: c_double, g: c_double, b: c_double) -> c_int; // :14
    pub fn rtb_tex_checker(s: *mut rtb_scene, odd: c_int, even: c_int) -> c_int; // :34
    pub fn rtb_tex_noise(s: *mut rtb_scene, scale: c_double, seed: u64) -> c_int; // :57
    pub fn rtb_tex_noise_tables(s: *mut rtb_scene, scale: c_double, ranvec: *const f64, perm_x: *const u8, perm_y: *const u8, perm_z: *const u8) -> c_int;
    pub fn rtb_tex_image(s: *mut rtb_scene, rgb8: *const u8, nx: u32, ny: u32) -> c_int; // :81
    // materials, src/material.rs
    pub fn rtb_mat_lambertian(s: *mut rtb_scene, albedo_tex: c_int) -> c_int; // :41
    pub fn rtb_mat_metal(s: *mut rtb_scene, albedo_tex: c_int, fuzz: c_double) -> c_int; // :67
    pub fn rtb_mat_dielectric(s: *mut rtb_scene, ref_idx: c_double) -> c_int; // :100
    pub fn rtb_mat_diffuse_light(s: *mut rtb_scene, emit_tex: c_int) -> c_int; // :138
    pub fn rtb_mat_isotropic(s: *mut rtb_scene, albedo_tex: c_int) -> c_int; // :159
    // hittables
    pub fn rtb_obj_sphere(s: *mut rtb_scene, center: *const f64, radius: c_double, mat: c_int) -> c_int; // src/sphere.rs:27
    pub fn rtb_obj_moving_sphere(s: *mut rtb_scene, c0: *const f64, c1: *const f64, t0: c_double, t1: c_double, radius: c_double, mat: c_int) -> c_int; // :97
    pub fn rtb_obj_rect(s: *mut rtb_scene, plane: c_int, x0: c_double, y0: c_double, x1: c_double, y1: c_double, k: c_double, mat: c_int) -> c_int; // src/rect.rs:25
    pub fn rtb_obj_cube(s: *mut rtb_scene, p_min: *const f64, p_max: *const f64, mat: c_int) -> c_int; // src/cube.rs:15
    pub fn rtb_obj_flip_normals(s: *mut rtb_scene, obj: c_int) -> c_int; // src/hittable.rs:72
    pub fn rtb_obj_rotate(s: *mut rtb_scene, axis: c_int, obj: c_int, angle_deg: c_double) -> c_int; // src/rotate.rs:30
    pub fn rtb_obj_translate(s: *mut rtb_scene, obj: c_int, offset: *const f64) -> c_int; // src/traslate.rs:12
    pub fn rtb_obj_constant_medium(s: *mut rtb_scene, boundary: c_int, density: c_double, tex: c_int) -> c_int; // src/medium.rs:18
    pub fn rtb_obj_list(s: *mut rtb_scene, objs: *const c_int, n: c_int) -> c_int; // src/hittable.rs:23-34
    pub fn rtb_obj_bvh(s: *mut rtb_scene, objs: *const c_int, n: c_int, time0: c_double, time1: c_double, seed: u64) -> c_int; // src/bvh.rs:17
    // scene files
    pub fn rtb_scene_save(s: *const rtb_scene, path: *const c_char) -> c_int;
    pub fn rtb_scene_load(path: *const c_char) -> *mut rtb_scene;
    pub fn rtb_scene_set_bvh_builder(s: *mut rtb_scene, builder: c_int) -> c_int;
    // rendering: create_image, tests/test.rs:55-85
    pub fn rtb_render_opts_default(o: *mut rtb_render_opts);
    pub fn rtb_render(s: *mut rtb_scene, world: c_int, cam: *const RtbCamera, nx: c_int, ny: c_int, ns: c_int, opts: *const rtb_render_opts,
                      out_sum_rgb: *mut f32, out_rgb8: *mut u8, stats: *mut rtb_stats) -> c_int;
    pub fn rtb_render_multi(s: *mut rtb_scene, world: c_int, cam: *const RtbCamera, nx: c_int, ny: c_int, ns: c_int, opts: *const rtb_render_opts,
                            devices: *const c_int, n_devices: c_int, out_sum_rgb: *mut f32, out_rgb8: *mut u8, stats: *mut rtb_stats) -> c_int;
    pub fn rtb_render_progressive(s: *mut rtb_scene, world: c_int, cam: *const RtbCamera, nx: c_int, ny: c_int, ns: c_int, opts: *const rtb_render_opts,
                                  chunk_spp: c_int, preview_every: c_int, cb: rtb_progress_fn, user: *mut c_void,
                                  out_sum_rgb: *mut f32, out_rgb8: *mut u8, samples_done: *mut c_int) -> c_int;
    // image writers: "P3\n{nx} {ny}\n255\n" + "{r} {g} {b}\n" per pixel, tests/test.rs:59,79
    pub fn rtb_ppm_p3(rgb8: *const u8, nx: c_int, ny: c_int, out: *mut c_char, cap: usize) -> usize;
    pub fn rtb_ppm_p6(rgb8: *const u8, nx: c_int, ny: c_int, out: *mut c_char, cap: usize) -> usize;
    pub fn rtb_image_write_ppm(path: *const c_char, rgb8: *const u8, nx: c_int, ny: c_int, binary: c_int) -> c_int;
    pub fn rtb_release_workspace() -> c_int;
    pub fn rtb_image_write_png(path: *const c_char, rgb8: *const u8, nx: c_int, ny: c_int) -> c_int;
    // JPEG input: image::open(path).to_rgb8(), tests/test.rs:201-205 (replaces the `image` crate dependency)
    pub fn rtb_jpeg_decode(data: *const u8, len: usize, rgb8_out: *mut *mut u8, nx: *mut u32, ny: *mut u32) -> c_int;
    pub fn rtb_jpeg_decode_file(path: *const c_char, rgb8_out: *mut *mut u8, nx: *mut u32, ny: *mut u32) -> c_int;
    pub fn rtb_image_free(rgb8: *mut u8);
    pub fn rtb_tex_image_file(s: *mut rtb_scene, path: *const c_char) -> c_int;
    // device-resident / one-process-per-GPU form
    pub fn rtb_scene_upload(s: *mut rtb_scene, world: c_int, device: c_int) -> c_int;
    pub fn rtb_render_device(s: *mut rtb_scene, cam: *const RtbCamera, nx: c_int, ny: c_int, ns: c_int, opts: *const rtb_render_opts,
                             d_sum_rgb: *mut f32, stream: *mut c_void, stats: *mut rtb_stats) -> c_int;
    pub fn rtb_device_alloc(bytes: usize, device: c_int, d_ptr: *mut *mut c_void) -> c_int;
    pub fn rtb_device_free(d_ptr: *mut c_void) -> c_int;
    pub fn rtb_peer_export(d_ptr: *const c_void, handle64: *mut u8) -> c_int;
    pub fn rtb_peer_open(handle64: *const u8, device: c_int, d_ptr: *mut *mut c_void) -> c_int;
    pub fn rtb_peer_close(d_ptr: *mut c_void) -> c_int;
    pub fn rtb_reduce_resolve_device(d_sums: *const *const f32, n_bufs: c_int, nx: c_int, ny: c_int, ns: c_int, d_total: *mut f32, d_rgb8: *mut u8, stream: *mut c_void) -> c_int;
    pub fn rtb_reduce_resolve_range_device(d_sums: *const *const f32, n_bufs: c_int, first_value: usize, n_values: usize, ns: c_int, d_total: *mut f32,
                                           d_rgb8: *mut u8, stream: *mut c_void) -> c_int;
    pub fn rtb_resolve_bands_device(d_sum_rgb: *const f32, nx: c_int, ny: c_int, ns: c_int, band_begin: c_int, band_stride: c_int, d_rgb8: *mut u8,
                                    compact: c_int, stream: *mut c_void) -> c_int;
    pub fn rtb_resolve_device(d_sum_rgb: *const f32, nx: c_int, ny: c_int, ns: c_int, d_rgb8: *mut u8, stream: *mut c_void) -> c_int;
    pub fn rtb_scene_device_bytes(s: *const rtb_scene) -> usize;
    // scene bytes (.rtscene, include/rtb_scene_format.h) and the named scene constructors of tests/test.rs:89-523
    pub fn rtb_scene_serialized_size(s: *const rtb_scene) -> usize;
    pub fn rtb_scene_serialize(s: *const rtb_scene, buf: *mut c_void, cap: usize) -> c_int;
    pub fn rtb_scene_deserialize(buf: *const c_void, len: usize) -> *mut rtb_scene;
    pub fn rtb_scene_prim_count(s: *mut rtb_scene, world: c_int) -> c_int;
    pub fn rtb_scene_named(s: *mut rtb_scene, name: *const c_char, seed: u64, earth_rgb8: *const u8, earth_nx: u32, earth_ny: u32, nx: c_int, ny: c_int,
                           camera: *mut RtbCamera) -> c_int;
    // parity hooks: Hittable::hit(ray, t_min, MAX) (src/color.rs:7) and single stages of the device path (csrc/debug.cu)
    pub fn rtb_closest_hit(s: *mut rtb_scene, world: c_int, rays: *const f64, n: usize, t_min: c_double, device: c_int, out_prim: *mut i32, out_t: *mut f32) -> c_int;
    pub fn rtb_closest_hit_ex(s: *mut rtb_scene, world: c_int, rays: *const f64, n: usize, t_min: c_double, self_id: *const i32, device: c_int,
                              out_prim: *mut i32, out_rec: *mut f32) -> c_int;
    pub fn rtb_primary_rays(cam: *const RtbCamera, nx: c_int, ny: c_int, s: c_int, seed: u64, out_rays: *mut f64) -> c_int;
    pub fn rtb_debug_camera_rays(cam: *const RtbCamera, nx: c_int, ny: c_int, s: c_int, seed: u64, device: c_int, out_rays: *mut f32) -> c_int;
    pub fn rtb_debug_camera_draws(nx: c_int, ny: c_int, s: c_int, seed: u64, out: *mut f64) -> c_int;
    pub fn rtb_debug_scatter(s: *mut rtb_scene, world: c_int, mat: c_int, input: *const f64, n: usize, seed: u64, device: c_int, out: *mut f32) -> c_int;
    pub fn rtb_debug_sample(kind: c_int, seed: u64, n: usize, device: c_int, out: *mut f32) -> c_int;
    pub fn rtb_debug_texture_value(s: *mut rtb_scene, world: c_int, tex: c_int, uvp: *const f64, n: usize, device: c_int, out: *mut f32) -> c_int;
    pub fn rtb_philox4x32_10(counter: *const u32, key: *const u32, out: *mut u32) -> c_int;
    pub fn rtb_closest_hit_noalloc(s: *mut rtb_scene, d_rays: *const f32, n: usize, t_min: c_double, d_self_id: *const i32, d_out_prim: *mut i32,
                                   d_out_rec: *mut f32, stream: *mut c_void) -> c_int;
    pub fn rtb_debug_check_failures(device: c_int, out: *mut u32) -> c_int;
    pub fn rtb_issue_peak(device: c_int, fma_tlane_per_s: *mut f64, mixed_tlane_per_s: *mut f64, slab_tlane_per_s: *mut f64) -> c_int;
}
