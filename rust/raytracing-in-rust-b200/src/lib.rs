//! The raytracing_rust scene API as *descriptions*: same module paths and constructor signatures as
//! the reference crate (src/lib.rs:1-18), nothing intersected on the host.  `create_image` lowers the
//! object graph into an `rtb_scene` — one C call per reference constructor call — and renders it with
//! librtb200 on the GPU(s).  Not compiled in this repository's image (no Rust toolchain); the same
//! lowering is what include/rtb200.hpp does in C++, which is compiled and tested.
pub use rtb200_sys as sys;

use std::collections::HashMap;
use std::ffi::CStr;
use std::os::raw::c_char;

pub mod bvh;
pub mod camera;
pub mod cube;
pub mod hittable;
pub mod image;
pub mod logger;
pub mod material;
pub mod medium;
pub mod rect;
pub mod rotate;
pub mod sphere;
pub mod texture;
pub mod traslate;

/// `color` runs on the device (src/color.rs:6-23 is what the kernels implement); the symbol exists so the
/// driver's `use crate::color::color;` line still resolves.
pub mod color {
    pub fn color<R, W>(_ray: &R, _world: &W, _depth: usize) -> nalgebra::Vector3<f64> {
        unimplemented!("color() is evaluated by librtb200 on the GPU")
    }
}

/// State of one lowering pass: the scene being filled and the handles of shared Perlin tables
/// (cloning a NoiseTexture upstream copies its tables, src/perlin.rs:58; here clones share a handle).
pub struct Lowering {
    pub scene: *mut sys::rtb_scene,
    pub(crate) noise: HashMap<(u64, u64), i32>,
}

/// Records `self` into the scene and returns its handle.  Textures, materials and hittables all lower.
pub trait Lower {
    fn lower(&self, cx: &mut Lowering) -> i32;
}

pub(crate) fn last_error() -> String {
    unsafe { CStr::from_ptr(sys::rtb_last_error()).to_string_lossy().into_owned() }
}

/// Upstream errors are panics (`expect`, src/bvh.rs:43-51); negative handles become panics here too.
pub(crate) fn check(h: i32, what: &str) -> i32 {
    if h < 0 {
        panic!("{}: {}", what, last_error());
    }
    h
}

/// create_image(ny, nx, ns, cam, world) -> String, tests/test.rs:55-85: same argument order (ny first),
/// same P3 text.  `RTB200_GPUS` = 1 (default) renders on device 0, 0 = every GPU, k = the first k.
pub fn create_image(ny: usize, nx: usize, ns: usize, cam: camera::Camera, world: Box<dyn hittable::Hittable>) -> String {
    let gpus: i32 = std::env::var("RTB200_GPUS").ok().and_then(|v| v.parse().ok()).unwrap_or(1);
    create_image_on(ny, nx, ns, cam, world, gpus, None)
}

/// The same with an explicit GPU count and render options (background, seed, ...).
pub fn create_image_on(ny: usize, nx: usize, ns: usize, cam: camera::Camera, world: Box<dyn hittable::Hittable>, n_gpus: i32,
                       opts: Option<sys::rtb_render_opts>) -> String {
    unsafe {
        let scene = sys::rtb_scene_create();
        let mut cx = Lowering { scene, noise: HashMap::new() };
        let w = world.lower(&mut cx);
        let o = opts.unwrap_or_else(|| {
            let mut o = std::mem::MaybeUninit::<sys::rtb_render_opts>::uninit();
            sys::rtb_render_opts_default(o.as_mut_ptr());
            o.assume_init()
        });
        let mut rgb8 = vec![0u8; 3 * nx * ny];
        let rc = if n_gpus == 1 {
            sys::rtb_render(scene, w, &cam.c, nx as i32, ny as i32, ns as i32, &o, std::ptr::null_mut(), rgb8.as_mut_ptr(), std::ptr::null_mut())
        } else {
            sys::rtb_render_multi(scene, w, &cam.c, nx as i32, ny as i32, ns as i32, &o, std::ptr::null(), n_gpus, std::ptr::null_mut(),
                                  rgb8.as_mut_ptr(), std::ptr::null_mut())
        };
        if rc != 0 {
            let msg = last_error();
            sys::rtb_scene_destroy(scene);
            panic!("create_image: {}", msg); // no CPU fallback: a missing GPU is an error, not a slow path
        }
        let n = sys::rtb_ppm_p3(rgb8.as_ptr(), nx as i32, ny as i32, std::ptr::null_mut(), 0);
        let mut out = vec![0u8; n + 1];
        sys::rtb_ppm_p3(rgb8.as_ptr(), nx as i32, ny as i32, out.as_mut_ptr() as *mut c_char, n + 1);
        sys::rtb_scene_destroy(scene);
        out.truncate(n);
        String::from_utf8(out).expect("P3 text is ASCII")
    }
}
