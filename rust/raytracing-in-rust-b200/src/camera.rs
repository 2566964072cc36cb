//! src/camera.rs of the reference: Camera::new :21-51 keeps its nine arguments; the basis vectors
//! and get_ray (:53-67) are evaluated by librtb200 (f64 on the host for the basis, on the device per ray).
use crate::sys::RtbCamera;
use nalgebra::Vector3;

#[derive(Clone, Copy, Debug)]
pub struct Camera {
    pub(crate) c: RtbCamera,
}
impl Camera {
    #[allow(clippy::too_many_arguments)]
    pub fn new(look_from: Vector3<f64>, look_at: Vector3<f64>, view_up: Vector3<f64>, vertical_fov: f64, aspect: f64, aperture: f64,
               focus_dist: f64, time0: f64, time1: f64) -> Self {
        let v = |a: Vector3<f64>| [a.x, a.y, a.z];
        Self {
            c: RtbCamera {
                look_from: v(look_from),
                look_at: v(look_at),
                view_up: v(view_up),
                vertical_fov,
                aspect,
                aperture,
                focus_dist,
                time0,
                time1,
            },
        }
    }
}
