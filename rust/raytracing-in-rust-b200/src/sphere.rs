//! src/sphere.rs of the reference: Sphere :20-85, MovingSphere :87-175.
use crate::hittable::Hittable;
use crate::material::Material;
use crate::{check, sys, Lower, Lowering};
use nalgebra::Vector3;

pub struct Sphere<M: Material> {
    center: Vector3<f64>,
    radius: f64,
    material: M,
}
impl<M: Material> Sphere<M> {
    pub fn new(center: Vector3<f64>, radius: f64, material: M) -> Self {
        Self { center, radius, material }
    }
}
impl<M: Material> Lower for Sphere<M> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let m = self.material.lower(cx);
        check(unsafe { sys::rtb_obj_sphere(cx.scene, self.center.as_ptr(), self.radius, m) }, "Sphere")
    }
}
impl<M: Material> Hittable for Sphere<M> {}

/// center(time) = c0 + ((time - t0) / (t1 - t0)) (c1 - c0), src/sphere.rs:115-118
pub struct MovingSphere<M: Material> {
    center0: Vector3<f64>,
    center1: Vector3<f64>,
    time0: f64,
    time1: f64,
    radius: f64,
    material: M,
}
impl<M: Material> MovingSphere<M> {
    pub fn new(center0: Vector3<f64>, center1: Vector3<f64>, time0: f64, time1: f64, radius: f64, material: M) -> Self {
        Self { center0, center1, time0, time1, radius, material }
    }
}
impl<M: Material> Lower for MovingSphere<M> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let m = self.material.lower(cx);
        check(
            unsafe { sys::rtb_obj_moving_sphere(cx.scene, self.center0.as_ptr(), self.center1.as_ptr(), self.time0, self.time1, self.radius, m) },
            "MovingSphere",
        )
    }
}
impl<M: Material> Hittable for MovingSphere<M> {}
