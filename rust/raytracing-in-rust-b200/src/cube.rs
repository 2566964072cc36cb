//! src/cube.rs of the reference: six Rects pushed in a fixed order, none flipped (:21-74); the device
//! intersects it as one box and reports the face, so a Cube still owns six primitive IDs.
use crate::hittable::Hittable;
use crate::material::Material;
use crate::{check, sys, Lower, Lowering};
use nalgebra::Vector3;

pub struct Cube {
    p_min: Vector3<f64>,
    p_max: Vector3<f64>,
    material: Box<dyn Material>,
}
impl Cube {
    pub fn new<M: Material + Clone + 'static>(p_min: Vector3<f64>, p_max: Vector3<f64>, material: M) -> Self {
        Self { p_min, p_max, material: Box::new(material) }
    }
}
impl Lower for Cube {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let m = self.material.lower(cx);
        check(unsafe { sys::rtb_obj_cube(cx.scene, self.p_min.as_ptr(), self.p_max.as_ptr(), m) }, "Cube")
    }
}
impl Hittable for Cube {}
