//! src/rect.rs of the reference: Plane :8-12, Rect :14-76.
use crate::hittable::Hittable;
use crate::material::Material;
use crate::{check, sys, Lower, Lowering};

#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Plane {
    YZ,
    ZX,
    XY,
}

pub struct Rect<M: Material> {
    plane: Plane,
    x0: f64,
    y0: f64,
    x1: f64,
    y1: f64,
    k: f64,
    material: M,
}
impl<M: Material> Rect<M> {
    pub fn new(plane: Plane, x0: f64, y0: f64, x1: f64, y1: f64, k: f64, material: M) -> Self {
        Self { plane, x0, y0, x1, y1, k, material }
    }
}
impl<M: Material> Lower for Rect<M> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let m = self.material.lower(cx);
        let plane = match self.plane {
            Plane::YZ => sys::RTB_PLANE_YZ,
            Plane::ZX => sys::RTB_PLANE_ZX,
            Plane::XY => sys::RTB_PLANE_XY,
        };
        check(unsafe { sys::rtb_obj_rect(cx.scene, plane, self.x0, self.y0, self.x1, self.y1, self.k, m) }, "Rect")
    }
}
impl<M: Material> Hittable for Rect<M> {}
