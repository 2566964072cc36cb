//! src/traslate.rs of the reference (sic): Traslate :6-33.
use crate::hittable::Hittable;
use crate::{check, sys, Lower, Lowering};
use nalgebra::Vector3;

pub struct Traslate<H: Hittable> {
    hitable: H,
    offset: Vector3<f64>,
}
impl<H: Hittable> Traslate<H> {
    pub fn new(hitable: H, offset: Vector3<f64>) -> Self {
        Self { hitable, offset }
    }
}
impl<H: Hittable> Lower for Traslate<H> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let h = self.hitable.lower(cx);
        check(unsafe { sys::rtb_obj_translate(cx.scene, h, self.offset.as_ptr()) }, "Traslate")
    }
}
impl<H: Hittable> Hittable for Traslate<H> {}
