//! src/rotate.rs of the reference: Axis :7-11, Rotate :21-118 (angle in degrees).
use crate::hittable::Hittable;
use crate::{check, sys, Lower, Lowering};

#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Axis {
    X,
    Y,
    Z,
}

pub struct Rotate<H: Hittable> {
    axis: Axis,
    hittable: H,
    angle: f64,
}
impl<H: Hittable> Rotate<H> {
    pub fn new(axis: Axis, hittable: H, angle: f64) -> Self {
        Self { axis, hittable, angle }
    }
}
impl<H: Hittable> Lower for Rotate<H> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let h = self.hittable.lower(cx);
        let axis = match self.axis {
            Axis::X => sys::RTB_AXIS_X,
            Axis::Y => sys::RTB_AXIS_Y,
            Axis::Z => sys::RTB_AXIS_Z,
        };
        check(unsafe { sys::rtb_obj_rotate(cx.scene, axis, h, self.angle) }, "Rotate")
    }
}
impl<H: Hittable> Hittable for Rotate<H> {}
