//! src/medium.rs of the reference: ConstantMedium :11-61 — a boundary hittable, a density and the
//! albedo texture of the Isotropic phase function it owns (:20-24).
use crate::hittable::Hittable;
use crate::texture::Texture;
use crate::{check, sys, Lower, Lowering};

pub struct ConstantMedium<H: Hittable, T: Texture> {
    boundary: H,
    density: f64,
    texture: T,
}
impl<H: Hittable, T: Texture> ConstantMedium<H, T> {
    pub fn new(boundary: H, density: f64, texture: T) -> Self {
        Self { boundary, density, texture }
    }
}
impl<H: Hittable, T: Texture> Lower for ConstantMedium<H, T> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let b = self.boundary.lower(cx);
        let t = self.texture.lower(cx);
        check(unsafe { sys::rtb_obj_constant_medium(cx.scene, b, self.density, t) }, "ConstantMedium")
    }
}
impl<H: Hittable, T: Texture> Hittable for ConstantMedium<H, T> {}
