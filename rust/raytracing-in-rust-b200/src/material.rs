//! src/material.rs of the reference: Lambertian :36-58, Metal :61-92, Dielectric :95-131,
//! DiffuseLight :133-151, Isotropic :154-173.  scatter/emitted run on the device (csrc/trace.cuh).
use crate::texture::Texture;
use crate::{check, sys, Lower, Lowering};

pub trait Material: Lower {}

#[derive(Clone)]
pub struct Lambertian<T: Texture> {
    albedo: T,
}
impl<T: Texture> Lambertian<T> {
    pub fn new(albedo: T) -> Self {
        Self { albedo }
    }
}
impl<T: Texture> Lower for Lambertian<T> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let t = self.albedo.lower(cx);
        check(unsafe { sys::rtb_mat_lambertian(cx.scene, t) }, "Lambertian")
    }
}
impl<T: Texture> Material for Lambertian<T> {}

#[derive(Clone)]
pub struct Metal<T: Texture> {
    albedo: T,
    fuzz: f64,
}
impl<T: Texture> Metal<T> {
    /// fuzz is clamped to <= 1 where it is used, as upstream (src/material.rs:70)
    pub fn new(albedo: T, fuzz: f64) -> Self {
        Self { albedo, fuzz }
    }
}
impl<T: Texture> Lower for Metal<T> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let t = self.albedo.lower(cx);
        check(unsafe { sys::rtb_mat_metal(cx.scene, t, self.fuzz) }, "Metal")
    }
}
impl<T: Texture> Material for Metal<T> {}

#[derive(Clone, Debug)]
pub struct Dielectric {
    ref_idx: f64,
}
impl Dielectric {
    pub fn new(ref_idx: f64) -> Self {
        Self { ref_idx }
    }
}
impl Lower for Dielectric {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        check(unsafe { sys::rtb_mat_dielectric(cx.scene, self.ref_idx) }, "Dielectric")
    }
}
impl Material for Dielectric {}

#[derive(Clone)]
pub struct DiffuseLight<T: Texture> {
    emit: T,
}
impl<T: Texture> DiffuseLight<T> {
    pub fn new(emit: T) -> Self {
        Self { emit }
    }
}
impl<T: Texture> Lower for DiffuseLight<T> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let t = self.emit.lower(cx);
        check(unsafe { sys::rtb_mat_diffuse_light(cx.scene, t) }, "DiffuseLight")
    }
}
impl<T: Texture> Material for DiffuseLight<T> {}

#[derive(Clone)]
pub struct Isotropic<T: Texture> {
    albedo: T,
}
impl<T: Texture> Isotropic<T> {
    pub fn new(albedo: T) -> Self {
        Self { albedo }
    }
}
impl<T: Texture> Lower for Isotropic<T> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let t = self.albedo.lower(cx);
        check(unsafe { sys::rtb_mat_isotropic(cx.scene, t) }, "Isotropic")
    }
}
impl<T: Texture> Material for Isotropic<T> {}
