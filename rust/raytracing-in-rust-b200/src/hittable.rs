//! src/hittable.rs of the reference: Hittable :18-21, HittableList :23-65, FlipNormals :67-88.
use crate::{check, sys, Lower, Lowering};
use std::rc::Rc;

/// Upstream: `hit(ray, t_min, t_max)` and `bounding_box(t0, t1)`.  Both run on the device; a
/// hittable here is something that can record itself into the scene.
pub trait Hittable: Lower {}

impl<H: Hittable + ?Sized> Lower for Box<H> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        (**self).lower(cx)
    }
}
impl<H: Hittable + ?Sized> Hittable for Box<H> {}
impl<H: Hittable + ?Sized> Lower for Rc<H> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        (**self).lower(cx)
    }
}
impl<H: Hittable + ?Sized> Hittable for Rc<H> {}

/// Linear scan in insertion order upstream (:37-47); the order is kept because ties between
/// coincident planes are decided by it (a later Rect replaces an earlier one at equal t).
#[derive(Default)]
pub struct HittableList {
    list: Vec<Box<dyn Hittable>>,
}
impl HittableList {
    pub fn push(&mut self, hittable: impl Hittable + 'static) {
        self.list.push(Box::new(hittable))
    }
}
impl Lower for HittableList {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let hs: Vec<i32> = self.list.iter().map(|h| h.lower(cx)).collect();
        check(unsafe { sys::rtb_obj_list(cx.scene, hs.as_ptr(), hs.len() as i32) }, "HittableList")
    }
}
impl Hittable for HittableList {}

pub struct FlipNormals<H: Hittable> {
    hittable: H,
}
impl<H: Hittable> FlipNormals<H> {
    pub fn new(hittable: H) -> Self {
        Self { hittable }
    }
}
impl<H: Hittable> Lower for FlipNormals<H> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let h = self.hittable.lower(cx);
        check(unsafe { sys::rtb_obj_flip_normals(cx.scene, h) }, "FlipNormals")
    }
}
impl<H: Hittable> Hittable for FlipNormals<H> {}
