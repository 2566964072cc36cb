//! src/bvh.rs of the reference: BVHNode::new(&mut [Rc<dyn Hittable>], time0, time1) :17-66 builds a
//! median-split tree on the host.  Here the node only remembers its children; librtb200 builds its
//! own SAH (or device LBVH) tree over the flattened primitives — closest-hit answers do not depend
//! on the tree.  An empty slice is an error as upstream (`expect`, :43-51).
use crate::hittable::Hittable;
use crate::{check, sys, Lower, Lowering};
use std::rc::Rc;

pub struct BVHNode {
    children: Vec<Rc<dyn Hittable>>,
    time0: f64,
    time1: f64,
    seed: u64,
}
impl BVHNode {
    pub fn new(hittable: &mut [Rc<dyn Hittable>], time0: f64, time1: f64) -> Self {
        Self { children: hittable.to_vec(), time0, time1, seed: rand::random::<u64>() }
    }
}
impl Lower for BVHNode {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let hs: Vec<i32> = self.children.iter().map(|h| h.lower(cx)).collect();
        check(unsafe { sys::rtb_obj_bvh(cx.scene, hs.as_ptr(), hs.len() as i32, self.time0, self.time1, self.seed) }, "BVHNode")
    }
}
impl Hittable for BVHNode {}
