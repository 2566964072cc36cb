//! src/texture.rs of the reference: SolidTexture :9-26, CheckerTexture :28-49, NoiseTexture :51-72,
//! ImageTexture :74-108.  `value(u, v, p)` is evaluated on the device (csrc/trace.cuh texture_value).
use crate::{check, sys, Lower, Lowering};
use std::sync::Arc;

pub trait Texture: Lower {}

#[derive(Clone, Debug)]
pub struct SolidTexture {
    r: f64,
    g: f64,
    b: f64,
}
impl SolidTexture {
    pub fn new(r: f64, g: f64, b: f64) -> Self {
        Self { r, g, b }
    }
}
impl Lower for SolidTexture {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        check(unsafe { sys::rtb_tex_solid(cx.scene, self.r, self.g, self.b) }, "SolidTexture")
    }
}
impl Texture for SolidTexture {}

/// `CheckerTexture::new(odd, even)`: sin(10x)sin(10y)sin(10z) < 0 selects `odd` (src/texture.rs:40-47).
#[derive(Clone)]
pub struct CheckerTexture<T: Texture, U: Texture> {
    odd: T,
    even: U,
}
impl<T: Texture, U: Texture> CheckerTexture<T, U> {
    pub fn new(odd: T, even: U) -> Self {
        Self { odd, even }
    }
}
impl<T: Texture, U: Texture> Lower for CheckerTexture<T, U> {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let o = self.odd.lower(cx);
        let e = self.even.lower(cx);
        check(unsafe { sys::rtb_tex_checker(cx.scene, o, e) }, "CheckerTexture")
    }
}
impl<T: Texture, U: Texture> Texture for CheckerTexture<T, U> {}

/// Perlin tables are drawn at construction upstream (Perlin::new, src/perlin.rs:67-74); here the
/// construction draws a table seed and librtb200 builds the same kind of tables from it.
#[derive(Clone, Debug)]
pub struct NoiseTexture {
    scale: f64,
    table_seed: u64,
}
impl NoiseTexture {
    pub fn new(scale: f64) -> Self {
        Self { scale, table_seed: rand::random::<u64>() }
    }
    /// reproducible tables (the reference is unseeded)
    pub fn with_seed(scale: f64, table_seed: u64) -> Self {
        Self { scale, table_seed }
    }
}
impl Lower for NoiseTexture {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        let key = (self.table_seed, self.scale.to_bits());
        if let Some(h) = cx.noise.get(&key) {
            return *h; // a clone of the same texture: same tables, same handle
        }
        let h = check(unsafe { sys::rtb_tex_noise(cx.scene, self.scale, self.table_seed) }, "NoiseTexture");
        cx.noise.insert(key, h);
        h
    }
}
impl Texture for NoiseTexture {}

/// `ImageTexture::new(data, nx, ny)`: RGB8 rows, nearest texel (src/texture.rs:94-107).
#[derive(Clone)]
pub struct ImageTexture {
    data: Arc<Vec<u8>>,
    nx: u32,
    ny: u32,
}
impl ImageTexture {
    pub fn new(data: Vec<u8>, nx: u32, ny: u32) -> Self {
        Self { data: Arc::new(data), nx, ny }
    }
}
impl Lower for ImageTexture {
    fn lower(&self, cx: &mut Lowering) -> i32 {
        assert!(self.data.len() >= 3 * self.nx as usize * self.ny as usize, "ImageTexture: buffer shorter than nx*ny*3");
        check(unsafe { sys::rtb_tex_image(cx.scene, self.data.as_ptr(), self.nx, self.ny) }, "ImageTexture")
    }
}
impl Texture for ImageTexture {}
