//! Stand-in for the three calls the reference makes into the `image` crate
//! (tests/test.rs:201-205, 489-493): `image::open(path)?.to_rgb8()`, `.dimensions()`, `.into_raw()`.
//! Decoding is done by librtb200's own baseline JPEG decoder (csrc/image_io.cpp, byte-identical to
//! libjpeg-turbo's default decode), so the driver can drop its `image` dependency with
//! `use raytracing_in_rust::image;` in place of `use image;`.
use crate::sys;
use std::ffi::CString;
use std::path::Path;

#[derive(Debug)]
pub struct ImageError(pub String);
impl std::fmt::Display for ImageError {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        f.write_str(&self.0)
    }
}
impl std::error::Error for ImageError {}

/// What `image::open` returns; only the RGB8 view is offered.
pub struct DynamicImage(RgbImage);
impl DynamicImage {
    pub fn to_rgb8(&self) -> RgbImage {
        self.0.clone()
    }
    pub fn into_rgb8(self) -> RgbImage {
        self.0
    }
}

#[derive(Clone)]
pub struct RgbImage {
    data: Vec<u8>,
    nx: u32,
    ny: u32,
}
impl RgbImage {
    pub fn dimensions(&self) -> (u32, u32) {
        (self.nx, self.ny)
    }
    pub fn into_raw(self) -> Vec<u8> {
        self.data
    }
}

pub fn open<P: AsRef<Path>>(path: P) -> Result<DynamicImage, ImageError> {
    let c = CString::new(path.as_ref().to_string_lossy().as_bytes()).map_err(|e| ImageError(e.to_string()))?;
    let (mut p, mut nx, mut ny) = (std::ptr::null_mut::<u8>(), 0u32, 0u32);
    let rc = unsafe { sys::rtb_jpeg_decode_file(c.as_ptr(), &mut p, &mut nx, &mut ny) };
    if rc != 0 {
        return Err(ImageError(crate::last_error()));
    }
    let data = unsafe { std::slice::from_raw_parts(p, 3 * nx as usize * ny as usize) }.to_vec();
    unsafe { sys::rtb_image_free(p) };
    Ok(DynamicImage(RgbImage { data, nx, ny }))
}
