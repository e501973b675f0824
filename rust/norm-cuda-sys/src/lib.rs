//! FFI to `libnorm_cuda.so` (C ABI: `include/norm_cuda.h`) and a safe wrapper with the shape of
//! norm's own effect API: `FeatureUniform` in, decoded image in, RGBA8 buffer out.
//! UNVERIFIED: written without a Rust toolchain; mirrors the header one-to-one.
#![allow(non_camel_case_types)]
use std::ffi::{c_char, c_void, CStr};

pub type norm_status = i32;
pub const NORM_OK: norm_status = 0;
pub const NORM_CS_REFERENCE: u32 = 0;
pub const NORM_CS_RAW: u32 = 1;
pub const NORM_EXPAND_COMPAT: i32 = 0;
pub const NORM_EXPAND_SPEC: i32 = 1;

#[repr(C)]
pub struct norm_ctx { _p: [u8; 0] }
#[repr(C)]
pub struct norm_img { _p: [u8; 0] }

/// `FeatureUniform` of norm (src/renderer/feature_uniform.rs:1-17), 112 bytes.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct NormFeatureUniform {
    pub grayscale: u32,
    pub invert: u32,
    pub gamma: u32,
    pub blur: u32,
    pub blur_radius: u32,
    pub width: u32,
    pub height: u32,
    pub sharpen: u32,
    pub sharpen_factor: u32,
    pub edge_detect: u32,
    pub _padding_1: [u8; 8],
    pub transform: [[f32; 4]; 4],
}
const _: () = assert!(std::mem::size_of::<NormFeatureUniform>() == 112);

#[repr(C)]
pub struct NormPngDesc {
    pub width: u32,
    pub height: u32,
    pub bit_depth: u8,
    pub color_type: u8,
    pub _reserved: [u8; 2],
    pub gamma: u32,
    pub pixels: *const u8,
    pub len: usize,
    pub plte: *const u8,
    pub n_plte: i32,
    pub trns: *const u8,
    pub n_trns: i32,
}

/// The inflated IDAT stream as `ScanlineReader::new` receives it (src/png/scanline_reader.rs:15-28).
#[repr(C)]
pub struct NormInflatedDesc {
    pub width: u32,
    pub height: u32,
    pub bit_depth: u8,
    pub color_type: u8,
    pub interlace: u8,
    pub _reserved: u8,
    pub data: *const u8,
    pub len: usize,
}

/// src/renderer/shape_uniform.rs:3-10
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct NormShapeUniform {
    pub width: u32,
    pub height: u32,
    pub num_circles: u32,
    pub selected_circle: u32,
}

/// src/renderer/shape_uniform.rs:38-45
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct NormCircleData {
    pub x: f32,
    pub y: f32,
    pub radius: f32,
    pub _padding: f32,
}

/// src/renderer/draw_uniform.rs:3-9
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct NormDrawUniform {
    pub crosshair: u32,
    pub circle_center_x: f32,
    pub circle_center_y: f32,
    pub circle_radius: f32,
    pub camera_view_proj: [[f32; 4]; 4],
}
const _: () = assert!(std::mem::size_of::<NormDrawUniform>() == 80);

extern "C" {
    pub fn norm_cuda_abi_version() -> i32;
    pub fn norm_cuda_create(device: i32, out: *mut *mut norm_ctx) -> norm_status;
    pub fn norm_cuda_destroy(ctx: *mut norm_ctx);
    pub fn norm_cuda_last_error(ctx: *const norm_ctx) -> *const c_char;
    pub fn norm_cuda_sync(ctx: *mut norm_ctx) -> norm_status;
    pub fn norm_cuda_image_create(ctx: *mut norm_ctx, w: u32, h: u32, frames: u32, out: *mut *mut norm_img) -> norm_status;
    pub fn norm_cuda_image_destroy(ctx: *mut norm_ctx, img: *mut norm_img);
    pub fn norm_cuda_upload(ctx: *mut norm_ctx, img: *mut norm_img, frame0: u32, n: u32, host: *const u8, pitch: usize) -> norm_status;
    pub fn norm_cuda_download(ctx: *mut norm_ctx, img: *const norm_img, frame0: u32, n: u32, host: *mut u8, pitch: usize) -> norm_status;
    pub fn norm_cuda_expand_rgba8(ctx: *mut norm_ctx, png: *const NormPngDesc, mode: i32, out: *mut norm_img, frame: u32) -> norm_status;
    pub fn norm_cuda_effects(ctx: *mut norm_ctx, f: *const NormFeatureUniform, flags: u32, input: *const norm_img, out: *mut norm_img) -> norm_status;
    pub fn norm_cuda_chain(ctx: *mut norm_ctx, f: *const NormFeatureUniform, flags: u32, input: *const norm_img, out: *mut norm_img, ssim: *mut f64) -> norm_status;
    pub fn norm_cuda_chain_host(ctx: *mut norm_ctx, f: *const NormFeatureUniform, flags: u32, input: *const u8, out: *mut u8, frames: u32, ssim: *mut f64) -> norm_status;
    pub fn norm_cuda_chain_band(ctx: *mut norm_ctx, f: *const NormFeatureUniform, flags: u32, input: *const norm_img, out: *mut norm_img, own_row0: u32, own_rows: u32, moments: *mut f64) -> norm_status;
    pub fn norm_cuda_ssim(ctx: *mut norm_ctx, a: *const norm_img, b: *const norm_img, ssim: *mut f64) -> norm_status;
    pub fn norm_cuda_ssim_moments(ctx: *mut norm_ctx, a: *const norm_img, b: *const norm_img, moments: *mut f64) -> norm_status;
    pub fn norm_ssim_from_moments(moments: *const f64, npix: f64) -> f64;
    pub fn norm_cuda_unfilter(ctx: *mut norm_ctx, inflated: *const NormInflatedDesc, pixel_buffer: *mut u8) -> norm_status;
    pub fn norm_cuda_unfilter_expand_rgba8(ctx: *mut norm_ctx, inflated: *const NormInflatedDesc, plte: *const u8, n_plte: i32, trns: *const u8, n_trns: i32, mode: i32, out: *mut norm_img, frame: u32) -> norm_status;
    pub fn norm_cuda_unfilter_expand_rgba8_batch(ctx: *mut norm_ctx, inflated: *const NormInflatedDesc, n: u32, plte: *const u8, n_plte: i32, trns: *const u8, n_trns: i32, mode: i32, out: *mut norm_img, frame0: u32) -> norm_status;
    pub fn norm_cuda_filter_scanlines(ctx: *mut norm_ctx, width: u32, height: u32, bit_depth: u8, color_type: u8, pixel_buffer: *const u8, paeth_compat: i32, stream: *mut u8, filter_types: *mut u8) -> norm_status;
    pub fn norm_cuda_filter_image(ctx: *mut norm_ctx, img: *const norm_img, frame: u32, paeth_compat: i32, stream: *mut u8, filter_types: *mut u8) -> norm_status;
    pub fn norm_cuda_box_blur(ctx: *mut norm_ctx, input: *const norm_img, out: *mut norm_img) -> norm_status;
    pub fn norm_cuda_overlay(ctx: *mut norm_ctx, shapes: *const NormShapeUniform, circles: *const NormCircleData, draw: *const NormDrawUniform, input: *const norm_img, out: *mut norm_img) -> norm_status;
    pub fn norm_cuda_ssim_static(a: *const c_void, b: *const c_void, w: u32, h: u32, frames: u32, pitch: usize, frame_stride: usize, ssim_dev: *mut f64, moments_dev: *mut f64, stream: *mut c_void) -> norm_status;
}

fn check(ctx: *const norm_ctx, status: norm_status) -> anyhow::Result<()> {
    if status == NORM_OK {
        return Ok(());
    }
    let msg = unsafe { CStr::from_ptr(norm_cuda_last_error(ctx)) }.to_string_lossy().into_owned();
    anyhow::bail!("norm_cuda status {status}: {msg}")
}

/// Decoded image as norm's `Png` holds it (src/png/grammar.rs:65-70).
pub struct DecodedImage<'a> {
    pub width: u32,
    pub height: u32,
    pub bit_depth: u8,
    pub color_type: u8,
    pub gamma: u32,
    pub pixel_buffer: &'a [u8],
    pub plte: Option<&'a [u8]>,
    pub trns: Option<&'a [u8]>,
}

/// Stands where `Texture::from_image` + `EffectPipeline` + the fragment effects do
/// (texture.rs:22-80, effect_pipeline.rs:39-51, image_shader.wgsl:158-174).
pub struct CudaEffectPipeline {
    ctx: *mut norm_ctx,
    input: *mut norm_img,
    output: *mut norm_img,
    width: u32,
    height: u32,
}

impl CudaEffectPipeline {
    pub fn new(device: i32, image: &DecodedImage<'_>, mode: i32) -> anyhow::Result<Self> {
        let mut ctx = std::ptr::null_mut();
        check(std::ptr::null(), unsafe { norm_cuda_create(device, &mut ctx) })?;
        let mut this = Self { ctx, input: std::ptr::null_mut(), output: std::ptr::null_mut(), width: image.width, height: image.height };
        check(ctx, unsafe { norm_cuda_image_create(ctx, image.width, image.height, 1, &mut this.input) })?;
        check(ctx, unsafe { norm_cuda_image_create(ctx, image.width, image.height, 1, &mut this.output) })?;
        let desc = NormPngDesc {
            width: image.width,
            height: image.height,
            bit_depth: image.bit_depth,
            color_type: image.color_type,
            _reserved: [0; 2],
            gamma: image.gamma,
            pixels: image.pixel_buffer.as_ptr(),
            len: image.pixel_buffer.len(),
            plte: image.plte.map_or(std::ptr::null(), |p| p.as_ptr()),
            n_plte: image.plte.map_or(0, |p| (p.len() / 3) as i32),
            trns: image.trns.map_or(std::ptr::null(), |t| t.as_ptr()),
            n_trns: image.trns.map_or(0, |t| t.len() as i32),
        };
        check(ctx, unsafe { norm_cuda_expand_rgba8(ctx, &desc, mode, this.input, 0) })?;
        Ok(this)
    }

    /// One frame of the renderer, headless: RGBA8 out.
    pub fn execute(&self, f: &NormFeatureUniform, flags: u32) -> anyhow::Result<Vec<u8>> {
        check(self.ctx, unsafe { norm_cuda_effects(self.ctx, f, flags, self.input, self.output) })?;
        self.read_back()
    }

    /// The chained path (front -> blur -> sharpen -> edge) and SSIM of the result against the input.
    pub fn chain(&self, f: &NormFeatureUniform, flags: u32) -> anyhow::Result<(Vec<u8>, f64)> {
        let mut ssim = 0f64;
        check(self.ctx, unsafe { norm_cuda_chain(self.ctx, f, flags, self.input, self.output, &mut ssim) })?;
        Ok((self.read_back()?, ssim))
    }

    fn read_back(&self) -> anyhow::Result<Vec<u8>> {
        let mut out = vec![0u8; 4 * self.width as usize * self.height as usize];
        check(self.ctx, unsafe { norm_cuda_download(self.ctx, self.output, 0, 1, out.as_mut_ptr(), 0) })?;
        Ok(out)
    }
}

impl Drop for CudaEffectPipeline {
    fn drop(&mut self) {
        unsafe {
            norm_cuda_image_destroy(self.ctx, self.input);
            norm_cuda_image_destroy(self.ctx, self.output);
            norm_cuda_destroy(self.ctx);
        }
    }
}
