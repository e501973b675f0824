//! FFI to `libnorm_cuda.so` (C ABI: `include/norm_cuda.h`) and a safe wrapper with the shape of
//! norm's own effect API: `FeatureUniform` in, decoded image in, RGBA8 buffer out.
//! UNVERIFIED: written without a Rust toolchain.  The `extern "C"` blocks (`ffi.rs`) are GENERATED from
//! `include/norm_cuda.h` and `include/norm_nccl.h` by `tools/gen_rust_ffi.py`: one declaration per exported symbol, and
//! `tests/test_host.py` checks that list against `nm -D` of the built libraries.  The `#[repr(C)]` structs below are
//! written by hand; their sizes are asserted at compile time against the sizes the C side asserts.
#![allow(non_camel_case_types)]
use std::ffi::{c_char, c_void, CStr};

pub type norm_status = i32;
pub const NORM_OK: norm_status = 0;
pub const NORM_CS_REFERENCE: u32 = 0;
pub const NORM_CS_RAW: u32 = 1;
pub const NORM_OUT_SRGB: u32 = 2;
pub const NORM_LUMA_RGB: i32 = 0;
pub const NORM_LUMA_GRAY: i32 = 1;
pub const NORM_LUMA_GRAY_UNIT: i32 = 2;
pub const NORM_EXPAND_COMPAT: i32 = 0;
pub const NORM_EXPAND_SPEC: i32 = 1;

#[repr(C)]
pub struct norm_ctx { _p: [u8; 0] }
#[repr(C)]
pub struct norm_img { _p: [u8; 0] }
#[repr(C)]
pub struct norm_band { _p: [u8; 0] }
#[repr(C)]
pub struct norm_nccl_comm { _p: [u8; 0] }

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

const _: () = assert!(std::mem::size_of::<NormPngDesc>() == 64);

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

const _: () = assert!(std::mem::size_of::<NormInflatedDesc>() == 32);

/// What one rank of a row-band sharded image tells the others (`norm_cuda_band_export`), 256 bytes.
#[repr(C)]
#[derive(Clone, Copy)]
pub struct NormBandHandle {
    pub mem: [u8; 64],
    pub ready_event: [u8; 64],
    pub halo_event: [u8; 64],
    pub pitch: u64,
    pub local_ptr: u64,
    pub local_ready: u64,
    pub local_halo: u64,
    pub pid: u64,
    pub row0: u32,
    pub rows: u32,
    pub halo_up: u32,
    pub halo_down: u32,
    pub rank: i32,
    pub device: i32,
}
const _: () = assert!(std::mem::size_of::<NormBandHandle>() == 256);

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

const _: () = assert!(std::mem::size_of::<NormShapeUniform>() == 16 && std::mem::size_of::<NormCircleData>() == 16);

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

mod ffi;
pub use ffi::*;

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
