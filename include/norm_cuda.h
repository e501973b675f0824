/*
 * norm_cuda.h -- C ABI of libnorm_cuda.so: a B200 (sm_100a) CUDA backend for the
 * data-parallel image-effect path of friendlymatthew/norm.
 *
 * The reference crate exports no FFI for this path (it is a cdylib + rlib with no
 * #[no_mangle] symbols, Cargo.toml:42-43); the seam this library replaces is the Rust-level
 * one between `renderer::app_state` and wgpu.  Each entry point below cites the reference
 * interface it stands in for (paths relative to the reference repository root).
 * INTEGRATION.md shows the `extern "C"` block + safe wrapper a norm maintainer would add as
 * `src/renderer/cuda_backend.rs`.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ / torch types cross this boundary;
 *   - every function returns a norm_status; nothing throws or aborts across the ABI
 *     (the reference's release profile is panic = "abort", Cargo.toml:47);
 *   - the caller owns every host buffer; the library owns device and pinned memory behind
 *     opaque handles with explicit *_destroy / *_free;
 *   - a context is bound to ONE device and is thread-compatible (one thread at a time);
 *     all work of a context is ordered on its stream; calls are synchronous unless named
 *     *_async (finish those with norm_cuda_sync);
 *   - pixels are RGBA8, row-major, top row first; device images are pitched (pitch % 16 == 0).
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     NORM_E_CUDA.
 */
#ifndef NORM_CUDA_H
#define NORM_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NORM_CUDA_ABI_VERSION 1

typedef int32_t norm_status;
enum {
    NORM_OK = 0,
    NORM_E_INVALID_ARG = 1,  /* `ensure!` / `bail!` class of errors (decoder.rs:30-33, ssim.rs:92-95) */
    NORM_E_UNSUPPORTED = 2,  /* where the reference hits unimplemented!/todo! (grammar.rs:119,153,195) */
    NORM_E_CUDA = 3,
    NORM_E_NCCL = 4,
    NORM_E_OOM = 5,
    NORM_E_DIM_MISMATCH = 6  /* ssim.rs:92-95 "identical dimensions" */
};

/* Effect parameters.  Field-for-field copy of `FeatureUniform`
 * (src/renderer/feature_uniform.rs:1-17; WGSL mirror image_shader.wgsl:2-15): 112 bytes. */
typedef struct NormFeatureUniform {
    uint32_t grayscale;      /* 1 = on (feature_uniform.rs:40-42)                       */
    uint32_t invert;
    uint32_t gamma;          /* raw gAMA * 1e5, 0 = pass-through (app_state.rs:65)      */
    uint32_t blur;
    uint32_t blur_radius;    /* UI range 3..39 step 2, default 21 (feature_uniform.rs:66-84); accepted 1..63 */
    uint32_t width;          /* must equal the image width  (app_state.rs:65,678-682)   */
    uint32_t height;         /* must equal the image height                             */
    uint32_t sharpen;
    uint32_t sharpen_factor; /* UI range 1..40, default 16 (feature_uniform.rs:88-106)  */
    uint32_t edge_detect;
    uint8_t _padding_1[8];
    float transform[4][4];   /* FlipX / FlipY matrix (feature_uniform.rs:143-175)       */
} NormFeatureUniform;

#define NORM_MAX_BLUR_RADIUS 63u

/* flags for the effect entry points */
enum {
    NORM_CS_REFERENCE = 0u, /* input is sRGB-encoded: decoded at first load as Rgba8UnormSrgb does
                               (texture.rs:36); intermediates are linear unorm8 (gpu_state.rs:608) */
    NORM_CS_RAW = 1u,       /* bytes/255 throughout (synthetic data) */
    NORM_OUT_SRGB = 2u      /* or-ed in: the finished frame is sRGB-encoded as the reference's surface does on
                               store (gpu_state.rs:94-99 picks the first sRGB surface format; texture.rs:36 is the
                               matching decode at load), i.e. what a norm user sees on screen.  Accepted by
                               norm_cuda_effects, norm_cuda_chain[_async|_host|_band*] and norm_cuda_band_step_async.
                               RGB through an exact 256-entry table after the last stage, alpha untouched; the
                               chain's SSIM is of the linear result.  One extra in-place pass (8 B/pixel). */
};

/* Decoded PNG as the reference's `Png` holds it (src/png/grammar.rs:65-70): header fields,
 * raw gAMA, and `pixel_buffer` = unfiltered scanlines without filter bytes, raster order,
 * row stride ceil(width*channels*bit_depth/8).  plte / trns carry the PLTE and tRNS chunk
 * bodies the reference drops (decoder.rs:64-76,204-209); NULL/0 when absent. */
typedef struct NormPngDesc {
    uint32_t width, height;
    uint8_t bit_depth;   /* 1,2,4,8,16 */
    uint8_t color_type;  /* 0,2,3,4,6 (src/image/grammar.rs:10-17) */
    uint8_t _reserved[2];
    uint32_t gamma;
    const uint8_t *pixels;
    size_t len;
    const uint8_t *plte;
    int32_t n_plte;      /* entries (3 bytes each), <= 256 */
    const uint8_t *trns;
    int32_t n_trns;      /* bytes, <= 256 */
} NormPngDesc;

enum {
    NORM_EXPAND_COMPAT = 0, /* bit-exact `Png::rgba8()` (grammar.rs:123-155): 8-bit ct 0/2/4/6, alpha 0 for opaque types */
    NORM_EXPAND_SPEC = 1    /* every PngSuite colour type / depth, PLTE + tRNS, 16->8 via (v+128)/257, opaque alpha 255 */
};

typedef struct norm_ctx norm_ctx;
typedef struct norm_img norm_img; /* `frames` device-resident RGBA8 images of equal size */

/* ---- library / context ------------------------------------------------------------- */

/* ABI version of the loaded library. */
int32_t norm_cuda_abi_version(void);
/* Static description of a status code. */
const char *norm_cuda_status_string(norm_status s);

/* Replaces GpuResourceAllocator::new (src/renderer/gpu_state.rs:49-132): binds a context to
 * CUDA device `device` and creates its streams.  Fails with NORM_E_CUDA if no device. */
norm_status norm_cuda_create(int32_t device, norm_ctx **out);
/* As above but all work is issued on the caller's stream (a cudaStream_t cast to void*),
 * so the caller's own CUDA events bracket the kernels. */
norm_status norm_cuda_create_on_stream(int32_t device, void *cuda_stream, norm_ctx **out);
void norm_cuda_destroy(norm_ctx *ctx);
/* Last error text of this context ("" if none); valid until the next call on the context.
 * ctx may be NULL: returns the text of the last failed norm_cuda_create on this thread. */
const char *norm_cuda_last_error(const norm_ctx *ctx);
/* Blocks until all work issued on the context has finished (queue.submit + wait analogue,
 * gpu_state.rs:593-595). */
norm_status norm_cuda_sync(norm_ctx *ctx);
/* Number of kernels this context has launched since creation. */
uint64_t norm_cuda_launch_count(const norm_ctx *ctx);
/* Device facts for reports: SM count, bytes of device memory, name (<= 127 chars). */
norm_status norm_cuda_device_info(const norm_ctx *ctx, int32_t *sm_count, size_t *total_mem,
                                  char *name, size_t name_cap);

/* GPU timing on the context's stream: start/stop bracket a region with CUDA events;
 * elapsed_ms synchronises on the stop event. */
norm_status norm_cuda_timer_start(norm_ctx *ctx);
norm_status norm_cuda_timer_stop(norm_ctx *ctx);
norm_status norm_cuda_timer_elapsed_ms(norm_ctx *ctx, float *ms);

/* ---- image buffers ----------------------------------------------------------------- */

/* Replaces create_storage_texture / Texture::from_image's device allocation
 * (gpu_state.rs:597-623, texture.rs:37-46): `frames` pitched RGBA8 images. */
norm_status norm_cuda_image_create(norm_ctx *ctx, uint32_t width, uint32_t height, uint32_t frames,
                                   norm_img **out);
/* Wraps caller-owned device memory (not freed by image_destroy).  pitch % 16 == 0,
 * frame_stride >= pitch*height and % 16 == 0, ptr 16-byte aligned. */
norm_status norm_cuda_image_wrap(norm_ctx *ctx, void *device_ptr, uint32_t width, uint32_t height,
                                 uint32_t frames, size_t pitch, size_t frame_stride, norm_img **out);
void norm_cuda_image_destroy(norm_ctx *ctx, norm_img *img);
norm_status norm_cuda_image_info(const norm_img *img, uint32_t *width, uint32_t *height,
                                 uint32_t *frames, size_t *pitch, size_t *frame_stride,
                                 void **device_ptr);

/* Replaces queue.write_texture in Texture::from_image (src/renderer/texture.rs:48-62):
 * copies `nframes` tightly-or-strided host RGBA8 frames (host_pitch bytes per row, frames
 * back to back) into frames [frame0, frame0+nframes).  Batched: one 2-D copy per call when
 * rows are contiguous.  Pinned host memory (norm_cuda_host_alloc) makes it asynchronous. */
norm_status norm_cuda_upload(norm_ctx *ctx, norm_img *img, uint32_t frame0, uint32_t nframes,
                             const uint8_t *host_rgba8, size_t host_pitch);
/* The "RGBA8 buffer out" half of the boundary (the reference never reads back,
 * app_state.rs:567-661; this is new). */
norm_status norm_cuda_download(norm_ctx *ctx, const norm_img *img, uint32_t frame0, uint32_t nframes,
                               uint8_t *host_rgba8, size_t host_pitch);

/* Pinned host memory for upload / download / *_host entry points. */
norm_status norm_cuda_host_alloc(norm_ctx *ctx, size_t bytes, void **out);
void norm_cuda_host_free(norm_ctx *ctx, void *p);

/* ---- pixel expansion ---------------------------------------------------------------- */

/* Replaces `ImageExt::rgba8()` for `Png` (src/png/grammar.rs:123-155): uploads the packed
 * scanlines and expands them on the GPU into frame `frame` of `out` (width/height must
 * match).  NORM_E_UNSUPPORTED in compat mode for inputs the reference cannot expand. */
norm_status norm_cuda_expand_rgba8(norm_ctx *ctx, const NormPngDesc *png, int32_t mode, norm_img *out,
                                   uint32_t frame);
/* As norm_cuda_expand_rgba8 with the scanlines already in DEVICE memory (png->pixels is a device pointer: the GPU
 * unfilter's output, or a decoder that inflates on the device): `frames` images of one geometry, rows row_pitch
 * bytes apart (a multiple of 16, base 16-byte aligned, and at least 48 readable bytes after the last row: the last
 * item of a ragged row reads past it), images frame_stride bytes apart, into frames [frame0, frame0 + frames) of
 * `out`.  ONE launch for the whole batch, asynchronous on the context's stream.  png->len is the length of one image. */
norm_status norm_cuda_expand_rgba8_device(norm_ctx *ctx, const NormPngDesc *png, size_t row_pitch, size_t frame_stride,
                                          uint32_t frames, int32_t mode, norm_img *out, uint32_t frame0);
/* Same, result copied to a host buffer of 4*width*height bytes (the Cow<[u8]> the
 * reference returns). */
norm_status norm_cuda_expand_rgba8_host(norm_ctx *ctx, const NormPngDesc *png, int32_t mode,
                                        uint8_t *out_rgba8);

/* ---- scanline unfilter + Adam7 (the step before expansion) ----------------- */

/* The zlib-inflated IDAT stream exactly as `ScanlineReader::new` receives it
 * (src/png/scanline_reader.rs:15-28): per row one filter-type byte followed by the filtered
 * bytes; for interlace == 1 the seven Adam7 pass images back to back (PNG spec 8.2). */
typedef struct NormInflatedDesc {
    uint32_t width, height;
    uint8_t bit_depth, color_type, interlace, _reserved;
    const uint8_t *data;
    size_t len;
} NormInflatedDesc;

/* Replaces ScanlineReader::read_lines (src/png/scanline_reader.rs:30-168; Adam7 geometry
 * src/png/interlace.rs:35-83): undoes the None/Sub/Up/Average/Paeth filters and the interlacing
 * on the GPU and returns `Png::pixel_buffer` (height * ceil(width*channels*bit_depth/8) bytes,
 * raster order) in host memory.  PNG-spec behaviour for the cases the reference leaves
 * unimplemented (interlaced Up/Average/Paeth, sub-byte depths).  NORM_E_INVALID_ARG for a
 * filter byte > 4 ("Unrecognized filter method", grammar.rs:48-62) or a wrong stream length. */
norm_status norm_cuda_unfilter(norm_ctx *ctx, const NormInflatedDesc *inflated, uint8_t *pixel_buffer);
/* Unfilter and expand without the pixel_buffer ever leaving the device: inflated stream in,
 * RGBA8 in frame `frame` of `out` (ScanlineReader::read_lines + ImageExt::rgba8 in one call). */
norm_status norm_cuda_unfilter_expand_rgba8(norm_ctx *ctx, const NormInflatedDesc *inflated, const uint8_t *plte,
                                            int32_t n_plte, const uint8_t *trns, int32_t n_trns, int32_t mode,
                                            norm_img *out, uint32_t frame);
/* Batch form: n streams of identical geometry (same width, height, bit depth, colour type and
 * interlace flag -- the frames of a sequence), image i expanded into frame frame0 + i of `out`.
 * One CTA per (image, Adam7 pass) runs the wavefront, so a batch scales over the SMs where a single
 * image cannot.  Uses n * (stream + unfiltered) bytes of scratch.  The reference decodes one file
 * at a time (decoder.rs:29-113); this is the same call made n times. */
norm_status norm_cuda_unfilter_expand_rgba8_batch(norm_ctx *ctx, const NormInflatedDesc *inflated, uint32_t n,
                                                  const uint8_t *plte, int32_t n_plte, const uint8_t *trns,
                                                  int32_t n_trns, int32_t mode, norm_img *out, uint32_t frame0);

/* ---- scanline filtering for the encoder (the step after the path) ---------- */

/* Replaces ScanlineWriter::write / test_filters (src/png/scanline_writer.rs:22-104): per row the five
 * filter candidates, the one with the smallest sum of output bytes wins (first minimum in the order
 * None, Sub, Up, Average, Paeth), and the stream `filter byte + filtered row` is returned in host memory
 * (height * (1 + row_bytes) bytes) ready for deflate.  filter_types (height bytes) may be NULL.
 * paeth_compat = 0 uses the PNG-spec Paeth predictor; 1 reproduces the reference's `a + b + c`
 * (scanline_writer.rs:8), whose type-4 rows no decoder can undo - parity checks only. */
norm_status norm_cuda_filter_scanlines(norm_ctx *ctx, uint32_t width, uint32_t height, uint8_t bit_depth,
                                       uint8_t color_type, const uint8_t *pixel_buffer, int32_t paeth_compat,
                                       uint8_t *stream, uint8_t *filter_types);
/* Same for frame `frame` of a device-resident RGBA8 image (colour type 6, 8 bit): an effect result can
 * leave as PNG without its raw pixels crossing PCIe first. */
norm_status norm_cuda_filter_image(norm_ctx *ctx, const norm_img *img, uint32_t frame, int32_t paeth_compat,
                                   uint8_t *stream, uint8_t *filter_types);

/* ---- effects ------------------------------------------------------------------------ */

/* One frame of the reference renderer, headless: EffectPipeline::execute
 * (src/renderer/effect_pipeline.rs:39-51) over the effects wired in app_state.rs:95-125
 * (gamma always; grayscale; invert) followed by the ONE fragment effect fs_main selects,
 * priority blur > sharpen > edge_detect (image_shader.wgsl:158-174).  Applied to every
 * frame of `in`; `out` must have the same geometry and may not alias `in`. */
norm_status norm_cuda_effects(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags,
                              const norm_img *in, norm_img *out);

/* The chained path of BASELINE.json: front (gamma/grayscale/invert) -> blur -> sharpen ->
 * edge_detect, every enabled stage reading and writing RGBA8 as the reference's rgba8unorm
 * storage textures do (gpu_state.rs:608), then SSIM(result, in) per frame
 * (Png::compute_sim, src/png/ssim.rs:91-106).  ssim: host array of `frames` doubles or NULL. */
norm_status norm_cuda_chain(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags,
                            const norm_img *in, norm_img *out, double *ssim);
/* As norm_cuda_chain but returns without waiting; ssim_dev (device array of `frames` doubles,
 * may be NULL) receives the SSIM values on the context's stream.
 * All `_async` entry points: scratch of the context (SSIM partial sums, stage temporaries) grows on demand, so the
 * first call with a batch larger than any before it may call cudaMalloc and synchronise the device; later calls of
 * that size or smaller allocate nothing. */
norm_status norm_cuda_chain_async(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags,
                                  const norm_img *in, norm_img *out, double *ssim_dev);

/* Host-buffer form of norm_cuda_chain: `frames` RGBA8 frames in host memory in, RGBA8 out
 * (tightly packed, 4*width bytes per row).  Uploads, runs and downloads in overlapping
 * chunks on three internal streams of the context (NOT the context's own stream: a context created with
 * norm_cuda_create_on_stream does not see this work on its stream).  Fully synchronous: on return, and on failure
 * too, no copy into or out of the caller's buffers is still in flight.  Use pinned buffers for full PCIe speed. */
norm_status norm_cuda_chain_host(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags,
                                 const uint8_t *in_rgba8, uint8_t *out_rgba8, uint32_t frames,
                                 double *ssim);

/* ---- one giant image, row-band sharded over several GPUs ------------------ */

/* The chain on one rank's band.  `in` / `out` hold the band plus its stencil halo rows
 * (norm_plan_band: halo_up + rows + halo_down rows of the full image, full width) and
 * FeatureUniform.height is that extended height.  Output rows of the halo are scratch; rows
 * [own_row0, own_row0 + own_rows) are exact (a halo row only ever feeds halo outputs, and
 * ClampToEdge applies at the true image edge because bands at the edge have no halo there).
 * moments (host, 5 doubles per frame, may be NULL): the SSIM moments of the OWNED rows only,
 * to be summed over ranks (ncclAllReduce, norm_nccl.h) and finished with
 * norm_ssim_from_moments(sum, width * full_height).  Stands where Png::compute_sim's passes
 * over the whole image do (src/png/ssim.rs:18-183). */
norm_status norm_cuda_chain_band(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags,
                                 const norm_img *in, norm_img *out, uint32_t own_row0, uint32_t own_rows,
                                 double *moments);

/* As norm_cuda_chain_band but returns without waiting; moments_dev (device, 5 doubles per frame, may
 * be NULL) receives the owned-row moments on the context's stream, ready for an in-place
 * norm_nccl_allreduce_sum_f64 on that same stream (norm_cuda_stream) - no host round trip. */
norm_status norm_cuda_chain_band_async(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags,
                                       const norm_img *in, norm_img *out, uint32_t own_row0,
                                       uint32_t own_rows, double *moments_dev);
/* The context's stream as a cudaStream_t (to order the caller's own work, e.g. NCCL, with it). */
norm_status norm_cuda_stream(norm_ctx *ctx, void **cuda_stream);

/* Peer access between the one-process-per-GPU ranks: export an owned image as a 64-byte CUDA
 * IPC handle, open a neighbour's handle as a device pointer valid in this process (peer access
 * over NVLink is enabled on open), close it when done.  No reference counterpart (the
 * reference is single-device). */
norm_status norm_cuda_ipc_export(norm_ctx *ctx, const norm_img *img, uint8_t handle[64]);
norm_status norm_cuda_ipc_open(norm_ctx *ctx, const uint8_t handle[64], void **device_ptr);
norm_status norm_cuda_ipc_close(norm_ctx *ctx, void *device_ptr);
/* Halo exchange: copies `nrows` rows starting at src_row of a (local or peer-mapped) RGBA8
 * buffer with pitch src_pitch into rows [dst_row, dst_row+nrows) of frame 0 of `dst`, on the
 * context's stream (asynchronous; order with norm_cuda_sync).  The copy is ordered with THIS
 * context's stream only: nothing orders it with the peer process that produces the source rows.
 * Callers whose source changes use the norm_cuda_band_* calls below, which carry that ordering
 * (interprocess events); with this call alone the source must be complete before it is queued. */
norm_status norm_cuda_copy_rows(norm_ctx *ctx, norm_img *dst, uint32_t dst_row, const void *src_device_ptr,
                                size_t src_pitch, uint32_t src_row, uint32_t nrows);

/* ---- the row-band scheduler behind one handle ----------------------------------------------
 * No reference counterpart (the reference is single-device); the contract is
 * BASELINE.json configs[4].  One norm_band per rank (one process per GPU, or several ranks in
 * one process).  A rank owns rows [row0, row0 + rows) of the width x height image; `src` / `dst`
 * (norm_cuda_band_src / _dst) are its extended band images: halo_up + rows + halo_down rows, the
 * owned rows starting at row halo_up.  Life cycle:
 *
 *   norm_cuda_band_create            plan (norm_plan_band), allocate src/dst, streams, events
 *   norm_cuda_band_export            256-byte handle: CUDA IPC handles of src and of two events
 *   [exchange the handles of all ranks through any host side channel]
 *   norm_cuda_band_connect           open the neighbours' memory and events (peer access on)
 *   [write the owned rows of src: norm_cuda_copy_rows / norm_cuda_upload / a kernel on the
 *    context's stream], then norm_cuda_band_publish
 *   [host barrier over the ranks: every rank has published]
 *   norm_cuda_band_step_async        one pass of the chain over the band, see below
 *
 * norm_cuda_band_step_async queues, without ever blocking the host:
 *   - on the band's halo stream: wait for the neighbours' published rows, copy the halo rows over
 *     NVLink (cudaMemcpy2DAsync, one copy per neighbour and side), record "halo rows are in";
 *   - on the context's stream: the chain on the INTERIOR output rows (their stencil touches no
 *     halo row) at once, then - after "halo rows are in" - the two edge strips;
 *   - the SSIM moments (5 doubles) of the owned rows into moments_dev (device memory, may be
 *     NULL), summed in the fixed order interior, top strip, bottom strip;
 *   - on the band's side stream (norm_cuda_band_side_stream): a wait for those moments, so that a
 *     norm_nccl_allreduce_sum_f64 queued there runs beside the next step instead of inside it.
 *     The caller alternates between two moments buffers when it queues steps back to back.
 * Settings no fused kernel covers run the whole extended band after the halo rows are in.
 *
 * When the owned rows CHANGE from step to step: norm_cuda_band_begin_update (the context's
 * stream waits until every neighbour has finished copying out of this band), write the rows,
 * norm_cuda_band_publish, host barrier, norm_cuda_band_step_async, host barrier.  The two host
 * barriers order the ENQUEUE of the event records and waits between processes (a CUDA event
 * carries no step count); they do not wait for the GPU. */
typedef struct norm_band norm_band;
typedef struct NormBandHandle {
    uint8_t mem[64];          /* cudaIpcMemHandle_t of the band's src image */
    uint8_t ready_event[64];  /* cudaIpcEventHandle_t: owned rows written (norm_cuda_band_publish) */
    uint8_t halo_event[64];   /* cudaIpcEventHandle_t: this rank's halo copies are done */
    uint64_t pitch;           /* bytes per row of src */
    uint64_t local_ptr, local_ready, local_halo, pid; /* the same three objects for ranks inside one process */
    uint32_t row0, rows, halo_up, halo_down;
    int32_t rank, device;
} NormBandHandle;
norm_status norm_cuda_band_create(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t width, uint32_t height,
                                  int32_t nranks, int32_t rank, norm_band **out);
norm_status norm_cuda_band_destroy(norm_band *band);
norm_status norm_cuda_band_layout(const norm_band *band, uint32_t *row0, uint32_t *rows, uint32_t *halo_up,
                                  uint32_t *halo_down);
norm_img *norm_cuda_band_src(norm_band *band);
norm_img *norm_cuda_band_dst(norm_band *band);
norm_status norm_cuda_band_export(norm_band *band, NormBandHandle *handle);
norm_status norm_cuda_band_connect(norm_band *band, const NormBandHandle *all_ranks, int32_t nranks);
norm_status norm_cuda_band_begin_update(norm_band *band);
norm_status norm_cuda_band_publish(norm_band *band);
norm_status norm_cuda_band_step_async(norm_band *band, uint32_t flags, double *moments_dev);
norm_status norm_cuda_band_side_stream(norm_band *band, void **cuda_stream);
/* Waits for everything the band has queued (halo, context and side streams). */
norm_status norm_cuda_band_sync(norm_band *band);

/* Single stages, for callers that compose their own pipeline (EffectPipelineBuilder::
 * add_effect, effect_pipeline.rs:92-123) and for per-stage parity tests. */
/* gamma_correct_compute.wgsl:15-42 + grayscale_compute.wgsl:11-29 + invert_compute.wgsl:11-33 */
norm_status norm_cuda_front(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags,
                            const norm_img *in, norm_img *out);
/* gaussian_blur, image_shader.wgsl:83-103 */
norm_status norm_cuda_blur(norm_ctx *ctx, uint32_t radius, const norm_img *in, norm_img *out);
/* detect_edge, image_shader.wgsl:132-156 */
norm_status norm_cuda_sobel(norm_ctx *ctx, const norm_img *in, norm_img *out);
/* sharpen, image_shader.wgsl:120-129 */
norm_status norm_cuda_sharpen(norm_ctx *ctx, uint32_t factor, const norm_img *in, norm_img *out);
/* box_blur, image_shader.wgsl:105-118: 3x3 mean of all four channels.  Defined but never called
 * in the reference (no FeatureUniform field selects it), so it is only reachable here. */
norm_status norm_cuda_box_blur(norm_ctx *ctx, const norm_img *in, norm_img *out);

/* ---- shape overlay, composite, crosshair ------------------------------- */

/* shape_uniform.rs:3-10 */
typedef struct NormShapeUniform {
    uint32_t width, height;    /* window size == image size (app_state.rs:79) */
    uint32_t num_circles;      /* <= NORM_MAX_CIRCLES (shape_uniform.rs:60) */
    uint32_t selected_circle;  /* index of the selected circle, 0xFFFFFFFF = none */
} NormShapeUniform;
/* shape_uniform.rs:38-45: centre and radius in normalised window coordinates
 * (x / width, y / height, radius / min(width, height); shape.rs:188-243) */
typedef struct NormCircleData {
    float x, y, radius, _padding;
} NormCircleData;
/* draw_uniform.rs:3-9.  circle_center_* / circle_radius are in pixels.  camera_view_proj belongs to
 * the vertex stage (pan / zoom of the presented quad) and is not read here. */
typedef struct NormDrawUniform {
    uint32_t crosshair;
    float circle_center_x, circle_center_y, circle_radius;
    float camera_view_proj[4][4];
} NormDrawUniform;
#define NORM_MAX_CIRCLES 256u

/* Replaces the shape render pass (shape_shader.wgsl:41-89: filled circles with smoothstep edge,
 * outline on the selected one, blended in list order) and the tail of the image fragment shader
 * (image_shader.wgsl:176-193: composite of the shape layer, then the crosshair disc) applied to
 * every frame of `in`; the result goes to the headless RGBA8 buffer `out` (no window, swapchain or
 * present).  circles: host array of shapes->num_circles entries.  draw may be NULL (no crosshair).
 * The shape layer passes through an sRGB 8-bit render texture as in the reference (gpu_state.rs:
 * 404-418); pixels that nothing touches keep their bytes.  NORM_E_DIM_MISMATCH unless
 * shapes->width/height equal the image size; NORM_E_INVALID_ARG for more than NORM_MAX_CIRCLES.
 * Texture-space operation: apply it before FlipX / FlipY, as the reference's vertex stage does. */
norm_status norm_cuda_overlay(norm_ctx *ctx, const NormShapeUniform *shapes, const NormCircleData *circles,
                              const NormDrawUniform *draw, const norm_img *in, norm_img *out);

/* ---- SSIM --------------------------------------------------------------------------- */

/* Replaces Png::compute_sim (src/png/ssim.rs:91-106) on expanded RGBA8 images (RGBA luma
 * branch, ssim.rs:65-81): one global-window SSIM per frame, a[i] against b[i].
 * NORM_E_DIM_MISMATCH if geometries differ (ssim.rs:92-95).  ssim: `frames` doubles. */
norm_status norm_cuda_ssim(norm_ctx *ctx, const norm_img *a, const norm_img *b, double *ssim);
/* Png::compute_sim for images of any colour type: Png::luma_buffer picks the luma by colour type
 * (src/png/ssim.rs:18-87) and the two images may differ in it.  `a` / `b` hold the RGBA8 expansion. */
enum {
    NORM_LUMA_RGB = 0,       /* RGB / RGBA: r*0.29891 + g*0.58661 + b*0.11448 on raw bytes (ssim.rs:56-58, 73-75) */
    NORM_LUMA_GRAY = 1,      /* Grayscale: the sample itself (ssim.rs:25) */
    NORM_LUMA_GRAY_UNIT = 2  /* GrayscaleAlpha: sample / 255.0 (ssim.rs:41: the reference mixes scales; kept) */
};
norm_status norm_cuda_ssim_luma(norm_ctx *ctx, const norm_img *a, int32_t luma_a, const norm_img *b, int32_t luma_b,
                                double *ssim);
/* The five raw moments per frame {sum x, sum y, sum x^2, sum y^2, sum xy} of the f32 lumas,
 * accumulated in f64 -- what a multi-GPU caller all-reduces before finalising.
 * moments: host array of 5*frames doubles. */
norm_status norm_cuda_ssim_moments(norm_ctx *ctx, const norm_img *a, const norm_img *b, double *moments);
/* Host-only: SSIM from (all-reduced) moments over npix pixels (ssim.rs:127-183 in f64). */
double norm_ssim_from_moments(const double moments[5], double npix);

/* Has no counterpart in the reference.  SSIM for callers that are not allowed to allocate device
 * memory: a render / processing loop that must stay clear of the device allocator while frames
 * are in flight (cudaMalloc and cudaFree stall the whole GPU), or a host that has given all
 * device memory to its own pool so that a cudaMalloc from this library would fail.
 * Same result as norm_cuda_ssim / norm_cuda_ssim_moments, but the partial sums go to a fixed
 * 160 KiB array built into the library (resident on every GPU that loads it, until exit) and
 * the call creates nothing: no context, no cudaMalloc/cudaFree, no pinned memory.
 *   a_dev, b_dev   device pointers to `frames` pitched RGBA8 frames (same geometry for both;
 *                  pitch % 16 == 0, frame_stride % 16 == 0, 16-byte aligned), e.g. from
 *                  norm_cuda_image_info or the caller's own pool;
 *   ssim_dev       device array of `frames` doubles, or NULL;
 *   moments_dev    device array of 5*frames doubles, or NULL (at least one of the two);
 *   cuda_stream    cudaStream_t cast to void* (NULL = default stream); the work is queued on
 *                  it and the call returns without waiting.
 * Limits and rules:
 *   - frames <= NORM_SSIM_STATIC_MAX_FRAMES per call, otherwise NORM_E_INVALID_ARG with a message
 *     saying so (split the batch); any width / height.
 *   - The device that is current when you call is the one used.  The first call on each device
 *     looks up the array's address (cudaGetSymbolAddress); make that first call outside any
 *     stream capture.
 *   - All calls on a device share the one array.  Two calls whose GPU work can overlap (different
 *     streams, different threads) corrupt each other's sums: keep them on one stream or order
 *     them yourself.  Nothing carries over between calls - each launch overwrites the slots it
 *     reads - so there is nothing to reset.
 *   - Error text: norm_cuda_last_error(NULL) on the calling thread. */
#define NORM_SSIM_STATIC_MAX_FRAMES 1024u
norm_status norm_cuda_ssim_static(const void *a_dev, const void *b_dev, uint32_t width, uint32_t height,
                                  uint32_t frames, size_t pitch, size_t frame_stride, double *ssim_dev,
                                  double *moments_dev, void *cuda_stream);

/* ---- host-only planning helpers (usable without a GPU) ------------------------------ */

/* Normalised 1-D taps of the separable blur and the texel offset of tap 0
 * (image_shader.wgsl:81-103 under the floor convention). */
norm_status norm_blur_taps(uint32_t radius, float *taps, int32_t *dmin);
/* The table NORM_OUT_SRGB applies: code of a stored linear unorm8 byte under the surface's sRGB encode. */
norm_status norm_srgb_encode_lut(uint8_t lut[256]);
/* 256-entry front table: sRGB decode at load + gamma pass + unorm8 store. */
norm_status norm_front_lut(uint32_t flags, uint32_t gamma, uint8_t lut[256]);
/* sRGB tables of the shape render texture: eotf[k] = decoded value of code k; thresh[j] = linear
 * value at which the encoder steps from code j-1 to code j (thresh[0] = 0). */
norm_status norm_srgb_tables(float eotf[256], float thresh[256]);
/* Row-band plan for one giant image over `nranks` GPUs: rank `rank` owns rows
 * [*row0, *row0 + *rows) and needs *halo_up / *halo_down source rows from its neighbours
 * for the enabled stencil stages of `f`. */
norm_status norm_plan_band(const NormFeatureUniform *f, uint32_t height, int32_t nranks, int32_t rank,
                           uint32_t *row0, uint32_t *rows, uint32_t *halo_up, uint32_t *halo_down);
/* Frame-shard plan: rank owns frames [*frame0, *frame0 + *nframes) of `frames`. */
norm_status norm_plan_frames(uint32_t frames, int32_t nranks, int32_t rank, uint32_t *frame0,
                             uint32_t *nframes);

#ifdef __cplusplus
}
#endif
#endif /* NORM_CUDA_H */
