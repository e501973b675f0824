/*
 * norm_oracle.h -- CPU restatement of friendlymatthew/norm's image-effect hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (norm_b200/libnorm_cuda.so) never links, imports or calls anything in oracle/.
 *
 * Parity status:
 *   - PNG decode + pixel expansion: PINNED by the reference's 13 golden blobs
 *     (test_suite/{basn0g08,basn4a08,basn6a08,f0[0-4]n0g08,f0[0-4]n2c08}, reference tests
 *     src/png/decoder.rs:295-365), by the Adam7 pass-size vectors of
 *     src/png/interlace.rs:89-124, and cross-checked against PIL / cv2.
 *   - Effect shaders (gamma / grayscale / invert / blur / sharpen / Sobel) and SSIM:
 *     PARITY UNPINNED -- the reference has no live test, golden vector or runnable
 *     build for them here (no rustc, no wgpu adapter); the WGSL / Rust source is the only
 *     specification and every function below cites the lines it follows.
 *
 * All citations are relative to /root/reference.
 */
#ifndef NORM_ORACLE_H
#define NORM_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORACLE_OK = 0, ORACLE_E_INVALID = 1, ORACLE_E_UNSUPPORTED = 2, ORACLE_E_CORRUPT = 3 };
enum { ORACLE_EXPAND_COMPAT = 0, ORACLE_EXPAND_SPEC = 1 };
enum { ORACLE_CS_REFERENCE = 0, ORACLE_CS_RAW = 1 };

/* Decoded image: mirrors `Png` (src/png/grammar.rs:65-70) + the PLTE / tRNS data the
 * reference drops (src/png/decoder.rs:64-76,204-209). */
typedef struct {
    uint32_t width, height;
    uint8_t bit_depth, color_type, compression, filter, interlace;
    uint32_t gamma;          /* raw gAMA value, 0 if absent (decoder.rs:62-69) */
    uint8_t *pixels;         /* unfiltered scanlines, no filter bytes, raster order,
                                row stride = ceil(width*channels*bit_depth/8) */
    size_t len;
    uint8_t plte[768];
    int32_t n_plte;          /* entries */
    uint8_t trns[256];
    int32_t n_trns;          /* bytes */
} oracle_png;

/* src/png/decoder.rs:29-113 (flow), 119-222 (chunks), scanline_reader.rs:40-168 (unfilter),
 * interlace.rs:35-83 (pass geometry).  Spec-conformant where the reference is buggy. */
int oracle_png_decode(const uint8_t *data, size_t n, oracle_png *out);
void oracle_png_free(oracle_png *p);
/* interlace.rs:35-83: per-pass (width,height) for the 7 Adam7 passes. */
void oracle_adam7_pass_sizes(uint32_t w, uint32_t h, uint32_t pw[7], uint32_t ph[7]);
size_t oracle_row_bytes(uint32_t w, int color_type, int bit_depth);

/* src/png/scanline_writer.rs:5-104: encoder-side filter selection.  stream = rows*(1+row_bytes) bytes. */
void oracle_filter_scanlines(const uint8_t *pixels, uint32_t rows, size_t row_bytes, size_t bpp, int paeth_compat,
                             uint8_t *stream, uint8_t *filter_types);

/* src/png/grammar.rs:123-155 (compat) / PNG-spec semantics (spec mode).  out = 4*w*h bytes. */
int oracle_expand_rgba8(const oracle_png *p, int mode, uint8_t *out);
/* src/png/grammar.rs:89-121: rgb8() -- the view the reference's own tests pin. */
int oracle_rgb8(const oracle_png *p, uint8_t *out);

/* Per-channel front LUT: texture-format sRGB decode (texture.rs:36) + gamma pass
 * (gamma_correct_compute.wgsl:25-40) + unorm8 store. */
void oracle_front_lut(int colorspace, uint32_t gamma, uint8_t lut[256]);
/* S-form front stage: LUT on rgb, optional luma (grayscale_compute.wgsl:22-28),
 * optional invert (invert_compute.wgsl:22-32); quantised between stages as the reference's
 * rgba8unorm storage textures do (gpu_state.rs:597-623). */
void oracle_front_s(const uint8_t *in, uint32_t w, uint32_t h, const uint8_t lut[256],
                    int grayscale, int invert, uint8_t *out);
/* R-form (literal unit-scale float arithmetic of the shaders). */
void oracle_front_r(const uint8_t *in, uint32_t w, uint32_t h, int colorspace, uint32_t gamma,
                    int grayscale, int invert, uint8_t *out);

/* image_shader.wgsl:81-103.  Normalised 1-D tap table of the S-form: taps[k] applies to
 * texel offset (*dmin + k); n = radius taps. */
int oracle_blur_taps(uint32_t radius, float *taps, int *dmin);
int oracle_blur_s(const uint8_t *in, uint32_t w, uint32_t h, uint32_t radius, uint8_t *out);
int oracle_blur_r(const uint8_t *in, uint32_t w, uint32_t h, uint32_t radius, uint8_t *out);

/* image_shader.wgsl:132-156 */
void oracle_sobel_s(const uint8_t *in, uint32_t w, uint32_t h, uint8_t *out);
void oracle_sobel_r(const uint8_t *in, uint32_t w, uint32_t h, uint8_t *out);
/* image_shader.wgsl:120-129 */
void oracle_sharpen_s(const uint8_t *in, uint32_t w, uint32_t h, uint32_t factor, uint8_t *out);
void oracle_sharpen_r(const uint8_t *in, uint32_t w, uint32_t h, uint32_t factor, uint8_t *out);

/* image_shader.wgsl:105-118 (dead code in the reference: defined, never called) */
void oracle_box_blur_s(const uint8_t *in, uint32_t w, uint32_t h, uint8_t *out);
void oracle_box_blur_r(const uint8_t *in, uint32_t w, uint32_t h, uint8_t *out);

/* Shape overlay (shape_shader.wgsl:41-89), composite and crosshair (image_shader.wgsl:176-193).
 * Structs are field-for-field shape_uniform.rs:3-10, :38-45 and draw_uniform.rs:3-9.
 * Parity unpinned: the reference has no test for it and its render-target format and
 * interpolated tex_coords are driver-defined; this is the restatement with exact pixel centres
 * and an sRGB render texture. */
typedef struct { uint32_t width, height, num_circles, selected_circle; } oracle_shape_uniform;
typedef struct { float x, y, radius, pad; } oracle_circle;
typedef struct {
    uint32_t crosshair;
    float circle_center_x, circle_center_y, circle_radius;
    float camera_view_proj[4][4];
} oracle_draw_uniform;
void oracle_srgb_tables(float eotf[256], float thresh[256]);
/* gpu_state.rs:94-99: the sRGB surface's encode of a stored linear unorm8 byte (table) and of a whole RGBA8 frame */
void oracle_srgb_encode_lut(uint8_t lut[256]);
void oracle_encode_srgb(const uint8_t *in, size_t npix, uint8_t *out);
int oracle_overlay(const uint8_t *in, uint32_t w, uint32_t h, const oracle_shape_uniform *su,
                   const oracle_circle *circles, const oracle_draw_uniform *du, uint8_t *out);

/* src/png/ssim.rs:9-184 on two RGBA8 buffers (RGBA luma branch, ssim.rs:65-81).
 * moments = {sum x, sum y, sum x^2, sum y^2, sum x*y} over f32 lumas, accumulated in f64. */
void oracle_ssim_moments(const uint8_t *a, const uint8_t *b, size_t npix, double m[5]);
double oracle_ssim_from_moments(const double m[5], double npix);
double oracle_ssim(const uint8_t *a, const uint8_t *b, size_t npix);
/* per-colour-type lumas of Png::luma_buffer (ssim.rs:18-87): 0 RGB weights, 1 gray byte, 2 gray byte / 255 */
double oracle_ssim_luma(const uint8_t *a, const uint8_t *b, size_t npix, int mode_a, int mode_b);
/* Literal sequential-f32 evaluation, pass structure of ssim.rs:18-183 (reported, not a target). */
float oracle_ssim_f32_literal(const uint8_t *a, const uint8_t *b, size_t npix);

/* FeatureUniform, field-for-field (src/renderer/feature_uniform.rs:1-17), 112 bytes. */
typedef struct {
    uint32_t grayscale, invert, gamma, blur, blur_radius, width, height, sharpen, sharpen_factor,
        edge_detect;
    uint8_t pad[8];
    float transform[4][4];
} oracle_feature_uniform;

/* Reference frame semantics: compute chain (app_state.rs:95-125) then ONE fragment effect,
 * priority blur > sharpen > edge (image_shader.wgsl:158-174). */
int oracle_effects(const oracle_feature_uniform *f, int colorspace, const uint8_t *in, uint8_t *out);
/* Derived chain: front -> blur -> Sobel, each stage RGBA8; SSIM(result, in). */
int oracle_chain(const oracle_feature_uniform *f, int colorspace, const uint8_t *in, uint8_t *out,
                 double *ssim);

/* Batch of independent frames (BASELINE config 4), frames in parallel. */
int oracle_chain_batch(const oracle_feature_uniform *f, int colorspace, const uint8_t *in,
                       uint8_t *out, double *ssim, int nframes);

int oracle_num_threads(void);
void oracle_set_num_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
