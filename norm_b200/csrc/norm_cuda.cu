// libnorm_cuda.so -- C ABI (include/norm_cuda.h) over the sm_100a kernels.
// Host side of the CUDA backend: contexts, pitched RGBA8 image buffers, uploads, effect
// dispatch (the EffectPipeline::execute analogue), SSIM, and the host-buffer pipeline.
// No CPU fallback exists anywhere in this file: without a device every compute call fails.
#include "../../include/norm_cuda.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <unistd.h>
#include <type_traits>
#include <vector>

#include "blur_kernels.cuh"
#include "chain_kernels.cuh"
#include "chain2_kernels.cuh"
#include "common.cuh"
#include "expand_kernels.cuh"
#include "filter_kernels.cuh"
#include "overlay_kernels.cuh"
#include "stencil_kernels.cuh"
#include "chainw_kernels.cuh"
#include "stage_kernels.cuh"
#include "unfilter_kernels.cuh"

using namespace normk;

// ---------------------------------------------------------------------------------------
// objects
// ---------------------------------------------------------------------------------------

struct norm_img {
    ImgView v;
    bool owned;
};

namespace {

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
};

// Everything one stream needs to run the chain: ping-pong temporaries, SSIM partials,
// and (for the host-buffer pipeline) device staging for a chunk of frames.
struct Slot {
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    DevBuf tmp[2];
    DevBuf partials;
    DevBuf stage_in, stage_out;
    // row window of the next fused launches (band mode): produce output rows [win_y0, win_y1) only (0, 0 = all rows)
    // and add the window's SSIM moments to what d_moments already holds instead of overwriting it
    uint32_t win_y0 = 0, win_y1 = 0;
    bool accumulate = false;
};

thread_local std::string g_create_error;

}  // namespace

struct norm_ctx {
    int device = 0;
    int sm_count = 0;
    size_t total_mem = 0;
    char name[128] = {0};
    Slot main;
    Slot pipe[3];
    bool pipe_ready = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    uint64_t launches = 0;
    DevBuf raw;         // expansion: packed scanlines
    DevBuf flip_tmp;    // effects: pre-flip frame
    DevBuf inflated, passes;  // unfilter: filtered stream, unfiltered Adam7 pass images
    int32_t *d_status = nullptr;
    DevBuf results;     // per-frame moments (5) + ssim (1), doubles
    double *h_results = nullptr;
    size_t h_results_cap = 0;
    int balanced_partition = 1;  // NORM_CUDA_BALANCED=0: static (band, strip) items only; 2: balanced ranges whenever possible (tests)
    uint32_t force_bands = 0;    // NORM_CUDA_BANDS: row bands per frame of the fused kernels (experiments)
    size_t chunk_bytes = 0;      // NORM_CUDA_CHUNK_BYTES: frames per chunk of the staged paths (tests: many small chunks)
    int fused_gen = 0;           // NORM_CUDA_FUSED_GEN=1 forces the first-generation fused kernel
    int win3 = 1;                // NORM_CUDA_WIN3=0: chain and blur at r = 3 run the pipelined kernel instead of the register-window kernels
    bool force_unfused = false;  // NORM_CUDA_FORCE_UNFUSED=1: run the chain as separate stage kernels
};

namespace {

norm_status fail(norm_ctx *ctx, norm_status s, const std::string &msg) {
    if (ctx) ctx->err = msg;
    return s;
}

#define CU_TRY(ctx, expr)                                                                           \
    do {                                                                                            \
        cudaError_t e__ = (expr);                                                                   \
        if (e__ != cudaSuccess) {                                                                   \
            (void)cudaGetLastError();                                                               \
            return fail(ctx, e__ == cudaErrorMemoryAllocation ? NORM_E_OOM : NORM_E_CUDA,           \
                        std::string(#expr) + ": " + cudaGetErrorString(e__));                       \
        }                                                                                           \
    } while (0)

#define NORM_TRY(expr)                     \
    do {                                   \
        norm_status s__ = (expr);          \
        if (s__ != NORM_OK) return s__;    \
    } while (0)

norm_status ensure_buf(norm_ctx *ctx, DevBuf &b, size_t need) {
    if (b.cap >= need) return NORM_OK;
    if (b.p) {
        // the buffer may still be in use by queued work
        CU_TRY(ctx, cudaDeviceSynchronize());
        CU_TRY(ctx, cudaFree(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    CU_TRY(ctx, cudaMalloc(&b.p, need));
    b.cap = need;
    return NORM_OK;
}

void free_buf(DevBuf &b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

bool same_geometry(const ImgView &a, const ImgView &b) {
    return a.w == b.w && a.h == b.h && a.frames == b.frames;
}

ImgView sub_frames(const ImgView &v, uint32_t frame0, uint32_t n) {
    ImgView s = v;
    s.ptr = v.ptr + (size_t)frame0 * v.frame_stride;
    s.frames = n;
    return s;
}

ImgView view_of(void *p, uint32_t w, uint32_t h, uint32_t frames) {
    ImgView v;
    v.ptr = static_cast<uint8_t *>(p);
    v.pitch = align_up((size_t)w * 4, 256);
    v.frame_stride = v.pitch * h;
    v.w = w, v.h = h, v.frames = frames;
    return v;
}

bool overlaps(const ImgView &a, const ImgView &b) {
    const uint8_t *a0 = a.ptr, *a1 = a.ptr + (size_t)a.frames * a.frame_stride;
    const uint8_t *b0 = b.ptr, *b1 = b.ptr + (size_t)b.frames * b.frame_stride;
    return a0 < b1 && b0 < a1;
}

unsigned grid_for(const norm_ctx *ctx, size_t blocks_needed, int per_sm) {
    const size_t cap = (size_t)ctx->sm_count * per_sm;
    return (unsigned)std::max<size_t>(1, std::min(blocks_needed, cap));
}

#define LAUNCH_CHECK(ctx)                                                                  \
    do {                                                                                   \
        ++(ctx)->launches;                                                                 \
        cudaError_t e__ = cudaGetLastError();                                              \
        if (e__ != cudaSuccess)                                                            \
            return fail(ctx, NORM_E_CUDA, std::string("kernel launch: ") + cudaGetErrorString(e__)); \
    } while (0)

// ---- host-side tables (plugin logic, not compute) -------------------------------------

// image_shader.wgsl:83-103 under the floor convention: tap k has offset k - r/2, texel
// delta floor(0.5 + offset), weight exp(-offset^2 / (2 sigma^2)), sigma = r/4; normalised.
norm_status make_taps(uint32_t radius, Taps &t) {
    if (radius < 1 || radius > NORM_MAX_BLUR_RADIUS) return NORM_E_INVALID_ARG;
    const double r = (double)radius, sigma = r * 0.25;
    double w[kMaxRadius + 1], sum = 0.0;
    for (uint32_t k = 0; k < radius; ++k) {
        const double off = -r / 2.0 + (double)k;
        w[k] = std::exp(-(off * off) / (2.0 * sigma * sigma));
        sum += w[k];
    }
    std::memset(&t, 0, sizeof(t));
    for (uint32_t k = 0; k < radius; ++k) t.t[k] = (float)(w[k] / sum);
    t.n = (int32_t)radius;
    t.dmin = (int32_t)std::floor(0.5 - r / 2.0);
    return NORM_OK;
}

double srgb_eotf(double c) { return c <= 0.04045 ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4); }

uint8_t unorm8(float v) {
    if (!(v > 0.0f)) return 0;
    if (v > 1.0f) v = 1.0f;
    return (uint8_t)std::rint(v * 255.0f);
}

// texture.rs:36 (sRGB decode at load) + gamma_correct_compute.wgsl:25-40 + unorm8 store
void make_lut(uint32_t flags, uint32_t gamma, Lut256 &lut, bool &identity) {
    identity = true;
    for (unsigned c = 0; c < 256; ++c) {
        const double u = (double)c / 255.0;
        float v = (float)((flags & NORM_CS_RAW) ? u : srgb_eotf(u));
        if (gamma != 0u) {
            const float gamma_value = (float)gamma / 100000.0f;
            const float inv_gamma = 1.0f / gamma_value;
            v = std::pow(v, inv_gamma);
        }
        lut.v[c] = unorm8(v);
        identity = identity && lut.v[c] == c;
    }
}

norm_status check_uniform(norm_ctx *ctx, const NormFeatureUniform *f, const ImgView &in) {
    if (!f) return fail(ctx, NORM_E_INVALID_ARG, "feature uniform is NULL");
    if (f->width != in.w || f->height != in.h)
        return fail(ctx, NORM_E_DIM_MISMATCH, "FeatureUniform width/height differ from the image");
    if (f->blur == 1 && (f->blur_radius < 1 || f->blur_radius > NORM_MAX_BLUR_RADIUS))
        return fail(ctx, NORM_E_INVALID_ARG, "blur_radius outside 1..63");
    return NORM_OK;
}

// ---- stage launchers -------------------------------------------------------------------

norm_status launch_front(norm_ctx *ctx, cudaStream_t s, const ImgView &in, const ImgView &out,
                         const Lut256 &lut, bool use_lut, bool gray, bool inv) {
    const size_t items = (size_t)((in.w + 3) / 4) * in.h * in.frames;
    const unsigned grid = grid_for(ctx, (items + 255) / 256, 8);
#define FRONT_CASE(L, G, I) \
    if (use_lut == L && gray == G && inv == I) front_kernel<L, G, I><<<grid, 256, 0, s>>>(in, out, lut)
    FRONT_CASE(false, false, false);
    FRONT_CASE(false, false, true);
    FRONT_CASE(false, true, false);
    FRONT_CASE(false, true, true);
    FRONT_CASE(true, false, false);
    FRONT_CASE(true, false, true);
    FRONT_CASE(true, true, false);
    FRONT_CASE(true, true, true);
#undef FRONT_CASE
    LAUNCH_CHECK(ctx);
    return NORM_OK;
}

template <int R>
norm_status launch_blur_r(norm_ctx *ctx, cudaStream_t s, const ImgView &in, const ImgView &out, const Taps &taps) {
    const uint32_t tx = (in.w + kBlurTW - 1) / kBlurTW, ty = (in.h + kBlurTH - 1) / kBlurTH;
    const size_t smem = blur_smem_bytes(taps.n);
    CU_TRY(ctx, cudaFuncSetAttribute(blur_tile_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    CU_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, blur_tile_kernel<R>, kBlurThreads, smem));
    if (per_sm < 1) per_sm = 1;
    const unsigned grid = grid_for(ctx, (size_t)tx * ty * in.frames, per_sm);
    blur_tile_kernel<R><<<grid, kBlurThreads, smem, s>>>(in, out, taps, tx, ty);
    LAUNCH_CHECK(ctx);
    return NORM_OK;
}

norm_status launch_blur(norm_ctx *ctx, cudaStream_t s, const ImgView &in, const ImgView &out, uint32_t radius) {
    Taps taps;
    if (make_taps(radius, taps) != NORM_OK) return fail(ctx, NORM_E_INVALID_ARG, "blur radius outside 1..63");
    switch (radius) {
    case 3: return launch_blur_r<3>(ctx, s, in, out, taps);
    case 7: return launch_blur_r<7>(ctx, s, in, out, taps);
    case 21: return launch_blur_r<21>(ctx, s, in, out, taps);
    case 39: return launch_blur_r<39>(ctx, s, in, out, taps);
    // RGBA blur above r = 21 has no pipelined kernel (4 x R running outputs): give the UI's radii the unrolled tile kernel
    case 23: return launch_blur_r<23>(ctx, s, in, out, taps);
    case 25: return launch_blur_r<25>(ctx, s, in, out, taps);
    case 27: return launch_blur_r<27>(ctx, s, in, out, taps);
    case 29: return launch_blur_r<29>(ctx, s, in, out, taps);
    case 31: return launch_blur_r<31>(ctx, s, in, out, taps);
    case 33: return launch_blur_r<33>(ctx, s, in, out, taps);
    case 35: return launch_blur_r<35>(ctx, s, in, out, taps);
    case 37: return launch_blur_r<37>(ctx, s, in, out, taps);
    default: return launch_blur_r<0>(ctx, s, in, out, taps);
    }
}

// 3x3 stencils on a register window (stencil_kernels.cuh), optionally with the front stage applied at load
template <int OP>
norm_status launch_stencil3(norm_ctx *ctx, cudaStream_t s, const ImgView &in, const ImgView &out, const Lut256 *lut,
                            bool use_lut, bool gray, bool inv, int factor) {
    const uint32_t tx = (in.w + kS3Cols - 1) / kS3Cols, ty = (in.h + kS3Rows - 1) / kS3Rows;
    const unsigned grid = grid_for(ctx, ((size_t)tx * ty * in.frames + 7) / 8, 8);
    static const Lut256 none{};
    const Lut256 &l = lut ? *lut : none;
    const uint32_t inv_mask = inv ? 0x00ffffffu : 0u;
    switch ((use_lut ? 1 : 0) | (gray ? 2 : 0)) {
    case 0: stencil3_kernel<OP, 0><<<grid, 256, 0, s>>>(in, out, l, inv_mask, factor, tx, ty); break;
    case 1: stencil3_kernel<OP, 1><<<grid, 256, 0, s>>>(in, out, l, inv_mask, factor, tx, ty); break;
    case 2: stencil3_kernel<OP, 2><<<grid, 256, 0, s>>>(in, out, l, inv_mask, factor, tx, ty); break;
    default: stencil3_kernel<OP, 3><<<grid, 256, 0, s>>>(in, out, l, inv_mask, factor, tx, ty); break;
    }
    LAUNCH_CHECK(ctx);
    return NORM_OK;
}

norm_status launch_sobel(norm_ctx *ctx, cudaStream_t s, const ImgView &in, const ImgView &out) {
    return launch_stencil3<kOpSobel>(ctx, s, in, out, nullptr, false, false, false, 0);
}

norm_status launch_sharpen(norm_ctx *ctx, cudaStream_t s, const ImgView &in, const ImgView &out, uint32_t factor) {
    return launch_stencil3<kOpSharpen>(ctx, s, in, out, nullptr, false, false, false, (int)factor);
}

norm_status launch_box_blur(norm_ctx *ctx, cudaStream_t s, const ImgView &in, const ImgView &out) {
    return launch_stencil3<kOpBox>(ctx, s, in, out, nullptr, false, false, false, 0);
}

// sRGB encode thresholds / decode values of the shape render texture (host, double -> f32)
void make_srgb_tables(float *eotf, float *thresh) {
    for (int k = 0; k < 256; ++k) {
        eotf[k] = (float)srgb_eotf((double)k / 255.0);
        thresh[k] = k == 0 ? 0.0f : (float)srgb_eotf(((double)k - 0.5) / 255.0);
    }
}

// The surface's linear -> sRGB encode (gpu_state.rs:94-99 picks an sRGB surface format, so every colour the fragment
// stage returns is encoded on store): code of a stored linear unorm8 byte c = the code whose decision interval holds
// c/255, found in the same threshold table the overlay uses (exact: no pow at run time).  Alpha is not encoded.
void make_srgb_encode_lut(Lut256 &lut) {
    float eotf[256], thresh[256];
    make_srgb_tables(eotf, thresh);
    for (int c = 0; c < 256; ++c) {
        const float v = (float)c / 255.0f;
        unsigned lo = 0, hi = 255;
        while (lo < hi) {
            const unsigned mid = (lo + hi + 1) >> 1;
            if (v >= thresh[mid]) lo = mid;
            else hi = mid - 1;
        }
        lut.v[c] = (uint8_t)lo;
    }
}

norm_status copy_view(norm_ctx *ctx, cudaStream_t s, const ImgView &in, const ImgView &out) {
    for (uint32_t f = 0; f < in.frames; ++f)
        CU_TRY(ctx, cudaMemcpy2DAsync(out.ptr + (size_t)f * out.frame_stride, out.pitch,
                                      in.ptr + (size_t)f * in.frame_stride, in.pitch, (size_t)in.w * 4, in.h,
                                      cudaMemcpyDeviceToDevice, s));
    return NORM_OK;
}

// SSIM of a[i] vs b[i]; d_moments / d_ssim are device arrays (5*frames / frames) or NULL.
norm_status launch_ssim(norm_ctx *ctx, Slot &slot, const ImgView &a, const ImgView &b, double *d_moments,
                        double *d_ssim, int luma_a = 0, int luma_b = 0) {
    const size_t chunks = (size_t)((a.w + 3) / 4) * a.h;
    size_t want = ((size_t)ctx->sm_count * 8 + a.frames - 1) / a.frames;  // ~8 CTAs per SM in total
    const size_t max_useful = (chunks + 256 * 4 - 1) / (256 * 4);
    uint32_t bpf = (uint32_t)std::max<size_t>(1, std::min(want, max_useful));
    NORM_TRY(ensure_buf(ctx, slot.partials, (size_t)a.frames * bpf * 5 * sizeof(double)));
    double *partials = static_cast<double *>(slot.partials.p);
    // grid.y is limited to 65535: walk the batch in slices
    for (uint32_t f0 = 0; f0 < a.frames; f0 += 65535u) {
        const uint32_t n = std::min(65535u, a.frames - f0);
        const ImgView va = sub_frames(a, f0, n), vb = sub_frames(b, f0, n);
        double *pp = partials + (size_t)f0 * bpf * 5;
        const dim3 grid(bpf, n);
        switch (luma_a * 3 + luma_b) {  // Png::luma_buffer's branch of each image (ssim.rs:18-87)
#define NORM_SSIM_CASE(A_, B_) case A_ * 3 + B_: ssim_moments_kernel<A_, B_><<<grid, 256, 0, slot.stream>>>(va, vb, pp); break;
            NORM_SSIM_CASE(0, 0) NORM_SSIM_CASE(0, 1) NORM_SSIM_CASE(0, 2) NORM_SSIM_CASE(1, 0) NORM_SSIM_CASE(1, 1)
            NORM_SSIM_CASE(1, 2) NORM_SSIM_CASE(2, 0) NORM_SSIM_CASE(2, 1) NORM_SSIM_CASE(2, 2)
#undef NORM_SSIM_CASE
        default: return fail(ctx, NORM_E_INVALID_ARG, "unknown luma mode");
        }
        LAUNCH_CHECK(ctx);
    }
    ssim_finalize_kernel<<<a.frames, 256, 0, slot.stream>>>(partials, bpf, (double)a.w * (double)a.h, d_moments,
                                                             d_ssim);
    LAUNCH_CHECK(ctx);
    return NORM_OK;
}

struct StagePlan {
    bool front = false, use_lut = false, gray = false, inv = false;
    bool blur = false, sharpen = false, edge = false;
    uint32_t radius = 0, factor = 0;
    Lut256 lut;
    int count() const { return (int)front + (int)blur + (int)sharpen + (int)edge; }
};

StagePlan make_plan(const NormFeatureUniform *f, uint32_t flags, bool chain) {
    StagePlan p;
    bool identity;
    make_lut(flags, f->gamma, p.lut, identity);
    p.use_lut = !identity;
    p.gray = f->grayscale == 1;  // feature_uniform.rs:40-42: `== 1`
    p.inv = f->invert == 1;
    p.front = p.use_lut || p.gray || p.inv;
    p.blur = f->blur == 1;
    p.sharpen = f->sharpen == 1;
    p.edge = f->edge_detect == 1;
    p.radius = f->blur_radius;
    p.factor = f->sharpen_factor;
    if (!chain) {
        // image_shader.wgsl:164-174: each effect overwrites `pixels` from texture A, so only
        // the last enabled one in the order edge, sharpen, blur survives.
        if (p.blur) p.sharpen = p.edge = false;
        else if (p.sharpen) p.edge = false;
    }
    return p;
}

// Runs the planned stages in -> out on slot.stream using the slot's ping-pong temporaries,
// `chunk` frames at a time so the temporaries stay small for large batches.
norm_status try_launch_blur_fused(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in, const ImgView &out,
                                  bool *fused);

norm_status run_stages_unfused(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in, const ImgView &out) {
    const int n = p.count();
    if (n == 0) return copy_view(ctx, slot.stream, in, out);
    {
        bool fused = false;  // front -> blur as one pipelined kernel where it is instantiated
        NORM_TRY(try_launch_blur_fused(ctx, slot, p, in, out, &fused));
        if (fused) return NORM_OK;
    }
    if (n == 2 && p.front && (p.edge != p.sharpen)) {
        // the reference frame "compute chain -> edge or sharpen": front applied at load, one pass over HBM
        if (p.edge) return launch_stencil3<kOpSobel>(ctx, slot.stream, in, out, &p.lut, p.use_lut, p.gray, p.inv, 0);
        return launch_stencil3<kOpSharpen>(ctx, slot.stream, in, out, &p.lut, p.use_lut, p.gray, p.inv, (int)p.factor);
    }
    const size_t frame_bytes = align_up((size_t)in.w * 4, 256) * in.h;
    uint32_t chunk = in.frames;
    if (n > 1) {
        const size_t budget = ctx->chunk_bytes ? ctx->chunk_bytes : ((size_t)256 << 20);
        chunk = (uint32_t)std::max<size_t>(1, std::min<size_t>(in.frames, budget / frame_bytes));
        NORM_TRY(ensure_buf(ctx, slot.tmp[0], frame_bytes * chunk));
        if (n > 2) NORM_TRY(ensure_buf(ctx, slot.tmp[1], frame_bytes * chunk));
    }
    for (uint32_t f0 = 0; f0 < in.frames; f0 += chunk) {
        const uint32_t nf = std::min(chunk, in.frames - f0);
        ImgView cur = sub_frames(in, f0, nf);
        const ImgView dst_final = sub_frames(out, f0, nf);
        int done = 0, pp = 0;
        auto next_dst = [&]() -> ImgView {
            if (done == n - 1) return dst_final;
            ImgView t = view_of(slot.tmp[pp].p, in.w, in.h, nf);
            pp ^= 1;
            return t;
        };
        if (p.front) {
            ImgView d = next_dst();
            NORM_TRY(launch_front(ctx, slot.stream, cur, d, p.lut, p.use_lut, p.gray, p.inv));
            cur = d, ++done;
        }
        if (p.blur) {
            ImgView d = next_dst();
            NORM_TRY(launch_blur(ctx, slot.stream, cur, d, p.radius));
            cur = d, ++done;
        }
        if (p.sharpen) {
            ImgView d = next_dst();
            NORM_TRY(launch_sharpen(ctx, slot.stream, cur, d, p.factor));
            cur = d, ++done;
        }
        if (p.edge) {
            ImgView d = next_dst();
            NORM_TRY(launch_sobel(ctx, slot.stream, cur, d));
            cur = d, ++done;
        }
    }
    return NORM_OK;
}

// Fused single-pass chain kernel (chain_kernels.cuh) covers front -> blur -> Sobel (+SSIM)
// at the templated radii; every other combination runs the stage kernels back to back.
// GEN 1: first-generation kernel; GEN 2: pipelined kernel, MODE 0 (blur+Sobel); GEN 3: pipelined, MODE 1 (blur only)
template <int R, bool GRAY, bool FRONT, int GEN>
norm_status launch_fused(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in, const ImgView &out,
                         double *d_ssim, double *d_moments, uint32_t y0, uint32_t y1) {
    using C = typename std::conditional<GEN == 1, FusedCfg<R, GRAY>, Fused2Cfg<R, GRAY, (GEN == 3 ? 1 : 0)>>::type;
    void (*kern)(const FusedParams) = nullptr;
    if constexpr (GEN == 3) kern = chain_fused2_kernel<R, GRAY, FRONT, 1>;
    else if constexpr (GEN == 2) kern = chain_fused2_kernel<R, GRAY, FRONT, 0>;
    else kern = chain_fused_kernel<R, GRAY, FRONT>;
    constexpr int kThreads = GEN == 1 ? kFusedThreads : kC2Threads;
    if constexpr (GEN != 1) {
        // The pipelined kernel re-divides its registers between the warp groups (setmaxnreg) on the assumption that it
        // was launched with exactly C::REGS_LAUNCH (80; 168 for the one-CTA-per-SM radii) per thread: with fewer, the groups that ask for more would wait forever for
        // registers nobody releases.  Checked once per instantiation instead of trusted.
        static int regs = -1;  // (per template instantiation)
        if (regs < 0) {
            cudaFuncAttributes fa;
            CU_TRY(ctx, cudaFuncGetAttributes(&fa, kern));
            regs = fa.numRegs;
        }
        if (!C::BIG && regs != C::REGS_LAUNCH)  // (the one-CTA-per-SM configuration does not re-divide its registers)
            return fail(ctx, NORM_E_UNSUPPORTED, "fused kernel was built with " + std::to_string(regs) +
                        " registers per thread; its setmaxnreg budgets need exactly " + std::to_string(C::REGS_LAUNCH) +
                        " (check the nvcc flags)");
    }
    CU_TRY(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::kSmem));
    int per_sm = 1;
    CU_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, C::kSmem));
    if (per_sm < 1) per_sm = 1;
    FusedParams fp;
    std::memset(&fp, 0, sizeof(fp));
    fp.in = in, fp.out = out;
    if (make_taps(R, fp.taps) != NORM_OK) return fail(ctx, NORM_E_INVALID_ARG, "bad radius");
    fp.lut = p.lut;
    for (int k = 0; k <= R; ++k)
        fp.tp[k] = make_float2(k < R ? fp.taps.t[k] : 0.0f, k >= 1 ? fp.taps.t[k - 1] : 0.0f);
    fp.strips = (in.w + C::SW - 1) / C::SW;
    // output rows of this launch: all of them, or the slot's window (band mode)
    fp.out_y0 = slot.win_y1 ? slot.win_y0 : 0u;
    fp.out_y1 = slot.win_y1 ? std::min(slot.win_y1, in.h) : in.h;
    if (fp.out_y0 >= fp.out_y1) return fail(ctx, NORM_E_INVALID_ARG, "empty output row window");
    const uint32_t out_rows = fp.out_y1 - fp.out_y0;
    // Row bands per frame: the persistent grid takes ceil(items / slots) rounds of one item per CTA, and an item costs
    // its rows plus a fixed pipeline fill (R - 1 extra rows of H pass and about one batch).  Pick the band count
    // (bands no shorter than 64 rows) with the smallest rounds x item cost: one band when there are many frames, a
    // count that fills the last round when one giant image gives only a few hundred strips.
    const size_t slots = (size_t)ctx->sm_count * per_sm;
    const size_t base_items = (size_t)fp.strips * in.frames;
    uint32_t bands = 1;
    {
        const uint32_t max_bands = std::max<uint32_t>(1, std::min<uint32_t>(out_rows / 64, 1024));
        const uint64_t fill = (uint64_t)(R - 1) + C::RB / 2;  // measured: 13 rounds of 294-row items cost 1.7 % more than one of 4118 rows
        uint64_t best = ~0ull;
        for (uint32_t nbands = 1; nbands <= max_bands; ++nbands) {
            const uint64_t rows_per = (out_rows + nbands - 1) / nbands;
            const uint64_t rounds = (base_items * nbands + slots - 1) / slots;
            const uint64_t cost = rounds * (rows_per + fill);
            if (cost < best) best = cost, bands = nbands;
        }
    }
    if (ctx->force_bands) bands = std::min<uint32_t>(ctx->force_bands, out_rows);
    fp.band_rows = (out_rows + bands - 1) / bands;
    fp.bands = (out_rows + fp.band_rows - 1) / fp.band_rows;
    // One giant frame (pipelined kernel): when the strips do not fill the last round, cut the strips' rows, laid end
    // to end, into one contiguous range per CTA instead (a range that crosses a strip boundary costs one more pipeline
    // fill).  274 strips of 4118 rows on 296 CTAs: 4148 row-equivalents per CTA become 3872.
    size_t grid_items = (size_t)fp.strips * fp.bands * in.frames;
    if (GEN != 1 && in.frames == 1 && !ctx->force_bands && ctx->balanced_partition) {
        const uint64_t fill = (uint64_t)(R - 1) + C::RB / 2;
        const uint64_t total_rows = (uint64_t)fp.strips * out_rows;
        const uint64_t rounds = ((uint64_t)fp.strips * fp.bands + slots - 1) / slots;
        const uint64_t cost_static = rounds * (fp.band_rows + fill);
        const uint64_t per = std::max<uint64_t>(ctx->balanced_partition == 2 ? 7 : 64, (total_rows + slots - 1) / slots);
        const uint64_t pieces = (per + out_rows - 1) / out_rows + 1;  // items a range can be cut into
        const uint64_t cost_bal = per + std::min<uint64_t>(pieces, 2) * fill;
        if ((cost_bal < cost_static || ctx->balanced_partition == 2) && total_rows < (1ull << 31)) {
            fp.balanced = 1, fp.per = (uint32_t)per, fp.maxi = (uint32_t)pieces;
            fp.bands = 1, fp.band_rows = out_rows;
            grid_items = (size_t)((total_rows + per - 1) / per);
        }
    }
    fp.invert = p.inv ? 1u : 0u;
    fp.lut_identity = p.use_lut ? 0u : 1u;
    fp.do_ssim = (d_ssim || d_moments) ? 1u : 0u;
    fp.ssim_y0 = std::max(y0, fp.out_y0), fp.ssim_y1 = std::max(fp.ssim_y0, std::min(y1, fp.out_y1));
    if (grid_items >= (1ull << 31)) return fail(ctx, NORM_E_UNSUPPORTED, "more than 2^31 work items in one launch: split the batch");
    const unsigned grid = grid_for(ctx, grid_items, per_sm);
    // partial-sum slots per frame: one per work item (balanced: maxi per CTA, unused ones are written as zero)
    const uint32_t items_per_frame = fp.balanced ? grid * fp.maxi : fp.strips * fp.bands;
    if (fp.do_ssim) {
        NORM_TRY(ensure_buf(ctx, slot.partials, (size_t)in.frames * items_per_frame * 5 * sizeof(double)));
        fp.partials = static_cast<double *>(slot.partials.p);
    }
    kern<<<grid, kThreads, C::kSmem, slot.stream>>>(fp);
    LAUNCH_CHECK(ctx);
    if (fp.do_ssim) {
        const double npix = (double)in.w * (double)(fp.ssim_y1 - fp.ssim_y0);
        ssim_finalize_kernel<<<in.frames, 256, 0, slot.stream>>>(fp.partials, items_per_frame, npix, d_moments, d_ssim,
                                                                 slot.accumulate ? 1 : 0);
        LAUNCH_CHECK(ctx);
    }
    return NORM_OK;
}

// Largest radius of the pipelined kernel (development builds lower these to fall back to the first-generation kernels).
// Every radius the reference's UI reaches runs on it: the shift-form V pass keeps 3 x (R - 1) (chain on RGB) or
// 2 x 2 x (R - 1) (RGBA blur, groups B and C) running outputs in registers, with one CTA per SM from r = 33 / 37.
#ifndef NORM_GEN2_RGB_MAX_R
#define NORM_GEN2_RGB_MAX_R 39
#endif
#ifndef NORM_GEN2_RGBA_MAX_R
#define NORM_GEN2_RGBA_MAX_R 39
#endif

template <int R>
norm_status launch_fused_r(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in, const ImgView &out,
                           double *d_ssim, double *d_moments, uint32_t y0, uint32_t y1) {
    // second-generation kernel (column-owner consumer) where its register budget allows
    // (the first-generation kernel stays selectable, NORM_CUDA_FUSED_GEN=1, at the BASELINE radii: cross-check tests)
    if constexpr (R <= NORM_GEN2_RGB_MAX_R) {
        if (ctx->fused_gen != 1) {
            if (p.gray) return launch_fused<R, true, true, 2>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
            if (p.front) return launch_fused<R, false, true, 2>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
            return launch_fused<R, false, false, 2>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
        }
        if (p.gray) return launch_fused<R, true, true, 1>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    } else {
        if (p.gray) return launch_fused<R, true, true, 2>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    }
    if (p.front) return launch_fused<R, false, true, 1>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    return launch_fused<R, false, false, 1>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
}

// The other radii the reference's UI can reach (odd, 3..39 in steps of 2, feature_uniform.rs:66-84): pipelined
// kernel only (the first-generation kernel is not instantiated for them unless NORM_GEN2_RGB_MAX_R is lowered).
template <int R>
norm_status launch_fused_rx(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in, const ImgView &out,
                            double *d_ssim, double *d_moments, uint32_t y0, uint32_t y1, bool *fused) {
    if (p.gray) return launch_fused<R, true, true, 2>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    if constexpr (R <= NORM_GEN2_RGB_MAX_R) {
        if (ctx->fused_gen == 1) {
            *fused = false;
            return NORM_OK;
        }
        if (p.front) return launch_fused<R, false, true, 2>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
        return launch_fused<R, false, false, 2>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    } else {
        if (p.front) return launch_fused<R, false, true, 1>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
        return launch_fused<R, false, false, 1>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    }
}

#if defined(NORM_NO_EXTRA_RADII) && defined(NORM_DEV_RADIUS)  // development builds: the BASELINE radii + one more
#define NORM_EXTRA_RADII(X) X(NORM_DEV_RADIUS)
#elif defined(NORM_NO_EXTRA_RADII)  // development builds: only the BASELINE radii (3, 7, 21, 39); a third of the compile time
#define NORM_EXTRA_RADII(X)
#else
#define NORM_EXTRA_RADII(X) X(5) X(9) X(11) X(13) X(15) X(17) X(19) X(23) X(25) X(27) X(29) X(31) X(33) X(35) X(37)
#endif

// front -> blur with nothing after it (norm_cuda_blur, the blur effect of norm_cuda_effects, blur-only chains):
// pipelined kernel in MODE 1.  CH*R accumulators per thread: R <= 21 on RGBA (4 channels), any R after grayscale (2).
template <int R>
norm_status launch_blur_only_r(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in, const ImgView &out,
                               bool *fused) {
    *fused = true;
    if (p.gray) return launch_fused<R, true, true, 3>(ctx, slot, p, in, out, nullptr, nullptr, 0, in.h);
    if constexpr (R <= NORM_GEN2_RGBA_MAX_R) {
        if (p.front) return launch_fused<R, false, true, 3>(ctx, slot, p, in, out, nullptr, nullptr, 0, in.h);
        return launch_fused<R, false, false, 3>(ctx, slot, p, in, out, nullptr, nullptr, 0, in.h);
    }
    *fused = false;
    return NORM_OK;
}

// The blur alone at r = 3 on register windows (blur_win3_kernel).
template <bool GRAY, int FRONT>
norm_status launch_blur_win3(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in, const ImgView &out) {
    FusedParams fp;
    std::memset(&fp, 0, sizeof(fp));
    fp.in = in, fp.out = out;
    if (make_taps(3, fp.taps) != NORM_OK || fp.taps.n != 3 || fp.taps.dmin != -1) return fail(ctx, NORM_E_INVALID_ARG, "bad radius");
    fp.lut = p.lut;
    fp.out_y0 = slot.win_y1 ? slot.win_y0 : 0u;
    fp.out_y1 = slot.win_y1 ? std::min(slot.win_y1, in.h) : in.h;
    if (fp.out_y0 >= fp.out_y1) return fail(ctx, NORM_E_INVALID_ARG, "empty output row window");
    const uint32_t out_rows = fp.out_y1 - fp.out_y0;
    fp.invert = p.inv ? 1u : 0u;
    const uint32_t tiles_x = (in.w + kCwCols - 1) / kCwCols;
    const size_t want = (size_t)ctx->sm_count * 2 * kCwWarps * 2;
    uint32_t tile_rows = 72;  // 2 of 74 source rows are halo; halved down to 18 rows for small jobs
    while (tile_rows > 18 && (size_t)tiles_x * ((out_rows + tile_rows - 1) / tile_rows) * in.frames < want) tile_rows /= 2;
    const uint32_t tiles_y = (out_rows + tile_rows - 1) / tile_rows;
    const size_t tiles = (size_t)tiles_x * tiles_y * in.frames;
    if (tiles >= (1ull << 31)) return fail(ctx, NORM_E_UNSUPPORTED, "more than 2^31 work items in one launch: split the batch");
    const unsigned grid = grid_for(ctx, (tiles + kCwWarps - 1) / kCwWarps, 2);
    blur_win3_kernel<GRAY, FRONT><<<grid, kCwThreads, 0, slot.stream>>>(fp, tiles_x, tiles_y, tile_rows);
    LAUNCH_CHECK(ctx);
    return NORM_OK;
}

norm_status try_launch_blur_fused(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in, const ImgView &out,
                                  bool *fused) {
    *fused = false;
    if (!p.blur || p.edge || p.sharpen || ctx->force_unfused || ctx->fused_gen == 1) return NORM_OK;
    if (p.radius == 3 && ctx->win3) {
        *fused = true;
        // front stage: none / with the table / grayscale or inversion alone (the identity table is not looked up)
        if (p.gray) return p.use_lut ? launch_blur_win3<true, 1>(ctx, slot, p, in, out) : launch_blur_win3<true, 2>(ctx, slot, p, in, out);
        if (p.use_lut) return launch_blur_win3<false, 1>(ctx, slot, p, in, out);
        if (p.front) return launch_blur_win3<false, 2>(ctx, slot, p, in, out);
        return launch_blur_win3<false, 0>(ctx, slot, p, in, out);
    }
    switch (p.radius) {
    case 3: return launch_blur_only_r<3>(ctx, slot, p, in, out, fused);
    case 7: return launch_blur_only_r<7>(ctx, slot, p, in, out, fused);
    case 21: return launch_blur_only_r<21>(ctx, slot, p, in, out, fused);
    case 39: return launch_blur_only_r<39>(ctx, slot, p, in, out, fused);
#define X(R_) case R_: return launch_blur_only_r<R_>(ctx, slot, p, in, out, fused);
        NORM_EXTRA_RADII(X)
#undef X
    default: return NORM_OK;
    }
}

// The chain at r = 3 on register windows (chainw_kernels.cuh): one warp per tile of 120 columns x tile_rows rows.
template <bool GRAY, int FRONT>
norm_status launch_chain_win3(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in, const ImgView &out,
                              double *d_ssim, double *d_moments, uint32_t y0, uint32_t y1) {
    FusedParams fp;
    std::memset(&fp, 0, sizeof(fp));
    fp.in = in, fp.out = out;
    if (make_taps(3, fp.taps) != NORM_OK || fp.taps.n != 3 || fp.taps.dmin != -1) return fail(ctx, NORM_E_INVALID_ARG, "bad radius");
    fp.lut = p.lut;
    fp.out_y0 = slot.win_y1 ? slot.win_y0 : 0u;
    fp.out_y1 = slot.win_y1 ? std::min(slot.win_y1, in.h) : in.h;
    if (fp.out_y0 >= fp.out_y1) return fail(ctx, NORM_E_INVALID_ARG, "empty output row window");
    const uint32_t out_rows = fp.out_y1 - fp.out_y0;
    fp.invert = p.inv ? 1u : 0u;
    fp.do_ssim = (d_ssim || d_moments) ? 1u : 0u;
    fp.ssim_y0 = std::max(y0, fp.out_y0), fp.ssim_y1 = std::max(fp.ssim_y0, std::min(y1, fp.out_y1));
    // tile height: 72 rows (4 of 76 source rows are halo) when that still gives every warp of the grid two tiles,
    // halved down to 18 for small jobs
    const uint32_t tiles_x = (in.w + kCwCols - 1) / kCwCols;
    const size_t want = (size_t)ctx->sm_count * 2 * kCwWarps * 2;
    uint32_t tile_rows = 72;
    while (tile_rows > 18 && (size_t)tiles_x * ((out_rows + tile_rows - 1) / tile_rows) * in.frames < want) tile_rows /= 2;
    const uint32_t tiles_y = (out_rows + tile_rows - 1) / tile_rows;
    const size_t tiles = (size_t)tiles_x * tiles_y * in.frames;
    if (tiles >= (1ull << 31)) return fail(ctx, NORM_E_UNSUPPORTED, "more than 2^31 work items in one launch: split the batch");
    const uint32_t items_per_frame = tiles_x * tiles_y;
    if (fp.do_ssim) {
        NORM_TRY(ensure_buf(ctx, slot.partials, (size_t)in.frames * items_per_frame * 5 * sizeof(double)));
        fp.partials = static_cast<double *>(slot.partials.p);
    }
    const unsigned grid = grid_for(ctx, (tiles + kCwWarps - 1) / kCwWarps, 2);
    chain_win3_kernel<GRAY, FRONT><<<grid, kCwThreads, 0, slot.stream>>>(fp, tiles_x, tiles_y, tile_rows);
    LAUNCH_CHECK(ctx);
    if (fp.do_ssim) {
        const double npix = (double)in.w * (double)(fp.ssim_y1 - fp.ssim_y0);
        ssim_finalize_kernel<<<in.frames, 256, 0, slot.stream>>>(fp.partials, items_per_frame, npix, d_moments, d_ssim,
                                                                 slot.accumulate ? 1 : 0);
        LAUNCH_CHECK(ctx);
    }
    return NORM_OK;
}

norm_status try_launch_chain_fused(norm_ctx *ctx, Slot &slot, const StagePlan &p, const ImgView &in,
                                   const ImgView &out, double *d_ssim, double *d_moments, uint32_t y0, uint32_t y1,
                                   bool *fused) {
    *fused = false;
    if (!(p.blur && p.edge) || p.sharpen || ctx->force_unfused) return NORM_OK;
    *fused = true;
    if (p.radius == 3 && ctx->win3 && ctx->fused_gen != 1) {
        if (p.gray)
            return p.use_lut ? launch_chain_win3<true, 1>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1)
                             : launch_chain_win3<true, 2>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
        if (p.use_lut) return launch_chain_win3<false, 1>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
        if (p.front) return launch_chain_win3<false, 2>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
        return launch_chain_win3<false, 0>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    }
    switch (p.radius) {
    case 3: return launch_fused_r<3>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    case 7: return launch_fused_r<7>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    case 21: return launch_fused_r<21>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
    case 39: return launch_fused_r<39>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1);
#define X(R_) case R_: return launch_fused_rx<R_>(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1, fused);
        NORM_EXTRA_RADII(X)
#undef X
    default: *fused = false; return NORM_OK;
    }
}

// Rows of SOURCE needed above / below an output row by the chained stencil stages (blur, sharpen, Sobel).
norm_status stencil_reach(const NormFeatureUniform *f, uint32_t *up_out, uint32_t *down_out) {
    uint32_t up = 0, down = 0;
    if (f->blur == 1) {
        Taps t;
        if (make_taps(f->blur_radius, t) != NORM_OK) return NORM_E_INVALID_ARG;
        up += (uint32_t)(-t.dmin);
        down += (uint32_t)(t.dmin + t.n - 1);
    }
    if (f->sharpen == 1) up += 1, down += 1;
    if (f->edge_detect == 1) up += 1, down += 1;
    *up_out = up, *down_out = down;
    return NORM_OK;
}

// Row sub-view of every frame: rows [y0, y1).
ImgView sub_rows(const ImgView &v, uint32_t y0, uint32_t y1) {
    ImgView r = v;
    r.ptr = v.ptr + (size_t)y0 * v.pitch;
    r.h = y1 - y0;
    return r;
}

// NORM_OUT_SRGB: one table pass over the finished frame, in place (RGB through the encode table, alpha kept).
norm_status encode_srgb_inplace(norm_ctx *ctx, cudaStream_t s, const ImgView &v) {
    static const Lut256 enc = [] {
        Lut256 l;
        make_srgb_encode_lut(l);
        return l;
    }();
    return launch_front(ctx, s, v, v, enc, /*use_lut=*/true, /*gray=*/false, /*inv=*/false);
}

// Chain in -> out; SSIM (d_ssim) and/or raw moments (d_moments) over output rows [y0, y1) only.
norm_status run_chain_rows(norm_ctx *ctx, Slot &slot, const NormFeatureUniform *f, uint32_t flags, const ImgView &in,
                           const ImgView &out, double *d_ssim, double *d_moments, uint32_t y0, uint32_t y1) {
    const StagePlan p = make_plan(f, flags, /*chain=*/true);
    bool fused = false;
    NORM_TRY(try_launch_chain_fused(ctx, slot, p, in, out, d_ssim, d_moments, y0, y1, &fused));
    if (!fused) {
        NORM_TRY(run_stages_unfused(ctx, slot, p, in, out));
        if (d_ssim || d_moments)
            NORM_TRY(launch_ssim(ctx, slot, sub_rows(out, y0, std::min(y1, in.h)), sub_rows(in, y0, std::min(y1, in.h)),
                                 d_moments, d_ssim));
    }
    // the SSIM above is of the linear result; the display encode comes last
    if (flags & NORM_OUT_SRGB) NORM_TRY(encode_srgb_inplace(ctx, slot.stream, out));
    return NORM_OK;
}

norm_status run_chain(norm_ctx *ctx, Slot &slot, const NormFeatureUniform *f, uint32_t flags, const ImgView &in,
                      const ImgView &out, double *d_ssim) {
    return run_chain_rows(ctx, slot, f, flags, in, out, d_ssim, nullptr, 0, in.h);
}

norm_status ensure_results(norm_ctx *ctx, uint32_t frames) {
    const size_t need = (size_t)frames * 6 * sizeof(double);
    NORM_TRY(ensure_buf(ctx, ctx->results, need));
    if (ctx->h_results_cap < need) {
        if (ctx->h_results) cudaFreeHost(ctx->h_results);
        ctx->h_results = nullptr;
        ctx->h_results_cap = 0;
        CU_TRY(ctx, cudaMallocHost(&ctx->h_results, need));
        ctx->h_results_cap = need;
    }
    return NORM_OK;
}

norm_status check_pair(norm_ctx *ctx, const norm_img *in, const norm_img *out, bool allow_alias) {
    if (!ctx) return NORM_E_INVALID_ARG;
    if (!in || !out) return fail(ctx, NORM_E_INVALID_ARG, "image handle is NULL");
    if (!same_geometry(in->v, out->v)) return fail(ctx, NORM_E_DIM_MISMATCH, "images differ in width/height/frames");
    if (!allow_alias && overlaps(in->v, out->v)) return fail(ctx, NORM_E_INVALID_ARG, "in and out may not alias");
    return NORM_OK;
}

norm_status bind(norm_ctx *ctx) {
    if (!ctx) return NORM_E_INVALID_ARG;
    ctx->err.clear();
    CU_TRY(ctx, cudaSetDevice(ctx->device));
    return NORM_OK;
}

norm_status create_impl(int32_t device, void *ext_stream, bool use_ext, norm_ctx **out) {
    if (!out) return NORM_E_INVALID_ARG;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        (void)cudaGetLastError();
        g_create_error = std::string("no CUDA device available: ") + cudaGetErrorString(e) +
                         " (libnorm_cuda has no CPU fallback)";
        return NORM_E_CUDA;
    }
    if (device < 0 || device >= count) {
        g_create_error = "device index out of range";
        return NORM_E_INVALID_ARG;
    }
    norm_ctx *ctx = new (std::nothrow) norm_ctx();
    if (!ctx) return NORM_E_OOM;
    ctx->device = device;
    if (const char *e = std::getenv("NORM_CUDA_FORCE_UNFUSED")) ctx->force_unfused = e[0] == '1';
    if (const char *e = std::getenv("NORM_CUDA_FUSED_GEN")) ctx->fused_gen = std::atoi(e);
    if (const char *e = std::getenv("NORM_CUDA_BALANCED")) ctx->balanced_partition = std::atoi(e);
    if (const char *e = std::getenv("NORM_CUDA_WIN3")) ctx->win3 = std::atoi(e);
    if (const char *e = std::getenv("NORM_CUDA_BANDS")) ctx->force_bands = (uint32_t)std::atoi(e);
    if (const char *e = std::getenv("NORM_CUDA_CHUNK_BYTES")) ctx->chunk_bytes = (size_t)std::strtoull(e, nullptr, 10);
    auto bail = [&](cudaError_t ce, const char *what) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(ce);
        norm_cuda_destroy(ctx);
        return NORM_E_CUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail(e, "cudaSetDevice");
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail(e, "cudaGetDeviceProperties");
    ctx->sm_count = prop.multiProcessorCount;
    ctx->total_mem = prop.totalGlobalMem;
    std::snprintf(ctx->name, sizeof(ctx->name), "%s", prop.name);
    if (prop.major < 10) {
        g_create_error = std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major) +
                         std::to_string(prop.minor) + "; libnorm_cuda is built for sm_100a only";
        norm_cuda_destroy(ctx);
        return NORM_E_UNSUPPORTED;
    }
    if (use_ext) {
        ctx->main.stream = static_cast<cudaStream_t>(ext_stream);
        ctx->main.own_stream = false;
    } else {
        if ((e = cudaStreamCreateWithFlags(&ctx->main.stream, cudaStreamNonBlocking)) != cudaSuccess)
            return bail(e, "cudaStreamCreate");
        ctx->main.own_stream = true;
    }
    if ((e = cudaEventCreate(&ctx->ev0)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreate(&ctx->ev1)) != cudaSuccess) return bail(e, "cudaEventCreate");
    *out = ctx;
    return NORM_OK;
}

norm_status ensure_pipe(norm_ctx *ctx) {
    if (ctx->pipe_ready) return NORM_OK;
    for (auto &s : ctx->pipe) {
        CU_TRY(ctx, cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        s.own_stream = true;
    }
    ctx->pipe_ready = true;
    return NORM_OK;
}

void free_slot(Slot &s) {
    free_buf(s.tmp[0]);
    free_buf(s.tmp[1]);
    free_buf(s.partials);
    free_buf(s.stage_in);
    free_buf(s.stage_out);
    if (s.own_stream && s.stream) cudaStreamDestroy(s.stream);
    s.stream = nullptr;
}

}  // namespace

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------

extern "C" {

int32_t norm_cuda_abi_version(void) { return NORM_CUDA_ABI_VERSION; }

const char *norm_cuda_status_string(norm_status s) {
    switch (s) {
    case NORM_OK: return "ok";
    case NORM_E_INVALID_ARG: return "invalid argument";
    case NORM_E_UNSUPPORTED: return "unsupported input";
    case NORM_E_CUDA: return "CUDA error";
    case NORM_E_NCCL: return "NCCL error";
    case NORM_E_OOM: return "out of memory";
    case NORM_E_DIM_MISMATCH: return "dimension mismatch";
    default: return "unknown status";
    }
}

norm_status norm_cuda_create(int32_t device, norm_ctx **out) { return create_impl(device, nullptr, false, out); }

norm_status norm_cuda_create_on_stream(int32_t device, void *cuda_stream, norm_ctx **out) {
    return create_impl(device, cuda_stream, true, out);
}

void norm_cuda_destroy(norm_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    free_slot(ctx->main);
    for (auto &s : ctx->pipe) free_slot(s);
    free_buf(ctx->raw);
    free_buf(ctx->flip_tmp);
    free_buf(ctx->inflated);
    free_buf(ctx->passes);
    if (ctx->d_status) cudaFree(ctx->d_status);
    free_buf(ctx->results);
    if (ctx->h_results) cudaFreeHost(ctx->h_results);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    (void)cudaGetLastError();
    delete ctx;
}

const char *norm_cuda_last_error(const norm_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

norm_status norm_cuda_sync(norm_ctx *ctx) {
    NORM_TRY(bind(ctx));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    if (ctx->pipe_ready)
        for (auto &s : ctx->pipe) CU_TRY(ctx, cudaStreamSynchronize(s.stream));
    return NORM_OK;
}

uint64_t norm_cuda_launch_count(const norm_ctx *ctx) { return ctx ? ctx->launches : 0; }

norm_status norm_cuda_device_info(const norm_ctx *ctx, int32_t *sm_count, size_t *total_mem, char *name,
                                  size_t name_cap) {
    if (!ctx) return NORM_E_INVALID_ARG;
    if (sm_count) *sm_count = ctx->sm_count;
    if (total_mem) *total_mem = ctx->total_mem;
    if (name && name_cap) std::snprintf(name, name_cap, "%s", ctx->name);
    return NORM_OK;
}

norm_status norm_cuda_timer_start(norm_ctx *ctx) {
    NORM_TRY(bind(ctx));
    CU_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->main.stream));
    return NORM_OK;
}
norm_status norm_cuda_timer_stop(norm_ctx *ctx) {
    NORM_TRY(bind(ctx));
    CU_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->main.stream));
    return NORM_OK;
}
norm_status norm_cuda_timer_elapsed_ms(norm_ctx *ctx, float *ms) {
    NORM_TRY(bind(ctx));
    if (!ms) return fail(ctx, NORM_E_INVALID_ARG, "ms is NULL");
    CU_TRY(ctx, cudaEventSynchronize(ctx->ev1));
    CU_TRY(ctx, cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    return NORM_OK;
}

// ---- images ----------------------------------------------------------------------------

norm_status norm_cuda_image_create(norm_ctx *ctx, uint32_t width, uint32_t height, uint32_t frames, norm_img **out) {
    NORM_TRY(bind(ctx));
    if (!out) return fail(ctx, NORM_E_INVALID_ARG, "out is NULL");
    *out = nullptr;
    if (width == 0 || height == 0 || frames == 0) return fail(ctx, NORM_E_INVALID_ARG, "zero-sized image");
    norm_img *img = new (std::nothrow) norm_img();
    if (!img) return NORM_E_OOM;
    img->v = view_of(nullptr, width, height, frames);
    img->owned = true;
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, img->v.frame_stride * frames);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        delete img;
        return fail(ctx, e == cudaErrorMemoryAllocation ? NORM_E_OOM : NORM_E_CUDA,
                    std::string("cudaMalloc: ") + cudaGetErrorString(e));
    }
    img->v.ptr = static_cast<uint8_t *>(p);
    *out = img;
    return NORM_OK;
}

norm_status norm_cuda_image_wrap(norm_ctx *ctx, void *device_ptr, uint32_t width, uint32_t height, uint32_t frames,
                                 size_t pitch, size_t frame_stride, norm_img **out) {
    NORM_TRY(bind(ctx));
    if (!out) return fail(ctx, NORM_E_INVALID_ARG, "out is NULL");
    *out = nullptr;
    if (!device_ptr || width == 0 || height == 0 || frames == 0) return fail(ctx, NORM_E_INVALID_ARG, "bad image");
    if (pitch % 16 || frame_stride % 16 || (uintptr_t)device_ptr % 16 || pitch < (size_t)width * 4 ||
        frame_stride < pitch * height)
        return fail(ctx, NORM_E_INVALID_ARG, "pitch / frame_stride / pointer must be 16-byte aligned and large enough");
    norm_img *img = new (std::nothrow) norm_img();
    if (!img) return NORM_E_OOM;
    img->v.ptr = static_cast<uint8_t *>(device_ptr);
    img->v.pitch = pitch;
    img->v.frame_stride = frame_stride;
    img->v.w = width, img->v.h = height, img->v.frames = frames;
    img->owned = false;
    *out = img;
    return NORM_OK;
}

void norm_cuda_image_destroy(norm_ctx *ctx, norm_img *img) {
    if (!img) return;
    if (img->owned && img->v.ptr) {
        if (ctx) cudaSetDevice(ctx->device);
        cudaFree(img->v.ptr);
    }
    delete img;
}

norm_status norm_cuda_image_info(const norm_img *img, uint32_t *width, uint32_t *height, uint32_t *frames,
                                 size_t *pitch, size_t *frame_stride, void **device_ptr) {
    if (!img) return NORM_E_INVALID_ARG;
    if (width) *width = img->v.w;
    if (height) *height = img->v.h;
    if (frames) *frames = img->v.frames;
    if (pitch) *pitch = img->v.pitch;
    if (frame_stride) *frame_stride = img->v.frame_stride;
    if (device_ptr) *device_ptr = img->v.ptr;
    return NORM_OK;
}

static norm_status copy_frames(norm_ctx *ctx, cudaStream_t s, const ImgView &v, uint32_t frame0, uint32_t nframes,
                               uint8_t *host, size_t host_pitch, bool to_device) {
    if (frame0 + (uint64_t)nframes > v.frames) return fail(ctx, NORM_E_INVALID_ARG, "frame range outside image");
    if (host_pitch == 0) host_pitch = (size_t)v.w * 4;
    if (host_pitch < (size_t)v.w * 4) return fail(ctx, NORM_E_INVALID_ARG, "host_pitch < 4*width");
    uint8_t *dev = v.ptr + (size_t)frame0 * v.frame_stride;
    const bool one_shot = v.frame_stride == v.pitch * v.h;  // rows of consecutive frames are equidistant
    const uint32_t groups = one_shot ? 1 : nframes;
    const size_t rows = one_shot ? (size_t)v.h * nframes : v.h;
    for (uint32_t g = 0; g < groups; ++g) {
        uint8_t *d = dev + (size_t)g * v.frame_stride;
        uint8_t *h = host + (size_t)g * host_pitch * v.h;
        if (to_device)
            CU_TRY(ctx, cudaMemcpy2DAsync(d, v.pitch, h, host_pitch, (size_t)v.w * 4, rows, cudaMemcpyHostToDevice, s));
        else
            CU_TRY(ctx, cudaMemcpy2DAsync(h, host_pitch, d, v.pitch, (size_t)v.w * 4, rows, cudaMemcpyDeviceToHost, s));
    }
    return NORM_OK;
}

norm_status norm_cuda_upload(norm_ctx *ctx, norm_img *img, uint32_t frame0, uint32_t nframes, const uint8_t *host,
                             size_t host_pitch) {
    NORM_TRY(bind(ctx));
    if (!img || !host) return fail(ctx, NORM_E_INVALID_ARG, "NULL argument");
    NORM_TRY(copy_frames(ctx, ctx->main.stream, img->v, frame0, nframes, const_cast<uint8_t *>(host), host_pitch, true));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_download(norm_ctx *ctx, const norm_img *img, uint32_t frame0, uint32_t nframes, uint8_t *host,
                               size_t host_pitch) {
    NORM_TRY(bind(ctx));
    if (!img || !host) return fail(ctx, NORM_E_INVALID_ARG, "NULL argument");
    NORM_TRY(copy_frames(ctx, ctx->main.stream, img->v, frame0, nframes, host, host_pitch, false));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_host_alloc(norm_ctx *ctx, size_t bytes, void **out) {
    NORM_TRY(bind(ctx));
    if (!out || bytes == 0) return fail(ctx, NORM_E_INVALID_ARG, "bad host_alloc request");
    CU_TRY(ctx, cudaMallocHost(out, bytes));
    return NORM_OK;
}

void norm_cuda_host_free(norm_ctx *ctx, void *p) {
    if (!p) return;
    if (ctx) cudaSetDevice(ctx->device);
    cudaFreeHost(p);
}

// ---- expansion -------------------------------------------------------------------------

static int channels_of(int ct) {
    switch (ct) {
    case 0: return 1;
    case 2: return 3;
    case 3: return 1;
    case 4: return 2;
    case 6: return 4;
    default: return 0;
    }
}

static bool valid_depth(int ct, int bd) {
    switch (ct) {
    case 0: return bd == 1 || bd == 2 || bd == 4 || bd == 8 || bd == 16;
    case 3: return bd == 1 || bd == 2 || bd == 4 || bd == 8;
    case 2:
    case 4:
    case 6: return bd == 8 || bd == 16;
    default: return false;
    }
}

extern "C++" {
template <int CT, int BD, int MODE>
static void launch_expand(const ExpandParams &p, const ImgView &out, unsigned grid, cudaStream_t s) {
    expand_kernel<CT, BD, MODE><<<grid, 256, 0, s>>>(p, out);
}
}

// `nframes` images of one geometry -> frames [frame, frame + nframes) of `out`, ONE launch.  Host pixels (one image)
// are uploaded to a 16-byte row pitch; device pixels (the unfilter's output) already have one.
static norm_status expand_impl(norm_ctx *ctx, const NormPngDesc *png, int32_t mode, const ImgView &out, uint32_t frame,
                               bool pixels_on_device, size_t device_row_stride = 0, uint32_t nframes = 1,
                               size_t device_frame_stride = 0) {
    if (!png || !png->pixels) return fail(ctx, NORM_E_INVALID_ARG, "png descriptor / pixels is NULL");
    const int ct = png->color_type, bd = png->bit_depth;
    if (!channels_of(ct) || !valid_depth(ct, bd)) return fail(ctx, NORM_E_INVALID_ARG, "invalid colour type / bit depth");
    if (png->width != out.w || png->height != out.h) return fail(ctx, NORM_E_DIM_MISMATCH, "image size differs from png");
    if (nframes == 0 || (uint64_t)frame + nframes > out.frames) return fail(ctx, NORM_E_INVALID_ARG, "frame out of range");
    if (mode != NORM_EXPAND_COMPAT && mode != NORM_EXPAND_SPEC) return fail(ctx, NORM_E_INVALID_ARG, "unknown mode");
    if (mode == NORM_EXPAND_COMPAT && (bd != 8 || ct == 3))
        return fail(ctx, NORM_E_UNSUPPORTED,
                    "compat mode mirrors Png::rgba8(): 8-bit grayscale / RGB / gray+alpha / RGBA only "
                    "(grammar.rs:123-155)");
    if (png->n_plte < 0 || png->n_plte > 256 || png->n_trns < 0 || png->n_trns > 256)
        return fail(ctx, NORM_E_INVALID_ARG, "PLTE / tRNS too long");
    if (ct == 3 && mode == NORM_EXPAND_SPEC && (!png->plte || png->n_plte == 0))
        return fail(ctx, NORM_E_INVALID_ARG, "palette image without PLTE");
    const size_t rowb = ((size_t)png->width * channels_of(ct) * bd + 7) / 8;
    if (png->len < rowb * png->height) return fail(ctx, NORM_E_INVALID_ARG, "pixel buffer shorter than height*row_bytes");

    ExpandParams p;
    std::memset(&p, 0, sizeof(p));
    if (!pixels_on_device) {
        // rows re-pitched to 16 bytes on the way up, so every load of the kernel is aligned; 64 bytes of slack: the
        // last item of a ragged row may read a few bytes past the row
        const size_t pitch = align_up(rowb, 16);
        NORM_TRY(ensure_buf(ctx, ctx->raw, align_up(pitch * png->height + 64, 256)));
        CU_TRY(ctx, cudaMemcpy2DAsync(ctx->raw.p, pitch, png->pixels, rowb, rowb, png->height, cudaMemcpyHostToDevice,
                                      ctx->main.stream));
        p.src = static_cast<const uint8_t *>(ctx->raw.p), p.src_pitch = pitch, p.src_frame_stride = 0;
    } else {
        if (device_row_stride % 16 != 0 || ((uintptr_t)png->pixels & 15u) != 0)
            return fail(ctx, NORM_E_INVALID_ARG, "device scanlines must have a 16-byte aligned base and pitch");
        p.src = png->pixels, p.src_pitch = device_row_stride, p.src_frame_stride = device_frame_stride;
    }
    p.w = png->width, p.h = png->height, p.frames = nframes, p.frame0 = frame;
    if (png->plte && png->n_plte > 0) {
        std::memcpy(p.plte, png->plte, (size_t)png->n_plte * 3);
        p.n_plte = png->n_plte;
    }
    if (png->trns && png->n_trns > 0) {
        std::memcpy(p.trns, png->trns, (size_t)png->n_trns);
        p.n_trns = png->n_trns;
    }
    if (ct != 3) p.n_trns = 0;  // per-index alpha only exists for palette images
    const uint32_t smask = bd == 16 ? 0xFFFFu : ((1u << bd) - 1u);
    if (mode == NORM_EXPAND_SPEC && png->trns) {
        if (ct == 0 && png->n_trns >= 2) {
            p.key[0] = (((uint32_t)png->trns[0] << 8) | png->trns[1]) & smask;
            p.has_key = 1;
        } else if (ct == 2 && png->n_trns >= 6) {
            for (int c = 0; c < 3; ++c) p.key[c] = (((uint32_t)png->trns[2 * c] << 8) | png->trns[2 * c + 1]) & smask;
            p.has_key = 1;
        }
    }
    const size_t items = (size_t)((png->width + 3) / 4) * png->height * nframes;
    const unsigned grid = grid_for(ctx, (items + 511) / 512, 8);  // two items per thread and round
    cudaStream_t s = ctx->main.stream;
    // one instantiation per {colour type, bit depth, mode}: compat = 8-bit ct 0/2/4/6; spec = everything PngSuite has
    bool launched = true;
    if (mode == NORM_EXPAND_COMPAT) {
        switch (ct) {
        case 0: launch_expand<0, 8, 0>(p, out, grid, s); break;
        case 2: launch_expand<2, 8, 0>(p, out, grid, s); break;
        case 4: launch_expand<4, 8, 0>(p, out, grid, s); break;
        case 6: launch_expand<6, 8, 0>(p, out, grid, s); break;
        default: launched = false;
        }
    } else {
        switch (ct * 100 + bd) {
#define NORM_EXPAND_CASE(CT_, BD_) case CT_ * 100 + BD_: launch_expand<CT_, BD_, 1>(p, out, grid, s); break;
            NORM_EXPAND_CASE(0, 1) NORM_EXPAND_CASE(0, 2) NORM_EXPAND_CASE(0, 4) NORM_EXPAND_CASE(0, 8) NORM_EXPAND_CASE(0, 16)
            NORM_EXPAND_CASE(2, 8) NORM_EXPAND_CASE(2, 16)
            NORM_EXPAND_CASE(3, 1) NORM_EXPAND_CASE(3, 2) NORM_EXPAND_CASE(3, 4) NORM_EXPAND_CASE(3, 8)
            NORM_EXPAND_CASE(4, 8) NORM_EXPAND_CASE(4, 16) NORM_EXPAND_CASE(6, 8) NORM_EXPAND_CASE(6, 16)
#undef NORM_EXPAND_CASE
        default: launched = false;
        }
    }
    if (!launched) return fail(ctx, NORM_E_INVALID_ARG, "invalid colour type / bit depth");
    LAUNCH_CHECK(ctx);
    return NORM_OK;
}

static norm_status expand_to(norm_ctx *ctx, const NormPngDesc *png, int32_t mode, const ImgView &out, uint32_t frame) {
    return expand_impl(ctx, png, mode, out, frame, false);
}

norm_status norm_cuda_expand_rgba8(norm_ctx *ctx, const NormPngDesc *png, int32_t mode, norm_img *out, uint32_t frame) {
    NORM_TRY(bind(ctx));
    if (!out) return fail(ctx, NORM_E_INVALID_ARG, "out is NULL");
    NORM_TRY(expand_to(ctx, png, mode, out->v, frame));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));  // png->pixels may be pageable: do not retain it
    return NORM_OK;
}

norm_status norm_cuda_expand_rgba8_device(norm_ctx *ctx, const NormPngDesc *png, size_t row_pitch, size_t frame_stride,
                                          uint32_t frames, int32_t mode, norm_img *out, uint32_t frame0) {
    NORM_TRY(bind(ctx));
    if (!out) return fail(ctx, NORM_E_INVALID_ARG, "out is NULL");
    if (!png) return fail(ctx, NORM_E_INVALID_ARG, "png descriptor is NULL");
    const size_t rowb = ((size_t)png->width * channels_of(png->color_type) * png->bit_depth + 7) / 8;
    if (row_pitch < rowb) return fail(ctx, NORM_E_INVALID_ARG, "row pitch shorter than a scanline");
    if (frames > 1 && frame_stride < row_pitch * png->height) return fail(ctx, NORM_E_INVALID_ARG, "frame stride shorter than an image");
    return expand_impl(ctx, png, mode, out->v, frame0, /*pixels_on_device=*/true, row_pitch, frames, frame_stride);
}

norm_status norm_cuda_expand_rgba8_host(norm_ctx *ctx, const NormPngDesc *png, int32_t mode, uint8_t *out_rgba8) {
    NORM_TRY(bind(ctx));
    if (!png || !out_rgba8) return fail(ctx, NORM_E_INVALID_ARG, "NULL argument");
    if (png->width == 0 || png->height == 0) return fail(ctx, NORM_E_INVALID_ARG, "zero-sized image");
    const size_t bytes = align_up((size_t)png->width * 4, 256) * png->height;
    NORM_TRY(ensure_buf(ctx, ctx->main.tmp[0], bytes));
    const ImgView v = view_of(ctx->main.tmp[0].p, png->width, png->height, 1);
    NORM_TRY(expand_to(ctx, png, mode, v, 0));
    NORM_TRY(copy_frames(ctx, ctx->main.stream, v, 0, 1, out_rgba8, 0, false));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

// ---- unfilter -------------------------------------------------------------

static void adam7_sizes(uint32_t w, uint32_t h, uint32_t pw[7], uint32_t ph[7]) {
    static const uint32_t x0[7] = {0, 4, 0, 2, 0, 1, 0}, y0[7] = {0, 0, 4, 0, 2, 0, 1};
    static const uint32_t dx[7] = {8, 8, 4, 4, 2, 2, 1}, dy[7] = {8, 8, 8, 4, 4, 2, 2};
    for (int p = 0; p < 7; ++p) {
        pw[p] = w > x0[p] ? (w - x0[p] + dx[p] - 1) / dx[p] : 0;
        ph[p] = h > y0[p] ? (h - y0[p] + dy[p] - 1) / dy[p] : 0;
    }
}

// Uploads n inflated streams of identical geometry and undoes the scanline filters (and Adam7) on
// the GPU: one CTA per (image, pass) runs the wavefront, so a batch fills the machine where a single
// image occupies one SM.  The raster pixel_buffers are left in device memory: *dev_out points at
// image 0, *pitch_out is the row stride (row_bytes for interlaced images, a 16-byte-aligned pitch
// otherwise) and image i starts *img_stride_out bytes after image i-1.
static norm_status unfilter_to_device(norm_ctx *ctx, const NormInflatedDesc *descs, uint32_t n, const uint8_t **dev_out,
                                      size_t *pitch_out, size_t *row_bytes_out, size_t *img_stride_out) {
    if (!descs || n == 0) return fail(ctx, NORM_E_INVALID_ARG, "inflated descriptor is NULL / empty batch");
    const NormInflatedDesc *d = &descs[0];
    const int ct = d->color_type, bd = d->bit_depth;
    if (!channels_of(ct) || !valid_depth(ct, bd)) return fail(ctx, NORM_E_INVALID_ARG, "invalid colour type / bit depth");
    if (d->width == 0 || d->height == 0 || d->interlace > 1) return fail(ctx, NORM_E_INVALID_ARG, "bad dimensions / interlace");
    const uint32_t w = d->width, h = d->height, bits = (uint32_t)channels_of(ct) * bd;
    auto rowb = [&](uint32_t ww) { return ((size_t)ww * bits + 7) / 8; };
    const uint32_t bpp = std::max(1u, bits / 8);
    uint32_t pw[7] = {w, 0, 0, 0, 0, 0, 0}, ph[7] = {h, 0, 0, 0, 0, 0, 0};
    int npass = 1;
    if (d->interlace) adam7_sizes(w, h, pw, ph), npass = 7;
    size_t need = 0, pass_bytes = 0, ft_bytes = 0;
    for (int p = 0; p < npass; ++p)
        if (pw[p] && ph[p]) {
            need += (size_t)ph[p] * (1 + rowb(pw[p]));
            pass_bytes += (size_t)ph[p] * align_up(rowb(pw[p]), 16);
            ft_bytes += align_up(ph[p], 16);
        }
    for (uint32_t i = 0; i < n; ++i) {
        const NormInflatedDesc &e = descs[i];
        if (!e.data) return fail(ctx, NORM_E_INVALID_ARG, "inflated data is NULL");
        if (e.width != w || e.height != h || e.bit_depth != d->bit_depth || e.color_type != d->color_type ||
            e.interlace != d->interlace)
            return fail(ctx, NORM_E_INVALID_ARG, "images of a batch must share width, height, bit depth, colour type and interlace");
        if (e.len != need) return fail(ctx, NORM_E_INVALID_ARG, "inflated length does not match the image geometry");
    }
    const size_t raster_pitch = align_up(rowb(w), 16);   // de-interlaced rows at a 16-byte pitch, + slack for ragged reads
    const size_t out_bytes = raster_pitch * h + 64;
    const size_t src_stride = align_up(need, 256), dst_stride = align_up(pass_bytes, 256), ft_stride = align_up(ft_bytes, 256),
                 raw_stride = align_up(out_bytes, 256);
    NORM_TRY(ensure_buf(ctx, ctx->inflated, src_stride * n));
    NORM_TRY(ensure_buf(ctx, ctx->passes, (dst_stride + ft_stride) * n + 64));  // + slack: expansion may read a few bytes past a ragged last row
    if (d->interlace) NORM_TRY(ensure_buf(ctx, ctx->raw, raw_stride * n));
    if (!ctx->d_status) CU_TRY(ctx, cudaMalloc(&ctx->d_status, sizeof(int32_t)));
    cudaStream_t s = ctx->main.stream;
    CU_TRY(ctx, cudaMemsetAsync(ctx->d_status, 0, sizeof(int32_t), s));
    for (uint32_t i = 0; i < n; ++i)
        CU_TRY(ctx, cudaMemcpyAsync(static_cast<uint8_t *>(ctx->inflated.p) + i * src_stride, descs[i].data, need,
                                    cudaMemcpyHostToDevice, s));

    UnfilterParams up;
    std::memset(&up, 0, sizeof(up));
    up.status = ctx->d_status;
    up.nimg = n, up.src_stride = src_stride, up.dst_stride = dst_stride, up.ft_stride = ft_stride;
    DeinterlaceParams dp;
    std::memset(&dp, 0, sizeof(dp));
    uint8_t *ft_base = static_cast<uint8_t *>(ctx->passes.p) + dst_stride * n;
    size_t soff = 0, poff = 0, foff = 0;
    uint32_t max_items = 1;
    for (int p = 0; p < npass; ++p) {
        if (!pw[p] || !ph[p]) continue;
        UnfilterPass &ps = up.pass[up.npass++];
        ps.src = static_cast<const uint8_t *>(ctx->inflated.p) + soff;
        ps.dst = static_cast<uint8_t *>(ctx->passes.p) + poff;
        ps.ftype = ft_base + foff;
        ps.rows = ph[p], ps.row_bytes = (uint32_t)rowb(pw[p]), ps.pitch = (uint32_t)align_up(rowb(pw[p]), 16), ps.bpp = bpp;
        dp.pass_px[p] = ps.dst, dp.pass_row_bytes[p] = ps.pitch;
        soff += (size_t)ph[p] * (1 + rowb(pw[p]));
        poff += (size_t)ph[p] * ps.pitch;
        foff += align_up(ph[p], 16);
        max_items = std::max<uint32_t>(max_items, (uint32_t)(((size_t)ps.pitch / 4 * ps.rows + 255) / 256));
    }
    // grid.y / grid.z are limited to 65535; larger batches go in slices
    for (uint32_t i0 = 0; i0 < n; i0 += 32768) {
        const uint32_t cnt = std::min<uint32_t>(32768, n - i0);
        UnfilterParams sl = up;
        sl.nimg = cnt;
        for (int p = 0; p < sl.npass; ++p)
            sl.pass[p].src += i0 * src_stride, sl.pass[p].dst += i0 * dst_stride, sl.pass[p].ftype += i0 * ft_stride;
        const uint32_t gx = std::max(1u, std::min<uint32_t>(max_items, (uint32_t)ctx->sm_count * 8 / std::min<uint32_t>(cnt, 64u)));
        unfilter_align_kernel<<<dim3(gx, sl.npass, cnt), 256, 0, s>>>(sl);
        LAUNCH_CHECK(ctx);
        const dim3 wg(sl.npass, cnt);
        switch (bpp) {
        case 1: unfilter_wave_kernel<1><<<wg, kUnfThreads, 0, s>>>(sl); break;
        case 2: unfilter_wave_kernel<2><<<wg, kUnfThreads, 0, s>>>(sl); break;
        case 3: unfilter_wave_kernel<3><<<wg, kUnfThreads, 0, s>>>(sl); break;
        case 4: unfilter_wave_kernel<4><<<wg, kUnfThreads, 0, s>>>(sl); break;
        case 6: unfilter_wave_kernel<6><<<wg, kUnfThreads, 0, s>>>(sl); break;
        default: unfilter_wave_kernel<8><<<wg, kUnfThreads, 0, s>>>(sl); break;
        }
        LAUNCH_CHECK(ctx);
        if (d->interlace) {
            DeinterlaceParams dq = dp;
            for (int p = 0; p < 7; ++p) {
                if (!dq.pass_px[p]) dq.pass_px[p] = static_cast<const uint8_t *>(ctx->passes.p);  // empty pass: never read
                else dq.pass_px[p] += i0 * dst_stride;
            }
            dq.dst = static_cast<uint8_t *>(ctx->raw.p) + i0 * raw_stride;
            dq.w = w, dq.h = h, dq.row_bytes = (uint32_t)rowb(w), dq.bits = bits, dq.dst_pitch = (uint32_t)raster_pitch;
            dq.pass_stride = dst_stride, dq.dst_stride = raw_stride;
            const unsigned gdx = std::max(1u, grid_for(ctx, (out_bytes + 255) / 256, 8) / std::min<uint32_t>(cnt, 64u));
            deinterlace_kernel<<<dim3(gdx, cnt), 256, 0, s>>>(dq);
            LAUNCH_CHECK(ctx);
        }
    }
    if (d->interlace) {
        *dev_out = static_cast<const uint8_t *>(ctx->raw.p), *pitch_out = raster_pitch, *img_stride_out = raw_stride;
    } else {
        *dev_out = up.pass[0].dst, *pitch_out = up.pass[0].pitch, *img_stride_out = dst_stride;
    }
    int32_t st = 0;
    CU_TRY(ctx, cudaMemcpyAsync(&st, ctx->d_status, sizeof(st), cudaMemcpyDeviceToHost, s));
    CU_TRY(ctx, cudaStreamSynchronize(s));
    if (st) return fail(ctx, NORM_E_INVALID_ARG, "Unrecognized filter method in the scanline stream");  // grammar.rs:48-62
    *row_bytes_out = rowb(w);
    return NORM_OK;
}

norm_status norm_cuda_unfilter(norm_ctx *ctx, const NormInflatedDesc *d, uint8_t *pixel_buffer) {
    NORM_TRY(bind(ctx));
    if (!pixel_buffer) return fail(ctx, NORM_E_INVALID_ARG, "pixel_buffer is NULL");
    size_t rb = 0, pitch = 0, stride = 0;
    const uint8_t *dev = nullptr;
    NORM_TRY(unfilter_to_device(ctx, d, 1, &dev, &pitch, &rb, &stride));
    CU_TRY(ctx, cudaMemcpy2DAsync(pixel_buffer, rb, dev, pitch, rb, d->height, cudaMemcpyDeviceToHost, ctx->main.stream));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_unfilter_expand_rgba8_batch(norm_ctx *ctx, const NormInflatedDesc *descs, uint32_t n, const uint8_t *plte,
                                                  int32_t n_plte, const uint8_t *trns, int32_t n_trns, int32_t mode,
                                                  norm_img *out, uint32_t frame0) {
    NORM_TRY(bind(ctx));
    if (!out) return fail(ctx, NORM_E_INVALID_ARG, "out is NULL");
    if (n == 0 || (uint64_t)frame0 + n > out->v.frames) return fail(ctx, NORM_E_INVALID_ARG, "frame range outside the image");
    size_t rb = 0, pitch = 0, stride = 0;
    const uint8_t *dev = nullptr;
    NORM_TRY(unfilter_to_device(ctx, descs, n, &dev, &pitch, &rb, &stride));
    NormPngDesc png;
    std::memset(&png, 0, sizeof(png));
    png.width = descs->width, png.height = descs->height, png.bit_depth = descs->bit_depth, png.color_type = descs->color_type;
    png.len = pitch * descs->height;
    png.plte = plte, png.n_plte = n_plte, png.trns = trns, png.n_trns = n_trns;
    png.pixels = dev;  // the n unfiltered images sit `stride` bytes apart: one launch expands them all
    NORM_TRY(expand_impl(ctx, &png, mode, out->v, frame0, /*pixels_on_device=*/true, pitch, n, stride));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_unfilter_expand_rgba8(norm_ctx *ctx, const NormInflatedDesc *d, const uint8_t *plte, int32_t n_plte,
                                            const uint8_t *trns, int32_t n_trns, int32_t mode, norm_img *out, uint32_t frame) {
    return norm_cuda_unfilter_expand_rgba8_batch(ctx, d, d ? 1u : 0u, plte, n_plte, trns, n_trns, mode, out, frame);
}

// ---- encoder-side scanline filtering ---------------------------------------

static norm_status filter_device_rows(norm_ctx *ctx, const uint8_t *dev_rows, size_t pitch, uint32_t rows,
                                      uint32_t row_bytes, uint32_t bpp, int32_t paeth_compat, uint8_t *stream,
                                      uint8_t *filter_types) {
    const size_t out_bytes = (size_t)rows * ((size_t)row_bytes + 1);
    NORM_TRY(ensure_buf(ctx, ctx->inflated, align_up(out_bytes, 256) + align_up(rows, 256)));
    FilterParams fp;
    fp.src = dev_rows, fp.src_pitch = pitch;
    fp.dst = static_cast<uint8_t *>(ctx->inflated.p);
    fp.ftypes = fp.dst + align_up(out_bytes, 256);
    fp.rows = rows, fp.row_bytes = row_bytes, fp.bpp = bpp, fp.paeth_compat = paeth_compat;
    cudaStream_t s = ctx->main.stream;
    filter_rows_kernel<<<grid_for(ctx, (rows + 7) / 8, 8), 256, 0, s>>>(fp);
    LAUNCH_CHECK(ctx);
    CU_TRY(ctx, cudaMemcpyAsync(stream, fp.dst, out_bytes, cudaMemcpyDeviceToHost, s));
    if (filter_types) CU_TRY(ctx, cudaMemcpyAsync(filter_types, fp.ftypes, rows, cudaMemcpyDeviceToHost, s));
    CU_TRY(ctx, cudaStreamSynchronize(s));
    return NORM_OK;
}

norm_status norm_cuda_filter_scanlines(norm_ctx *ctx, uint32_t width, uint32_t height, uint8_t bit_depth,
                                       uint8_t color_type, const uint8_t *pixel_buffer, int32_t paeth_compat,
                                       uint8_t *stream, uint8_t *filter_types) {
    NORM_TRY(bind(ctx));
    if (!pixel_buffer || !stream) return fail(ctx, NORM_E_INVALID_ARG, "NULL argument");
    if (!channels_of(color_type) || !valid_depth(color_type, bit_depth) || width == 0 || height == 0)
        return fail(ctx, NORM_E_INVALID_ARG, "invalid image header");
    const uint32_t bits = (uint32_t)channels_of(color_type) * bit_depth;
    const size_t rowb = ((size_t)width * bits + 7) / 8;
    NORM_TRY(ensure_buf(ctx, ctx->raw, align_up(rowb * height, 256)));
    CU_TRY(ctx, cudaMemcpyAsync(ctx->raw.p, pixel_buffer, rowb * height, cudaMemcpyHostToDevice, ctx->main.stream));
    return filter_device_rows(ctx, static_cast<const uint8_t *>(ctx->raw.p), rowb, height, (uint32_t)rowb,
                              std::max(1u, bits / 8), paeth_compat, stream, filter_types);
}

norm_status norm_cuda_filter_image(norm_ctx *ctx, const norm_img *img, uint32_t frame, int32_t paeth_compat,
                                   uint8_t *stream, uint8_t *filter_types) {
    NORM_TRY(bind(ctx));
    if (!img || !stream) return fail(ctx, NORM_E_INVALID_ARG, "NULL argument");
    if (frame >= img->v.frames) return fail(ctx, NORM_E_INVALID_ARG, "frame out of range");
    return filter_device_rows(ctx, img->v.ptr + (size_t)frame * img->v.frame_stride, img->v.pitch, img->v.h,
                              img->v.w * 4, 4, paeth_compat, stream, filter_types);
}

// ---- effects ---------------------------------------------------------------------------

norm_status norm_cuda_effects(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags, const norm_img *in,
                              norm_img *out) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    NORM_TRY(check_uniform(ctx, f, in->v));
    const StagePlan p = make_plan(f, flags, /*chain=*/false);
    // FlipX / FlipY: sign of the transform's x / y scale (feature_uniform.rs:143-175)
    const bool flip_x = f->transform[0][0] < 0.0f, flip_y = f->transform[1][1] < 0.0f;
    if (!flip_x && !flip_y) {
        NORM_TRY(run_stages_unfused(ctx, ctx->main, p, in->v, out->v));
    } else {
        const size_t bytes = in->v.frames * align_up((size_t)in->v.w * 4, 256) * in->v.h;
        NORM_TRY(ensure_buf(ctx, ctx->flip_tmp, bytes));
        const ImgView tmp = view_of(ctx->flip_tmp.p, in->v.w, in->v.h, in->v.frames);
        NORM_TRY(run_stages_unfused(ctx, ctx->main, p, in->v, tmp));
        const size_t items = (size_t)((in->v.w + 3) / 4) * in->v.h * in->v.frames;
        flip_kernel<<<grid_for(ctx, (items + 255) / 256, 8), 256, 0, ctx->main.stream>>>(tmp, out->v, flip_x, flip_y);
        LAUNCH_CHECK(ctx);
    }
    if (flags & NORM_OUT_SRGB) NORM_TRY(encode_srgb_inplace(ctx, ctx->main.stream, out->v));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_chain_async(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags, const norm_img *in,
                                  norm_img *out, double *ssim_dev) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    NORM_TRY(check_uniform(ctx, f, in->v));
    return run_chain(ctx, ctx->main, f, flags, in->v, out->v, ssim_dev);
}

norm_status norm_cuda_chain(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags, const norm_img *in,
                            norm_img *out, double *ssim) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    NORM_TRY(check_uniform(ctx, f, in->v));
    double *d_ssim = nullptr;
    if (ssim) {
        NORM_TRY(ensure_results(ctx, in->v.frames));
        d_ssim = static_cast<double *>(ctx->results.p);
    }
    NORM_TRY(run_chain(ctx, ctx->main, f, flags, in->v, out->v, d_ssim));
    if (ssim)
        CU_TRY(ctx, cudaMemcpyAsync(ctx->h_results, d_ssim, (size_t)in->v.frames * sizeof(double),
                                    cudaMemcpyDeviceToHost, ctx->main.stream));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    if (ssim) std::memcpy(ssim, ctx->h_results, (size_t)in->v.frames * sizeof(double));
    return NORM_OK;
}

norm_status norm_cuda_chain_host(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags, const uint8_t *in_rgba8,
                                 uint8_t *out_rgba8, uint32_t frames, double *ssim) {
    NORM_TRY(bind(ctx));
    if (!f || !in_rgba8 || !out_rgba8 || frames == 0) return fail(ctx, NORM_E_INVALID_ARG, "NULL / empty argument");
    if (f->width == 0 || f->height == 0) return fail(ctx, NORM_E_INVALID_ARG, "zero-sized frames");
    const uint32_t w = f->width, h = f->height;
    ImgView probe = view_of(nullptr, w, h, 1);
    NORM_TRY(check_uniform(ctx, f, probe));
    NORM_TRY(ensure_pipe(ctx));
    NORM_TRY(ensure_results(ctx, frames));
    double *d_ssim = ssim ? static_cast<double *>(ctx->results.p) : nullptr;

    // chunk: ~64 MB of frames per stage buffer, at least 1 frame; three chunks in flight
    const size_t frame_bytes = probe.frame_stride, host_frame = (size_t)w * 4 * h;
    const size_t budget = ctx->chunk_bytes ? ctx->chunk_bytes : ((size_t)64 << 20);
    const uint32_t chunk = (uint32_t)std::max<size_t>(1, std::min<size_t>(frames, budget / frame_bytes));
    for (auto &s : ctx->pipe) {
        NORM_TRY(ensure_buf(ctx, s.stage_in, frame_bytes * chunk));
        NORM_TRY(ensure_buf(ctx, s.stage_out, frame_bytes * chunk));
    }
    // From the first enqueue on, a failure must not return while copies into / out of the caller's buffers are
    // still in flight on the pipe streams: drain them first, then report.
    norm_status st = NORM_OK;
    uint32_t ci = 0;
    for (uint32_t f0 = 0; f0 < frames && st == NORM_OK; f0 += chunk, ++ci) {
        Slot &s = ctx->pipe[ci % 3];
        const uint32_t nf = std::min(chunk, frames - f0);
        const ImgView vin = view_of(s.stage_in.p, w, h, nf), vout = view_of(s.stage_out.p, w, h, nf);
        st = copy_frames(ctx, s.stream, vin, 0, nf, const_cast<uint8_t *>(in_rgba8) + (size_t)f0 * host_frame, 0, true);
        if (st == NORM_OK) st = run_chain(ctx, s, f, flags, vin, vout, d_ssim ? d_ssim + f0 : nullptr);
        if (st == NORM_OK) st = copy_frames(ctx, s.stream, vout, 0, nf, out_rgba8 + (size_t)f0 * host_frame, 0, false);
    }
    if (st != NORM_OK) {
        const std::string why = ctx->err;
        for (auto &s : ctx->pipe) cudaStreamSynchronize(s.stream);
        (void)cudaGetLastError();
        return fail(ctx, st, why);
    }
    for (auto &s : ctx->pipe) CU_TRY(ctx, cudaStreamSynchronize(s.stream));
    if (ssim) {
        CU_TRY(ctx, cudaMemcpyAsync(ctx->h_results, d_ssim, (size_t)frames * sizeof(double), cudaMemcpyDeviceToHost,
                                    ctx->main.stream));
        CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
        std::memcpy(ssim, ctx->h_results, (size_t)frames * sizeof(double));
    }
    return NORM_OK;
}

norm_status norm_cuda_chain_band(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags, const norm_img *in,
                                 norm_img *out, uint32_t own_row0, uint32_t own_rows, double *moments) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    NORM_TRY(check_uniform(ctx, f, in->v));
    if (own_rows == 0 || (uint64_t)own_row0 + own_rows > in->v.h)
        return fail(ctx, NORM_E_INVALID_ARG, "owned row range outside the band image");
    double *d_m = nullptr;
    if (moments) {
        NORM_TRY(ensure_results(ctx, in->v.frames));
        d_m = static_cast<double *>(ctx->results.p);
    }
    NORM_TRY(run_chain_rows(ctx, ctx->main, f, flags, in->v, out->v, nullptr, d_m, own_row0, own_row0 + own_rows));
    if (moments)
        CU_TRY(ctx, cudaMemcpyAsync(ctx->h_results, d_m, (size_t)in->v.frames * 5 * sizeof(double),
                                    cudaMemcpyDeviceToHost, ctx->main.stream));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    if (moments) std::memcpy(moments, ctx->h_results, (size_t)in->v.frames * 5 * sizeof(double));
    return NORM_OK;
}

norm_status norm_cuda_chain_band_async(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags, const norm_img *in,
                                       norm_img *out, uint32_t own_row0, uint32_t own_rows, double *moments_dev) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    NORM_TRY(check_uniform(ctx, f, in->v));
    if (own_rows == 0 || (uint64_t)own_row0 + own_rows > in->v.h)
        return fail(ctx, NORM_E_INVALID_ARG, "owned row range outside the band image");
    return run_chain_rows(ctx, ctx->main, f, flags, in->v, out->v, nullptr, moments_dev, own_row0, own_row0 + own_rows);
}

norm_status norm_cuda_stream(norm_ctx *ctx, void **cuda_stream) {
    if (!ctx || !cuda_stream) return NORM_E_INVALID_ARG;
    *cuda_stream = ctx->main.stream;
    return NORM_OK;
}

norm_status norm_cuda_ipc_export(norm_ctx *ctx, const norm_img *img, uint8_t handle[64]) {
    NORM_TRY(bind(ctx));
    if (!img || !handle) return fail(ctx, NORM_E_INVALID_ARG, "NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle is 64 bytes");
    cudaIpcMemHandle_t h;
    CU_TRY(ctx, cudaIpcGetMemHandle(&h, img->v.ptr));
    std::memcpy(handle, &h, 64);
    return NORM_OK;
}

norm_status norm_cuda_ipc_open(norm_ctx *ctx, const uint8_t handle[64], void **device_ptr) {
    NORM_TRY(bind(ctx));
    if (!handle || !device_ptr) return fail(ctx, NORM_E_INVALID_ARG, "NULL argument");
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    CU_TRY(ctx, cudaIpcOpenMemHandle(device_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return NORM_OK;
}

norm_status norm_cuda_ipc_close(norm_ctx *ctx, void *device_ptr) {
    NORM_TRY(bind(ctx));
    if (!device_ptr) return NORM_OK;
    CU_TRY(ctx, cudaIpcCloseMemHandle(device_ptr));
    return NORM_OK;
}

norm_status norm_cuda_copy_rows(norm_ctx *ctx, norm_img *dst, uint32_t dst_row, const void *src_device_ptr,
                                size_t src_pitch, uint32_t src_row, uint32_t nrows) {
    NORM_TRY(bind(ctx));
    if (!dst || !src_device_ptr) return fail(ctx, NORM_E_INVALID_ARG, "NULL argument");
    if (nrows == 0) return NORM_OK;
    if ((uint64_t)dst_row + nrows > dst->v.h || src_pitch < (size_t)dst->v.w * 4)
        return fail(ctx, NORM_E_INVALID_ARG, "row range outside the destination image / source pitch too small");
    // device -> device; when src is a peer mapping (norm_cuda_ipc_open) this travels over NVLink
    CU_TRY(ctx, cudaMemcpy2DAsync(dst->v.ptr + (size_t)dst_row * dst->v.pitch, dst->v.pitch,
                                  static_cast<const uint8_t *>(src_device_ptr) + (size_t)src_row * src_pitch, src_pitch,
                                  (size_t)dst->v.w * 4, nrows, cudaMemcpyDeviceToDevice, ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_front(norm_ctx *ctx, const NormFeatureUniform *f, uint32_t flags, const norm_img *in,
                            norm_img *out) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, true));
    NORM_TRY(check_uniform(ctx, f, in->v));
    StagePlan p = make_plan(f, flags, true);
    NORM_TRY(launch_front(ctx, ctx->main.stream, in->v, out->v, p.lut, p.use_lut, p.gray, p.inv));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_blur(norm_ctx *ctx, uint32_t radius, const norm_img *in, norm_img *out) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    if (radius < 1 || radius > NORM_MAX_BLUR_RADIUS) return fail(ctx, NORM_E_INVALID_ARG, "blur radius outside 1..63");
    StagePlan p;
    bool identity;
    make_lut(NORM_CS_RAW, 0, p.lut, identity);
    p.blur = true, p.radius = radius;
    NORM_TRY(run_stages_unfused(ctx, ctx->main, p, in->v, out->v));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_sobel(norm_ctx *ctx, const norm_img *in, norm_img *out) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    NORM_TRY(launch_sobel(ctx, ctx->main.stream, in->v, out->v));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_box_blur(norm_ctx *ctx, const norm_img *in, norm_img *out) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    NORM_TRY(launch_box_blur(ctx, ctx->main.stream, in->v, out->v));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_overlay(norm_ctx *ctx, const NormShapeUniform *shapes, const NormCircleData *circles,
                              const NormDrawUniform *draw, const norm_img *in, norm_img *out) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    if (!shapes) return fail(ctx, NORM_E_INVALID_ARG, "shape uniform is NULL");
    if (shapes->num_circles > NORM_MAX_CIRCLES)
        return fail(ctx, NORM_E_INVALID_ARG, "num_circles exceeds NORM_MAX_CIRCLES (256, shape_uniform.rs:60)");
    if (shapes->num_circles && !circles) return fail(ctx, NORM_E_INVALID_ARG, "circles is NULL");
    if (shapes->width != in->v.w || shapes->height != in->v.h)
        return fail(ctx, NORM_E_DIM_MISMATCH, "shape uniform width/height differ from the image");
    static const SrgbTables tables = [] {
        SrgbTables t;
        make_srgb_tables(t.eotf, t.thresh);
        return t;
    }();
    OverlayParams op{};
    // shape_shader.wgsl:52-57: back to pixel coordinates, one f32 product each
    const float W = (float)shapes->width, H = (float)shapes->height,
                M = (float)std::min(shapes->width, shapes->height);
    for (uint32_t i = 0; i < shapes->num_circles; ++i) {
        op.cx[i] = circles[i].x * W;
        op.cy[i] = circles[i].y * H;
        op.rp[i] = circles[i].radius * M;
    }
    op.num_circles = shapes->num_circles;
    op.selected = shapes->selected_circle;
    if (draw && draw->crosshair == 1u) {
        op.crosshair = 1;
        op.mx = draw->circle_center_x, op.my = draw->circle_center_y, op.mr = draw->circle_radius;
    }
    const uint32_t tx = (in->v.w + kOvW - 1) / kOvW, ty = (in->v.h + kOvH - 1) / kOvH;
    const unsigned grid = grid_for(ctx, (size_t)tx * ty * in->v.frames, 8);
    overlay_kernel<<<grid, 256, 0, ctx->main.stream>>>(in->v, out->v, op, tables, tx, ty);
    LAUNCH_CHECK(ctx);
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

norm_status norm_cuda_sharpen(norm_ctx *ctx, uint32_t factor, const norm_img *in, norm_img *out) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, in, out, false));
    NORM_TRY(launch_sharpen(ctx, ctx->main.stream, in->v, out->v, factor));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    return NORM_OK;
}

// ---- SSIM ------------------------------------------------------------------------------

static norm_status ssim_impl(norm_ctx *ctx, const norm_img *a, const norm_img *b, double *moments, double *ssim,
                             int luma_a = 0, int luma_b = 0) {
    NORM_TRY(bind(ctx));
    NORM_TRY(check_pair(ctx, a, b, true));
    const uint32_t n = a->v.frames;
    NORM_TRY(ensure_results(ctx, n));
    double *d_m = static_cast<double *>(ctx->results.p), *d_s = d_m + (size_t)n * 5;
    NORM_TRY(launch_ssim(ctx, ctx->main, a->v, b->v, d_m, d_s, luma_a, luma_b));
    CU_TRY(ctx, cudaMemcpyAsync(ctx->h_results, d_m, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost,
                                ctx->main.stream));
    CU_TRY(ctx, cudaStreamSynchronize(ctx->main.stream));
    if (moments) std::memcpy(moments, ctx->h_results, (size_t)n * 5 * sizeof(double));
    if (ssim) std::memcpy(ssim, ctx->h_results + (size_t)n * 5, (size_t)n * sizeof(double));
    return NORM_OK;
}

norm_status norm_cuda_ssim(norm_ctx *ctx, const norm_img *a, const norm_img *b, double *ssim) {
    if (ctx && !ssim) return fail(ctx, NORM_E_INVALID_ARG, "ssim is NULL");
    return ssim_impl(ctx, a, b, nullptr, ssim);
}

norm_status norm_cuda_ssim_luma(norm_ctx *ctx, const norm_img *a, int32_t luma_a, const norm_img *b, int32_t luma_b, double *ssim) {
    if (ctx && !ssim) return fail(ctx, NORM_E_INVALID_ARG, "ssim is NULL");
    if (luma_a < 0 || luma_a > 2 || luma_b < 0 || luma_b > 2) return fail(ctx, NORM_E_INVALID_ARG, "unknown luma mode");
    return ssim_impl(ctx, a, b, nullptr, ssim, luma_a, luma_b);
}

norm_status norm_cuda_ssim_moments(norm_ctx *ctx, const norm_img *a, const norm_img *b, double *moments) {
    if (ctx && !moments) return fail(ctx, NORM_E_INVALID_ARG, "moments is NULL");
    return ssim_impl(ctx, a, b, moments, nullptr);
}

double norm_ssim_from_moments(const double moments[5], double npix) { return ssim_from_moments(moments, npix); }

// ---- SSIM for callers that may not allocate ----------------------------------------------
//
// norm_cuda_ssim needs scratch for its per-CTA partial sums and takes it from a context, which
// gets it from cudaMalloc.  Two kinds of caller cannot live with that: a frame loop that must
// never touch the device allocator once it is running (cudaMalloc/cudaFree synchronise the
// whole device and would stall every frame in flight), and a host application that has already
// handed all device memory to its own pool, where a library-side cudaMalloc simply fails.
// For them the partial sums live in this fixed array inside the library image.  Every GPU that
// loads the library carries these 160 KiB until the process exits.
//
// `g_ssim_static_scratch` below is, for host code, only a handle: its host-side address is not
// a device address.  The device address is looked up once per device with
// cudaGetSymbolAddress (that device being current) and that pointer is what the kernels get.
#define NORM_SSIM_STATIC_SLOTS 4096  /* partial-sum slots of 5 doubles */
static __device__ __align__(256) double g_ssim_static_scratch[NORM_SSIM_STATIC_SLOTS * 5];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
double *g_static_scratch_dev[kMaxDevices] = {nullptr};

norm_status static_fail(norm_status s, const std::string &msg) {
    g_create_error = msg;
    return s;
}
}  // namespace

norm_status norm_cuda_ssim_static(const void *a_dev, const void *b_dev, uint32_t width, uint32_t height,
                                  uint32_t frames, size_t pitch, size_t frame_stride, double *ssim_dev,
                                  double *moments_dev, void *cuda_stream) {
    if (!a_dev || !b_dev || (!ssim_dev && !moments_dev))
        return static_fail(NORM_E_INVALID_ARG, "norm_cuda_ssim_static: NULL image or no output pointer");
    if (width == 0 || height == 0 || frames == 0)
        return static_fail(NORM_E_INVALID_ARG, "norm_cuda_ssim_static: zero-sized input");
    if (frames > NORM_SSIM_STATIC_MAX_FRAMES)
        return static_fail(NORM_E_INVALID_ARG,
                           "norm_cuda_ssim_static: " + std::to_string(frames) + " frames exceed the fixed scratch; at most " +
                                   std::to_string(NORM_SSIM_STATIC_MAX_FRAMES) +
                                   " frames per call (split the batch, or use norm_cuda_ssim which allocates)");
    if (pitch % 16 || frame_stride % 16 || (uintptr_t)a_dev % 16 || (uintptr_t)b_dev % 16 ||
        pitch < (size_t)width * 4 || frame_stride < pitch * height)
        return static_fail(NORM_E_INVALID_ARG, "norm_cuda_ssim_static: pitch / frame_stride / pointers must be 16-byte aligned and large enough");

    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        (void)cudaGetLastError();
        return static_fail(NORM_E_CUDA, std::string("norm_cuda_ssim_static: cudaGetDevice: ") + cudaGetErrorString(e));
    }
    double *scratch = nullptr;
    {
        // one-time set-up for this device: resolve the array's device address
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (!g_static_scratch_dev[dev]) {
            double *ptr = nullptr;
            e = cudaGetSymbolAddress(reinterpret_cast<void **>(&ptr), g_ssim_static_scratch);
            if (e != cudaSuccess || !ptr) {
                (void)cudaGetLastError();
                return static_fail(NORM_E_CUDA, std::string("norm_cuda_ssim_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
            }
            g_static_scratch_dev[dev] = ptr;
        }
        scratch = g_static_scratch_dev[dev];
    }

    ImgView a, b;
    a.ptr = static_cast<uint8_t *>(const_cast<void *>(a_dev));
    a.pitch = pitch, a.frame_stride = frame_stride, a.w = width, a.h = height, a.frames = frames;
    b = a;
    b.ptr = static_cast<uint8_t *>(const_cast<void *>(b_dev));

    // partial-sum slots per frame: enough CTAs to fill the GPU, never more than the array holds.
    // Every slot a launch reads is written by that same launch first, so nothing has to be cleared.
    const size_t chunks = (size_t)((width + 3) / 4) * height;
    const size_t max_useful = (chunks + 256 * 4 - 1) / (256 * 4);
    size_t want = (size_t)NORM_SSIM_STATIC_SLOTS / frames;
    want = std::min(want, max_useful);
    want = std::min(want, ((size_t)1184 + frames - 1) / frames);
    const uint32_t bpf = (uint32_t)std::max<size_t>(1, want);
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    ssim_moments_kernel<0, 0><<<dim3(bpf, frames), 256, 0, s>>>(a, b, scratch);
    if ((e = cudaGetLastError()) != cudaSuccess)
        return static_fail(NORM_E_CUDA, std::string("norm_cuda_ssim_static: launch: ") + cudaGetErrorString(e));
    ssim_finalize_kernel<<<frames, 256, 0, s>>>(scratch, bpf, (double)width * (double)height, moments_dev, ssim_dev);
    if ((e = cudaGetLastError()) != cudaSuccess)
        return static_fail(NORM_E_CUDA, std::string("norm_cuda_ssim_static: launch: ") + cudaGetErrorString(e));
    return NORM_OK;
}

// ---- host-only helpers -----------------------------------------------------------------

norm_status norm_blur_taps(uint32_t radius, float *taps, int32_t *dmin) {
    Taps t;
    if (!taps || !dmin || make_taps(radius, t) != NORM_OK) return NORM_E_INVALID_ARG;
    std::memcpy(taps, t.t, radius * sizeof(float));
    *dmin = t.dmin;
    return NORM_OK;
}

norm_status norm_srgb_tables(float eotf[256], float thresh[256]) {
    if (!eotf || !thresh) return NORM_E_INVALID_ARG;
    make_srgb_tables(eotf, thresh);
    return NORM_OK;
}

norm_status norm_srgb_encode_lut(uint8_t lut[256]) {
    if (!lut) return NORM_E_INVALID_ARG;
    Lut256 l;
    make_srgb_encode_lut(l);
    std::memcpy(lut, l.v, 256);
    return NORM_OK;
}

norm_status norm_front_lut(uint32_t flags, uint32_t gamma, uint8_t lut[256]) {
    if (!lut) return NORM_E_INVALID_ARG;
    Lut256 l;
    bool identity;
    make_lut(flags, gamma, l, identity);
    std::memcpy(lut, l.v, 256);
    return NORM_OK;
}

norm_status norm_plan_frames(uint32_t frames, int32_t nranks, int32_t rank, uint32_t *frame0, uint32_t *nframes) {
    if (nranks < 1 || rank < 0 || rank >= nranks || !frame0 || !nframes) return NORM_E_INVALID_ARG;
    const uint32_t base = frames / (uint32_t)nranks, rem = frames % (uint32_t)nranks;
    *frame0 = (uint32_t)rank * base + std::min<uint32_t>((uint32_t)rank, rem);
    *nframes = base + ((uint32_t)rank < rem ? 1u : 0u);
    return NORM_OK;
}

norm_status norm_plan_band(const NormFeatureUniform *f, uint32_t height, int32_t nranks, int32_t rank, uint32_t *row0,
                           uint32_t *rows, uint32_t *halo_up, uint32_t *halo_down) {
    if (!f || !row0 || !rows || !halo_up || !halo_down) return NORM_E_INVALID_ARG;
    NORM_TRY(norm_plan_frames(height, nranks, rank, row0, rows));
    uint32_t up = 0, down = 0;
    NORM_TRY(stencil_reach(f, &up, &down));
    *halo_up = std::min(up, *row0);
    *halo_down = std::min(down, height - (*row0 + *rows));
    return NORM_OK;
}

}  // extern "C"

#include "band.inc"
