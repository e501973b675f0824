// Shape overlay, composite and crosshair: the last arithmetic between the effect
// output and the final pixel.  Follows shape_shader.wgsl:41-89 (circle fill with smoothstep AA,
// selected-circle outline, sequential mix blending), the ALPHA_BLENDING store into the sRGB shape
// texture (gpu_state.rs:321,404-418; cleared transparent, app_state.rs:589), and the composite +
// crosshair of image_shader.wgsl:176-193.  Same fixed operation order as oracle_overlay().
//
// HBM-bound for the pixels nothing touches (16-byte load -> 16-byte store); every CTA first
// culls the circle list against its tile so the per-pixel loop only walks circles that can
// reach the tile (culling is conservative, so the result equals the full loop).
#pragma once
#include "common.cuh"

namespace normk {

constexpr int kMaxCircles = 256;              // shape_uniform.rs:60 MAX_CIRCLES
constexpr int kOvW = 128, kOvH = 64;          // CTA tile (one cull); walked as 8 slabs of 4 x 2 warp blocks of 32 x 4 pixels
constexpr int kOvSlab = 8;

struct OverlayParams {
    float cx[kMaxCircles], cy[kMaxCircles], rp[kMaxCircles];  // pixel-space centre / radius
    uint32_t num_circles, selected;
    uint32_t crosshair;
    float mx, my, mr;  // draw_uniform.rs: circle_center_x/y, circle_radius (pixels)
};

struct SrgbTables {
    float eotf[256];    // decoded value of code k
    float thresh[256];  // thresh[j]: encoder steps from code j-1 to j at this linear value
};

__device__ __forceinline__ float smoothstep_rn(float e0, float e1, float x) {
    float t = __fdiv_rn(__fsub_rn(x, e0), __fsub_rn(e1, e0));
    t = fminf(fmaxf(t, 0.0f), 1.0f);
    return __fmul_rn(__fmul_rn(t, t), __fsub_rn(3.0f, __fmul_rn(2.0f, t)));
}

__device__ __forceinline__ float mix_rn(float a, float b, float t) {
    return __fmaf_rn(b, t, __fmul_rn(a, __fsub_rn(1.0f, t)));
}

__device__ __forceinline__ uint32_t unorm8_rn(float v) {
    v = fminf(fmaxf(v, 0.0f), 1.0f);
    return (uint32_t)__float2int_rn(__fmul_rn(v, 255.0f));
}

// sRGB code of a linear value = the largest j with v >= thresh[j] (thresh[0] = -inf side: every v, NaN included, gives >= 0).
__device__ __forceinline__ uint32_t srgb_code_search(const float *thresh, float v) {
    uint32_t lo = 0, hi = 255;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const uint32_t mid = (lo + hi + 1) >> 1;
        if (v >= thresh[mid]) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}
// The same code from a first guess: guess[i] is the code of the value i / 4095, a lower bound for every v in that cell
// (at most two codes short next to zero, where the curve is steepest), then stepped against the thresholds until it
// is exact -- two dependent look-ups on average instead of eight.
constexpr int kSrgbGuess = 4096;
__device__ __forceinline__ uint32_t srgb_code(const float *thresh, const uint8_t *guess, float v) {
    const int cell = min(max(__float2int_rd(__fmul_rn(v, (float)(kSrgbGuess - 1))), 0), kSrgbGuess - 1);  // NaN -> 0
    uint32_t c = guess[cell];
    while (c < 255u && v >= thresh[c + 1]) ++c;
    while (c > 0u && v < thresh[c]) --c;  // (the cell index is a rounded product: it can land one cell high)
    return c;
}

// Shape layer + composite + crosshair for the four pixels (x .. x+3, y) of one lane; (wbx, wby) is the warp's
// 32 x 4 block, whose pixel-centre box drives the warp-uniform reach test.  Not inlined: the caller walks
// eight slabs with its source chunks in registers and only tiles with a non-empty circle list get here.
struct OverlayLists {
    const float *cx, *cy, *rp;
    const uint16_t *idx;
    const float *eotf, *thresh, *div255;  // div255[b] = b / 255 (IEEE division, tabulated once per CTA)
    const uint8_t *guess;
    uint32_t cnt, selected;
    bool cross;
    float mx, my, mr;
    int w, h;
};

__device__ __noinline__ uint4 overlay_shade(uint4 src, int x, int y, int wbx, int wby, const OverlayLists &L) {
    uint32_t px[4] = {src.x, src.y, src.z, src.w};
    const float *s_cx = L.cx, *s_cy = L.cy, *s_rp = L.rp, *s_eotf = L.eotf, *s_thresh = L.thresh, *s_div = L.div255;
    const uint8_t *s_guess = L.guess;
    const uint16_t *s_idx = L.idx;
    const uint32_t cnt = L.cnt, selected = L.selected;
    const bool cross = L.cross;
    const float mx = L.mx, my = L.my, mr = L.mr;
    struct { int w, h; } in = {L.w, L.h};
    const float py = __fadd_rn((float)y, 0.5f);
    // pixel-centre box of the warp's block; distance from a point to it (0 inside)
    const float bx0 = (float)wbx + 0.5f, bx1 = (float)min(wbx + 32, (int)in.w) - 0.5f;
    const float by0 = (float)wby + 0.5f, by1 = (float)min(wby + 4, (int)in.h) - 0.5f;
    auto box_dist = [&](float qx, float qy) {
        const float ddx = fmaxf(fmaxf(bx0 - qx, qx - bx1), 0.0f), ddy = fmaxf(fmaxf(by0 - qy, qy - by1), 0.0f);
        return sqrtf(ddx * ddx + ddy * ddy);
    };
    const bool blk_cross = cross && box_dist(mx, my) <= mr + 1.0f;
    float c0[4], c1[4], c2[4], c3[4];  // shape colour of the lane's four pixels
#pragma unroll
    for (int j = 0; j < 4; ++j) c0[j] = c1[j] = c2[j] = c3[j] = 0.0f;
    bool any = blk_cross;
    for (uint32_t k = 0; k < cnt; ++k) {
        const float rp = s_rp[k], ccx = s_cx[k], ccy = s_cy[k];
        // warp-uniform: neither the disc nor the 3-px outline reaches this block (1 px of slack for rounding;
        // false for NaN parameters, which touch nothing)
        if (!(box_dist(ccx, ccy) <= rp + 4.0f)) continue;
        any = true;
        const bool sel = s_idx[k] == selected;
        const float dy = __fsub_rn(py, ccy);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float pxc = __fadd_rn((float)(x + j), 0.5f);
            const float dx = __fsub_rn(pxc, ccx);
            const float d = __fsqrt_rn(__fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
            if (d <= rp) {
                const float a = __fsub_rn(1.0f, smoothstep_rn(__fsub_rn(rp, 1.0f), rp, d));
                c0[j] = mix_rn(c0[j], 0.0f, a), c1[j] = mix_rn(c1[j], 1.0f, a);
                c2[j] = mix_rn(c2[j], 1.0f, a), c3[j] = mix_rn(c3[j], a, a);
            }
            if (sel) {
                const float od = fabsf(__fsub_rn(d, rp));
                if (od <= 3.0f) {
                    const float a = __fsub_rn(1.0f, smoothstep_rn(0.0f, 3.0f, od));
                    c0[j] = mix_rn(c0[j], 1.0f, a), c1[j] = mix_rn(c1[j], 1.0f, a);
                    c2[j] = mix_rn(c2[j], 0.0f, a), c3[j] = mix_rn(c3[j], a, a);
                }
            }
        }
    }
    if (any) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (x + j >= (int)in.w) break;
            const uint32_t ca = unorm8_rn(c3[j]);
            bool hit = false;
            if (blk_cross) {
                const float dx = __fsub_rn(__fadd_rn((float)(x + j), 0.5f), mx), dy = __fsub_rn(py, my);
                hit = __fsqrt_rn(__fmaf_rn(dy, dy, __fmul_rn(dx, dx))) < mr;
            }
            if (ca == 0 && !hit) continue;
            float p[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) p[c] = s_div[(px[j] >> (8 * c)) & 0xffu];
            if (ca > 0) {
                const float sa = s_div[ca];
                p[0] = mix_rn(p[0], s_eotf[srgb_code(s_thresh, s_guess, __fmul_rn(c0[j], c3[j]))], sa);
                p[1] = mix_rn(p[1], s_eotf[srgb_code(s_thresh, s_guess, __fmul_rn(c1[j], c3[j]))], sa);
                p[2] = mix_rn(p[2], s_eotf[srgb_code(s_thresh, s_guess, __fmul_rn(c2[j], c3[j]))], sa);
                p[3] = mix_rn(p[3], sa, sa);
            }
            if (hit) p[0] = 0.0f, p[1] = 1.0f, p[2] = 1.0f;
            px[j] = pack_rgba(unorm8_rn(p[0]), unorm8_rn(p[1]), unorm8_rn(p[2]), unorm8_rn(p[3]));
        }
    }
    return make_uint4(px[0], px[1], px[2], px[3]);
}

__global__ void __launch_bounds__(256) overlay_kernel(ImgView in, ImgView out, const __grid_constant__ OverlayParams op,
                                                      const __grid_constant__ SrgbTables tab, uint32_t tiles_x,
                                                      uint32_t tiles_y) {
    __shared__ float s_cx[kMaxCircles], s_cy[kMaxCircles], s_rp[kMaxCircles];
    __shared__ uint16_t s_idx[kMaxCircles];
    __shared__ uint32_t s_warp_cnt[8];
    __shared__ float s_eotf[256], s_thresh[256], s_div255[256];
    __shared__ uint8_t s_guess[kSrgbGuess];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    s_eotf[tid] = tab.eotf[tid];
    s_thresh[tid] = tab.thresh[tid];
    s_div255[tid] = __fdiv_rn((float)tid, 255.0f);
    __syncthreads();
    for (int i = tid; i < kSrgbGuess; i += 256) s_guess[i] = (uint8_t)srgb_code_search(s_thresh, __fdiv_rn((float)i, (float)(kSrgbGuess - 1)));
    const uint32_t n = op.num_circles, selected = op.selected;
    const bool cross = op.crosshair == 1u;
    const float mx = op.mx, my = op.my, mr = op.mr;
    __shared__ OverlayLists lists;  // read by overlay_shade(); in shared memory so the callee's loads are cheap
    if (tid == 0) lists = OverlayLists{s_cx, s_cy, s_rp, s_idx, s_eotf, s_thresh, s_div255, s_guess, 0u, selected, cross, mx, my, mr, (int)in.w, (int)in.h};
    const uint32_t tiles_per_frame = tiles_x * tiles_y;
    const size_t total = (size_t)tiles_per_frame * in.frames;
    for (size_t tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const uint32_t f = (uint32_t)(tile / tiles_per_frame);
        const uint32_t tr = (uint32_t)(tile - (size_t)f * tiles_per_frame);
        const int x0 = (int)(tr % tiles_x) * kOvW, y0 = (int)(tr / tiles_x) * kOvH;
        // this thread's eight 16-byte source chunks (one per slab), all in flight while the circles are culled
        uint4 srcs[kOvH / kOvSlab];
        {
            const int lx = x0 + (warp & 3) * 32 + 4 * (lane & 7);
#pragma unroll
            for (int slab = 0; slab < kOvH / kOvSlab; ++slab) {
                const int ly = y0 + slab * kOvSlab + (warp >> 2) * 4 + (lane >> 3);
                srcs[slab] = make_uint4(0, 0, 0, 0);
                if (ly < (int)in.h && lx < (int)in.w)
                    srcs[slab] = ldg_stream(reinterpret_cast<const uint4 *>(row_ptr(in, f, ly)) + (lx >> 2));
            }
        }
        __syncthreads();
        // ---- cull: thread i looks at circle i; ordered compaction keeps the blending order
        bool keep = false;
        if ((uint32_t)tid < n) {
            const float cx = op.cx[tid], cy = op.cy[tid], rp = op.rp[tid];
            // distance from the centre to the tile's pixel-centre box, with 4 px of margin for
            // the outline (3 px) and rounding
            const float bx0 = (float)x0 + 0.5f, bx1 = (float)min(x0 + kOvW, (int)in.w) - 0.5f;
            const float by0 = (float)y0 + 0.5f, by1 = (float)min(y0 + kOvH, (int)in.h) - 0.5f;
            const float ddx = fmaxf(fmaxf(bx0 - cx, cx - bx1), 0.0f), ddy = fmaxf(fmaxf(by0 - cy, cy - by1), 0.0f);
            const float dmin = sqrtf(ddx * ddx + ddy * ddy);
            // NaN parameters never satisfy d <= rp, so dropping them here changes nothing
            keep = dmin <= rp + 4.0f;
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_warp_cnt[warp] = __popc(bal);
        __syncthreads();
        uint32_t base = 0, cnt = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            if (w < warp) base += s_warp_cnt[w];
            cnt += s_warp_cnt[w];
        }
        if (tid == 0) lists.cnt = cnt;  // visible after the barrier below
        if (keep) {
            const uint32_t pos = base + __popc(bal & ((1u << lane) - 1u));
            s_idx[pos] = (uint16_t)tid;
            s_cx[pos] = op.cx[tid], s_cy[pos] = op.cy[tid], s_rp[pos] = op.rp[tid];
        }
        __syncthreads();

#pragma unroll
        for (int slab = 0; slab < kOvH / kOvSlab; ++slab) {
        // a warp owns a block of 32 x 4 pixels of the slab (lane = 4 pixels of one row): small enough that the
        // warp-uniform reach test below skips most circles of the tile's list
        const int wbx = x0 + (warp & 3) * 32, wby = y0 + slab * kOvSlab + (warp >> 2) * 4;
        const int y = wby + (lane >> 3), x = wbx + 4 * (lane & 7);
        if (y >= (int)in.h || x >= (int)in.w) continue;
        const uint4 src = srcs[slab];
        uint32_t px[4] = {src.x, src.y, src.z, src.w};
        if (cnt != 0 || cross) {
            const uint4 r = overlay_shade(make_uint4(px[0], px[1], px[2], px[3]), x, y, wbx, wby, lists);
            px[0] = r.x, px[1] = r.y, px[2] = r.z, px[3] = r.w;
        }
        stg_stream(reinterpret_cast<uint4 *>(const_cast<uint8_t *>(row_ptr(out, f, y))) + (x >> 2),
                   make_uint4(px[0], px[1], px[2], px[3]));
        }  // slab
    }
}

}  // namespace normk
