// Fused chain for the smallest blur (r = 3): front -> 3-tap separable blur -> Sobel -> SSIM on register windows.
//
// At r = 3 the pipelined kernel (chain2_kernels.cuh) spends most of its ~120 instructions per pixel on hand-offs
// between its warp groups: shared-memory planes, barriers, three address streams.  Three taps need none of that.
// One warp owns a tile of 120 output columns x tile_rows rows, lane i loads source columns x0 - 4 + 4i .. + 3 with one
// 16-byte load per row (lanes 0 and 31 only feed their neighbours' halos, as in stencil_kernels.cuh) and marches down
// the rows keeping in registers: three source rows in flight (a ring refilled three steps ahead), the last two H-pass
// rows, intensity rows and source-luma rows, and the five f64 SSIM moments.  Horizontal neighbours come by warp shuffle.  No
// shared memory beyond the two 256-entry tables, no barriers, one pass over HBM (4 B read + 4 B written per pixel).
//
// Arithmetic: the S-form of oracle/norm_oracle.c, operation for operation what chain_fused2_kernel executes
// (H and V pass acc = t0*v0, fma(t1, v1, acc), fma(t2, v2, acc) on source column / row offsets -1, 0, +1 with
// ClampToEdge; quantise; intensity from the exact integer radicand; sobel_mag_bits; f32 lumas, f64 moments).
// image_shader.wgsl:81-103 (blur), :132-156 (Sobel), ssim.rs:65-81 (luma).
#pragma once
#include "stencil_kernels.cuh"

namespace normk {

constexpr int kCwCols = 120;  // output columns per warp tile (lanes 1..30 x 4 pixels)
#ifndef NORM_CW_WARPS
#define NORM_CW_WARPS 8       // warps per CTA; two CTAs per SM
#endif
constexpr int kCwWarps = NORM_CW_WARPS, kCwThreads = 32 * kCwWarps;

template <bool GRAY, int FRONT>  // FRONT: 0 no front stage, 1 with the 256-entry table, 2 grayscale / inversion without a table
__global__ void __launch_bounds__(kCwThreads, 2) chain_win3_kernel(const __grid_constant__ FusedParams p, uint32_t tiles_x, uint32_t tiles_y,
                                                            uint32_t tile_rows) {
    constexpr int CH = GRAY ? 1 : 3;                       // after grayscale the three colour channels coincide
    constexpr int FR = (FRONT == 1 ? 1 : 0) | (GRAY ? 2 : 0);   // s3_front: bit 0 table, bit 1 grayscale
    __shared__ double lumaLutD[256];  // luma of the output pixel (c, c, c), as the f64 the moments take
    __shared__ uint8_t slut[256];
    for (int i = threadIdx.x; i < 256; i += kCwThreads) {
        const float c = (float)i;
        lumaLutD[i] = (double)__fadd_rn(__fadd_rn(__fmul_rn(c, kLumaR), __fmul_rn(c, kLumaG)), __fmul_rn(c, kLumaB));
        slut[i] = p.lut.v[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int W = (int)p.in.w, H = (int)p.in.h;
    const float t0 = p.taps.t[0], t1 = p.taps.t[1], t2 = p.taps.t[2];
    const float2 t0p = make_float2(t0, t0), t1p = make_float2(t1, t1), t2p = make_float2(t2, t2);
    const uint32_t inv_mask = p.invert ? 0x00ffffffu : 0u;
    const uint32_t tiles_per_frame = tiles_x * tiles_y;
    const size_t total = (size_t)tiles_per_frame * p.in.frames;
    const uint32_t ipitch = (uint32_t)p.in.pitch, opitch = (uint32_t)p.out.pitch;

    struct Raw {
        uint32_t p[4];
    };
    struct HRow {
        float2 v[CH][2];  // H-pass outputs of this lane's four columns, as column pairs
    };
    struct IRow {
        float v[6];       // intensity: left neighbour, four own columns, right neighbour
    };
    struct LRow {
        float v[4];       // SSIM luma of the raw source pixels
    };
    // steps per tile, the same for every tile: tile_rows + 2 blurred rows (+ 1: the image's bottom row), in whole trips of three
    const uint32_t steps = (tile_rows + 3 + 2) / 3 * 3;

    for (size_t tile = (size_t)blockIdx.x * kCwWarps + (threadIdx.x >> 5); tile < total; tile += (size_t)gridDim.x * kCwWarps) {
        const uint32_t f = (uint32_t)(tile / tiles_per_frame);
        const uint32_t tr = (uint32_t)(tile - (size_t)f * tiles_per_frame);
        const int x0 = (int)(tr % tiles_x) * kCwCols;
        const int ya = (int)p.out_y0 + (int)(tr / tiles_x) * (int)tile_rows, yb = min(ya + (int)tile_rows, (int)p.out_y1);
        const int col = x0 - 4 + 4 * lane;
        const bool vec = col >= 0 && col + 3 < W;
        const bool writes = lane >= 1 && lane <= 30 && col < W;
        const int colc = clampi(col, 0, ((W - 1) >> 2) << 2);  // every lane loads ONE aligned chunk (rows are pitched to 16 bytes)
        const bool border_tile = x0 == 0 || x0 + kCwCols + 4 > W;
        const int elast = W - 1 - col;                          // this lane's element that holds column W-1 (may lie outside 0..3)
        const uint8_t *lbase = p.in.ptr + (size_t)f * p.in.frame_stride + (size_t)colc * 4;
        uint8_t *obase = p.out.ptr + (size_t)f * p.out.frame_stride;
        // blurred rows this tile needs: one above and one below its output rows, inside the image
        const int o_lo = max(ya - 1, 0), o_hi = min(yb, H - 1);
        const int NB = o_hi - o_lo + 1;
        // step n computes blurred row o_lo + n from the source rows q = n, n+1, n+2 of the clamped sequence
        // row(q) = clamp(o_lo - 1 + q), and emits output row o_lo + n - 1; at the image's bottom row step NB emits row H-1.

        auto fetch = [&](int q) {
            const uint4 v = ldg_stream(reinterpret_cast<const uint4 *>(lbase + (uint64_t)(uint32_t)clampi(o_lo - 1 + q, 0, H - 1) * ipitch));
            Raw r;
            r.p[0] = v.x, r.p[1] = v.y, r.p[2] = v.z, r.p[3] = v.w;
            return r;
        };
        // H pass of one source row: SSIM luma of the raw pixels (kept for the step that emits this row), front stage,
        // unpack, neighbours by shuffle, three taps per output
        auto hrow = [&](const Raw &rin, HRow &h, LRow &l) {
            Raw q = rin;
            if (border_tile && !vec) {  // ClampToEdge on the source columns
                const int le = (W - 1) & 3;
                const uint32_t last = le == 0 ? q.p[0] : le == 1 ? q.p[1] : le == 2 ? q.p[2] : q.p[3];
                const uint32_t edge = col < 0 ? q.p[0] : last;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (col + j < 0 || col + j > W - 1) q.p[j] = edge;
            }
            uint32_t px[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) l.v[j] = luma_ssim(q.p[j]), px[j] = FRONT == 0 ? q.p[j] : s3_front<FR>(q.p[j], slut, inv_mask);  // (no front stage, no inversion either)
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                float v[6];
#pragma unroll
                for (int j = 0; j < 4; ++j) v[1 + j] = byte_f(px[j], c);
                v[0] = __shfl_up_sync(0xffffffffu, v[4], 1);
                v[5] = __shfl_down_sync(0xffffffffu, v[1], 1);
                float o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) o[j] = __fmaf_rn(t2, v[j + 2], __fmaf_rn(t1, v[j + 1], __fmul_rn(t0, v[j])));
                h.v[c][0] = make_float2(o[0], o[1]), h.v[c][1] = make_float2(o[2], o[3]);
            }
        };
        // V pass + quantise + intensity of one blurred row, neighbours by shuffle, ClampToEdge of the BLURRED image
        auto irow = [&](const HRow &ha, const HRow &hb, const HRow &hc, IRow &ir) {
            uint32_t qb[CH][4];
#pragma unroll
            for (int c = 0; c < CH; ++c)
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const float2 o = __ffma2_rn(t2p, hc.v[c][k], __ffma2_rn(t1p, hb.v[c][k], __fmul2_rn(t0p, ha.v[c][k])));
                    qb[c][2 * k] = __float_as_uint(__fadd_rn(o.x, 12582912.0f));
                    qb[c][2 * k + 1] = __float_as_uint(__fadd_rn(o.y, 12582912.0f));
                }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint32_t rgb;
                if (GRAY) {
                    rgb = __byte_perm(qb[0][j], 0u, 0x4000);  // (y, y, y, 0)
                } else {
                    const uint32_t rg = __byte_perm(qb[0][j], qb[CH > 1 ? 1 : 0][j], 0x7740);
                    rgb = __byte_perm(rg, qb[CH - 1][j], 0x5410);
                }
                ir.v[1 + j] = sqrt_rn_inrange0((float)__dp4a(rgb, rgb, 0u));
            }
            ir.v[0] = __shfl_up_sync(0xffffffffu, ir.v[4], 1);
            ir.v[5] = __shfl_down_sync(0xffffffffu, ir.v[1], 1);
            if (border_tile) {  // the Sobel taps sample the blurred image with ClampToEdge: column -1 = column 0, W = W-1
                if (col == 0) ir.v[0] = ir.v[1];
                if (elast == 0) ir.v[2] = ir.v[1];
                if (elast == 1) ir.v[3] = ir.v[2];
                if (elast == 2) ir.v[4] = ir.v[3];
                if (elast == 3) ir.v[5] = ir.v[4];
            }
        };

        double m5[5] = {0, 0, 0, 0, 0};
        Raw r0 = fetch(0), r1 = fetch(1), r2 = fetch(2);
        HRow ha, hb, hc;
        IRow ia, ib, ic;
        LRow la, lb, lc;  // SSIM lumas of the source rows q = n, n+1, n+2
#pragma unroll
        for (int j = 0; j < 6; ++j) ia.v[j] = ib.v[j] = ic.v[j] = 0.0f;
        hrow(r0, ha, la);
        r0 = fetch(3);
        hrow(r1, hb, lb);
        r1 = fetch(4);

// One step.  RC holds source row q = n + 2 (unpacked here, then refilled with row q = n + 5: three loads in flight);
// LN the lumas of row q = n, the row of the output emitted here.  Every step of every tile executes the same
// instruction sequence (the step count comes from a kernel parameter), so the shuffles sit in uniform control flow;
// steps past the tile's last row compute on clamped rows and emit nothing.
#define CW_STEP(N_, RC, HA, HB, HC, IA, IB, IC, LN, LC)                                                           \
    {                                                                                                             \
        const int n_ = (N_), y_ = o_lo + n_ - 1;                                                                  \
        hrow(RC, HC, LC);                                                                                         \
        RC = fetch(n_ + 5);                                                                                       \
        irow(HA, HB, HC, IC);                                                                                     \
        if (n_ == NB) IC = IB; /* below the image's last row: the row itself */                                   \
        const bool emit = y_ >= ya && y_ < yb;                                                                    \
        const bool srow = emit && p.do_ssim && (uint32_t)(y_ - (int)p.ssim_y0) < p.ssim_y1 - p.ssim_y0;           \
        uint32_t o[4], cbs[4];                                                                                    \
        _Pragma("unroll") for (int j = 0; j < 4; ++j) {                                                           \
            cbs[j] = sobel_mag_bits(IC.v[j], IC.v[j + 1], IC.v[j + 2], IB.v[j], IB.v[j + 2],                      \
                                    IA.v[j], IA.v[j + 1], IA.v[j + 2]);                                           \
            o[j] = __byte_perm(cbs[j], 0xffu, 0x4000);                                                            \
        }                                                                                                         \
        if (srow) { /* warp-uniform; lanes that own no pixel accumulate too and are cleared at the end of the tile */ \
            _Pragma("unroll") for (int j = 0; j < 4; ++j) {                                                       \
                double xv = lumaLutD[cbs[j] & 0xffu], yv = (double)LN.v[j];                                \
                if (border_tile && col + j >= W) xv = 0.0, yv = 0.0;                                              \
                m5[0] += xv;                                                                                      \
                m5[1] += yv;                                                                                      \
                m5[2] = fma(xv, xv, m5[2]);                                                                       \
                m5[3] = fma(yv, yv, m5[3]);                                                                       \
                m5[4] = fma(xv, yv, m5[4]);                                                                       \
            }                                                                                                     \
        }                                                                                                         \
        if (emit && writes) {                                                                                     \
            uint8_t *orow = obase + (uint64_t)(uint32_t)y_ * opitch;                                              \
            if (vec) {                                                                                            \
                stg_stream(reinterpret_cast<uint4 *>(orow + (size_t)col * 4), make_uint4(o[0], o[1], o[2], o[3])); \
            } else {                                                                                              \
                _Pragma("unroll") for (int j = 0; j < 4; ++j) if (col + j < W)                                    \
                        reinterpret_cast<uint32_t *>(orow)[col + j] = o[j];                                       \
            }                                                                                                     \
        }                                                                                                         \
        if (n_ == 0 && o_lo == 0) IB = IC; /* row -1 of the blurred image = row 0 (IB is the next step's IA) */   \
    }
#pragma unroll 1
        for (int n = 0; n < (int)steps; n += 3) {
            CW_STEP(n, r2, ha, hb, hc, ia, ib, ic, la, lc)
            CW_STEP(n + 1, r0, hb, hc, ha, ib, ic, ia, lb, la)
            CW_STEP(n + 2, r1, hc, ha, hb, ic, ia, ib, lc, lb)
        }
#undef CW_STEP
        if (p.do_ssim) {
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const double s = warp_sum(writes ? m5[k] : 0.0);
                if (lane == 0) p.partials[tile * 5 + k] = s;
            }
        }
    }
}


// The blur alone at r = 3 (norm_cuda_blur, the blur effect of norm_cuda_effects; image_shader.wgsl:88-103 blurs all four
// channels) on the same register windows: no Sobel, no moments, so a step is unpack -> H pass -> V pass -> quantise ->
// one 16-byte store, and the kernel runs at the memory system's pace instead of the pipelined kernel's hand-offs.
// Channels travel as pairs (r, g) and (b, a) so that every tap is one FFMA2; after grayscale the pairs are (y, a) only.
template <bool GRAY, int FRONT>  // FRONT: 0 no front stage, 1 with the 256-entry table, 2 grayscale / inversion without a table
__global__ void __launch_bounds__(kCwThreads, 2) blur_win3_kernel(const __grid_constant__ FusedParams p, uint32_t tiles_x, uint32_t tiles_y,
                                                                  uint32_t tile_rows) {
    constexpr int NP = GRAY ? 1 : 2;                       // channel pairs: (y, a) after grayscale, else (r, g) and (b, a)
    constexpr int FR = (FRONT == 1 ? 1 : 0) | (GRAY ? 2 : 0);
    __shared__ uint8_t slut[256];
    for (int i = threadIdx.x; i < 256; i += kCwThreads) slut[i] = p.lut.v[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int W = (int)p.in.w, H = (int)p.in.h;
    const float t0 = p.taps.t[0], t1 = p.taps.t[1], t2 = p.taps.t[2];
    const float2 t0p = make_float2(t0, t0), t1p = make_float2(t1, t1), t2p = make_float2(t2, t2);
    const uint32_t inv_mask = p.invert ? 0x00ffffffu : 0u;
    const uint32_t tiles_per_frame = tiles_x * tiles_y;
    const size_t total = (size_t)tiles_per_frame * p.in.frames;
    const uint32_t ipitch = (uint32_t)p.in.pitch, opitch = (uint32_t)p.out.pitch;
    struct Raw {
        uint32_t p[4];
    };
    struct HRow {
        float2 v[4][NP];  // H-pass outputs of this lane's four columns, per channel pair
    };
    const uint32_t steps = (tile_rows + 2) / 3 * 3;  // the same for every tile: the shuffles sit in uniform control flow

    for (size_t tile = (size_t)blockIdx.x * kCwWarps + (threadIdx.x >> 5); tile < total; tile += (size_t)gridDim.x * kCwWarps) {
        const uint32_t f = (uint32_t)(tile / tiles_per_frame);
        const uint32_t tr = (uint32_t)(tile - (size_t)f * tiles_per_frame);
        const int x0 = (int)(tr % tiles_x) * kCwCols;
        const int ya = (int)p.out_y0 + (int)(tr / tiles_x) * (int)tile_rows, yb = min(ya + (int)tile_rows, (int)p.out_y1);
        const int col = x0 - 4 + 4 * lane;
        const bool vec = col >= 0 && col + 3 < W;
        const bool writes = lane >= 1 && lane <= 30 && col < W;
        const int colc = clampi(col, 0, ((W - 1) >> 2) << 2);
        const bool border_tile = x0 == 0 || x0 + kCwCols + 4 > W;
        const uint8_t *lbase = p.in.ptr + (size_t)f * p.in.frame_stride + (size_t)colc * 4;
        uint8_t *obase = p.out.ptr + (size_t)f * p.out.frame_stride;
        // step n: source row q = n + 2 of the clamped sequence row(q) = clamp(ya - 1 + q) arrives, output row ya + n leaves
        auto fetch = [&](int q) {
            const uint4 v = ldg_stream(reinterpret_cast<const uint4 *>(lbase + (uint64_t)(uint32_t)clampi(ya - 1 + q, 0, H - 1) * ipitch));
            Raw r;
            r.p[0] = v.x, r.p[1] = v.y, r.p[2] = v.z, r.p[3] = v.w;
            return r;
        };
        auto hrow = [&](const Raw &rin, HRow &h) {
            Raw q = rin;
            if (border_tile && !vec) {  // ClampToEdge on the source columns
                const int le = (W - 1) & 3;
                const uint32_t last = le == 0 ? q.p[0] : le == 1 ? q.p[1] : le == 2 ? q.p[2] : q.p[3];
                const uint32_t edge = col < 0 ? q.p[0] : last;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (col + j < 0 || col + j > W - 1) q.p[j] = edge;
            }
            float2 v[6][NP];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t px = FRONT == 0 ? q.p[j] : s3_front<FR>(q.p[j], slut, inv_mask);
                if (GRAY) {
                    v[1 + j][0] = make_float2(byte_f(px, 0), byte_f(px, 3));
                } else {
                    v[1 + j][0] = make_float2(byte_f(px, 0), byte_f(px, 1));
                    v[1 + j][NP - 1] = make_float2(byte_f(px, 2), byte_f(px, 3));
                }
            }
#pragma unroll
            for (int c = 0; c < NP; ++c) {
                v[0][c].x = __shfl_up_sync(0xffffffffu, v[4][c].x, 1), v[0][c].y = __shfl_up_sync(0xffffffffu, v[4][c].y, 1);
                v[5][c].x = __shfl_down_sync(0xffffffffu, v[1][c].x, 1), v[5][c].y = __shfl_down_sync(0xffffffffu, v[1][c].y, 1);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    h.v[j][c] = __ffma2_rn(t2p, v[j + 2][c], __ffma2_rn(t1p, v[j + 1][c], __fmul2_rn(t0p, v[j][c])));
            }
        };
        Raw r0 = fetch(0), r1 = fetch(1), r2 = fetch(2);
        HRow ha, hb, hc;
        hrow(r0, ha);
        r0 = fetch(3);
        hrow(r1, hb);
        r1 = fetch(4);
#define BW_STEP(N_, RC, HA, HB, HC)                                                                               \
    {                                                                                                             \
        const int n_ = (N_), y_ = ya + n_;                                                                        \
        hrow(RC, HC);                                                                                             \
        RC = fetch(n_ + 5);                                                                                       \
        uint32_t o[4];                                                                                            \
        _Pragma("unroll") for (int j = 0; j < 4; ++j) {                                                           \
            uint32_t qb[NP][2];                                                                                   \
            _Pragma("unroll") for (int c = 0; c < NP; ++c) {                                                      \
                const float2 b = __ffma2_rn(t2p, HC.v[j][c], __ffma2_rn(t1p, HB.v[j][c], __fmul2_rn(t0p, HA.v[j][c]))); \
                qb[c][0] = __float_as_uint(__fadd_rn(b.x, 12582912.0f));                                          \
                qb[c][1] = __float_as_uint(__fadd_rn(b.y, 12582912.0f));                                          \
            }                                                                                                     \
            if (GRAY) { /* (y, y, y, a) */                                                                        \
                o[j] = __byte_perm(qb[0][0], qb[0][1], 0x4000);                                                   \
            } else {   /* low bytes of the four biased sums -> (r, g, b, a) */                                    \
                o[j] = __byte_perm(__byte_perm(qb[0][0], qb[0][1], 0x7740), __byte_perm(qb[NP - 1][0], qb[NP - 1][1], 0x7740), 0x5410); \
            }                                                                                                     \
        }                                                                                                         \
        if (y_ < yb && writes) {                                                                                  \
            uint8_t *orow = obase + (uint64_t)(uint32_t)y_ * opitch;                                              \
            if (vec) {                                                                                            \
                stg_stream(reinterpret_cast<uint4 *>(orow + (size_t)col * 4), make_uint4(o[0], o[1], o[2], o[3])); \
            } else {                                                                                              \
                _Pragma("unroll") for (int j = 0; j < 4; ++j) if (col + j < W)                                    \
                        reinterpret_cast<uint32_t *>(orow)[col + j] = o[j];                                       \
            }                                                                                                     \
        }                                                                                                         \
    }
#pragma unroll 1
        for (int n = 0; n < (int)steps; n += 3) {
            BW_STEP(n, r2, ha, hb, hc)
            BW_STEP(n + 1, r0, hb, hc, ha)
            BW_STEP(n + 2, r1, hc, ha, hb)
        }
#undef BW_STEP
    }
}

}  // namespace normk
