// Separable Gaussian blur (gaussian_blur, image_shader.wgsl:83-103) -- tile kernel.
//
// One CTA (256 threads) produces a 64 x 32 output tile:
//   1. the (64+R-1) x (32+R-1) RGBA8 source tile is staged in shared memory with ClampToEdge
//      addressing (gpu_state.rs:614-616);
//   2. horizontal pass: each thread takes one row and 8 consecutive outputs and walks the
//      8+R-1 source pixels once, keeping the 8x4 accumulators in registers (register sliding
//      window: 1 LDS per (8+R-1)/8 pixels instead of 1 per tap);
//   3. vertical pass: each thread owns one (column, channel) and walks the 32+R-1 rows of
//      the f32 intermediate once with <= R live accumulators;
//   4. results are quantised, re-interleaved through shared memory and stored as 16-byte
//      vectors.
// The f32 intermediate never leaves shared memory and is never quantised.
// Accumulation order is tap 0..R-1 by explicit fmaf from 0, identical to oracle_blur_s.
#pragma once
#include "common.cuh"

namespace normk {

constexpr int kBlurTW = 64, kBlurTH = 32, kBlurThreads = 256, kBlurSeg = 8;

__host__ __device__ constexpr int blur_src_stride(int r) { return (kBlurTW + r - 1) | 1; }  // odd: conflict-free
constexpr int kBlurHStride = kBlurTW + 1;

__host__ __device__ constexpr size_t blur_smem_bytes(int r) {
    return (size_t)(kBlurTH + r - 1) * blur_src_stride(r) * 4 +           // source tile (u32)
           (size_t)4 * (kBlurTH + r - 1) * kBlurHStride * sizeof(float);  // H planes
}

// R > 0: compile-time radius (fully unrolled register windows).  R == 0: runtime radius.
template <int R>
__global__ void __launch_bounds__(kBlurThreads) blur_tile_kernel(ImgView in, ImgView out,
                                                                 const __grid_constant__ Taps taps,
                                                                 uint32_t tiles_x, uint32_t tiles_y) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const int r = R > 0 ? R : taps.n;
    const int dmin = taps.dmin;
    const int rows = kBlurTH + r - 1, cols = kBlurTW + r - 1;
    const int sstride = blur_src_stride(r);
    uint32_t *src = reinterpret_cast<uint32_t *>(smem_raw);
    float *H = reinterpret_cast<float *>(smem_raw + (size_t)rows * sstride * 4);
    const int hplane = rows * kBlurHStride;

    const uint32_t tiles_per_frame = tiles_x * tiles_y;
    const size_t total = (size_t)tiles_per_frame * in.frames;
    for (size_t tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const uint32_t f = (uint32_t)(tile / tiles_per_frame);
        const uint32_t trem = (uint32_t)(tile - (size_t)f * tiles_per_frame);
        const int x0 = (int)(trem % tiles_x) * kBlurTW, y0 = (int)(trem / tiles_x) * kBlurTH;

        __syncthreads();  // previous tile's output staging is done with `src`
        for (int i = threadIdx.x; i < rows * cols; i += kBlurThreads) {
            const int ry = i / cols, rx = i - ry * cols;
            const int gy = clampi(y0 + dmin + ry, 0, (int)in.h - 1);
            const int gx = clampi(x0 + dmin + rx, 0, (int)in.w - 1);
            src[ry * sstride + rx] = reinterpret_cast<const uint32_t *>(row_ptr(in, f, gy))[gx];
        }
        __syncthreads();

        // ---- horizontal pass: item = (row, 8-pixel segment); lanes walk consecutive rows
        for (int item = threadIdx.x; item < rows * (kBlurTW / kBlurSeg); item += kBlurThreads) {
            const int row = item % rows, seg = item / rows;
            const uint32_t *sp = src + row * sstride + seg * kBlurSeg;
            float acc[kBlurSeg][4];
#pragma unroll
            for (int o = 0; o < kBlurSeg; ++o) acc[o][0] = acc[o][1] = acc[o][2] = acc[o][3] = 0.0f;
            if (R > 0) {
#pragma unroll
                for (int j = 0; j < kBlurSeg + R - 1; ++j) {
                    const uint32_t p = sp[j];
                    const float c0 = byte_f(p, 0), c1 = byte_f(p, 1), c2 = byte_f(p, 2), c3 = byte_f(p, 3);
#pragma unroll
                    for (int o = 0; o < kBlurSeg; ++o) {
                        const int k = j - o;  // compile-time after unrolling
                        if (k >= 0 && k < R) {
                            const float t = taps.t[k];
                            acc[o][0] = __fmaf_rn(t, c0, acc[o][0]);
                            acc[o][1] = __fmaf_rn(t, c1, acc[o][1]);
                            acc[o][2] = __fmaf_rn(t, c2, acc[o][2]);
                            acc[o][3] = __fmaf_rn(t, c3, acc[o][3]);
                        }
                    }
                }
            } else {
                for (int k = 0; k < r; ++k) {
                    const float t = taps.t[k];
#pragma unroll
                    for (int o = 0; o < kBlurSeg; ++o) {
                        const uint32_t p = sp[o + k];
                        acc[o][0] = __fmaf_rn(t, byte_f(p, 0), acc[o][0]);
                        acc[o][1] = __fmaf_rn(t, byte_f(p, 1), acc[o][1]);
                        acc[o][2] = __fmaf_rn(t, byte_f(p, 2), acc[o][2]);
                        acc[o][3] = __fmaf_rn(t, byte_f(p, 3), acc[o][3]);
                    }
                }
            }
            float *hp = H + row * kBlurHStride + seg * kBlurSeg;
#pragma unroll
            for (int o = 0; o < kBlurSeg; ++o) {
                hp[o] = acc[o][0];
                hp[hplane + o] = acc[o][1];
                hp[2 * hplane + o] = acc[o][2];
                hp[3 * hplane + o] = acc[o][3];
            }
        }
        __syncthreads();

        // ---- vertical pass: thread = (column, channel); output staged as bytes over `src`
        {
            const int col = threadIdx.x & (kBlurTW - 1), ch = threadIdx.x >> 6;
            const float *hp = H + ch * hplane + col;
            uint8_t *stage = reinterpret_cast<uint8_t *>(src);
            if (R > 0) {
                float acc[kBlurTH];
#pragma unroll
                for (int o = 0; o < kBlurTH; ++o) acc[o] = 0.0f;
#pragma unroll
                for (int j = 0; j < kBlurTH + R - 1; ++j) {
                    const float h = hp[j * kBlurHStride];
#pragma unroll
                    for (int o = 0; o < kBlurTH; ++o) {
                        const int k = j - o;
                        if (k >= 0 && k < R) acc[o] = __fmaf_rn(taps.t[k], h, acc[o]);
                    }
                    if (j >= R - 1) {
                        const int o = j - (R - 1);
                        stage[(o * kBlurTW + col) * 4 + ch] = (uint8_t)quant_u8(acc[o]);
                    }
                }
            } else {
                for (int o = 0; o < kBlurTH; ++o) {
                    float a = 0.0f;
                    for (int k = 0; k < r; ++k) a = __fmaf_rn(taps.t[k], hp[(o + k) * kBlurHStride], a);
                    stage[(o * kBlurTW + col) * 4 + ch] = (uint8_t)quant_u8(a);
                }
            }
        }
        __syncthreads();

        // ---- 16-byte coalesced stores
        const uint4 *st = reinterpret_cast<const uint4 *>(src);
        for (int i = threadIdx.x; i < kBlurTH * (kBlurTW / 4); i += kBlurThreads) {
            const int ry = i / (kBlurTW / 4), cx = i - ry * (kBlurTW / 4);
            const int y = y0 + ry, x = x0 + 4 * cx;
            if (y < (int)in.h && x < (int)in.w)
                stg_stream(reinterpret_cast<uint4 *>(const_cast<uint8_t *>(row_ptr(out, f, y))) + (x >> 2), st[i]);
        }
    }
}

}  // namespace normk
