// Shared device-side definitions for the norm_b200 CUDA kernels (sm_100a).
//
// Numerical contract (DESIGN.md "S-form"): every kernel reproduces, bit for bit, the
// fixed-order raw-scale f32 arithmetic that oracle/norm_oracle.c states for the reference
// shaders.  All roundings are explicit (__f*_rn intrinsics); the library is built with
// -fmad=false and without fast-math so nothing is contracted or reassociated behind our back.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace normk {

// luma weights: grayscale_compute.wgsl:25, ssim.rs:73-75
constexpr float kLumaR = 0.29891f, kLumaG = 0.58661f, kLumaB = 0.11448f;

constexpr int kMaxRadius = 63;  // NORM_MAX_BLUR_RADIUS

// A batch of pitched RGBA8 frames in device memory.
struct ImgView {
    uint8_t *ptr;
    size_t pitch;         // bytes per row, % 16 == 0
    size_t frame_stride;  // bytes per frame, % 16 == 0
    uint32_t w, h, frames;
};

struct Taps {
    float t[kMaxRadius + 1];
    int32_t n;     // number of taps == blur radius
    int32_t dmin;  // texel offset of tap 0
};

struct Lut256 {
    uint8_t v[256];
};

__device__ __forceinline__ float byte_f(uint32_t p, int i) {
    return (float)((p >> (8 * i)) & 0xffu);  // exact
}

// raw-scale float -> u8: clamp, round half to even (== oracle quant_raw)
__device__ __forceinline__ uint32_t quant_u8(float v) {
    v = fminf(fmaxf(v, 0.0f), 255.0f);
    return (uint32_t)__float2int_rn(v);
}

__device__ __forceinline__ uint32_t pack_rgba(uint32_t r, uint32_t g, uint32_t b, uint32_t a) {
    return r | (g << 8) | (b << 16) | (a << 24);
}

// grayscale_compute.wgsl:25 in raw scale, fused in the oracle's fixed order
__device__ __forceinline__ float luma_fused(float r, float g, float b) {
    return __fmaf_rn(b, kLumaB, __fmaf_rn(g, kLumaG, __fmul_rn(r, kLumaR)));
}

// ssim.rs:73-75: f32, one rounding per operation, left to right (Rust never contracts)
__device__ __forceinline__ float luma_ssim(uint32_t p) {
    const float r = byte_f(p, 0), g = byte_f(p, 1), b = byte_f(p, 2);
    return __fadd_rn(__fadd_rn(__fmul_rn(r, kLumaR), __fmul_rn(g, kLumaG)), __fmul_rn(b, kLumaB));
}

// image_shader.wgsl:132-135 in raw scale: the radicand is an exact integer < 2^24
__device__ __forceinline__ float intensity_raw(uint32_t p) {
    const float r = byte_f(p, 0), g = byte_f(p, 1), b = byte_f(p, 2);
    return __fsqrt_rn(__fmaf_rn(b, b, __fmaf_rn(g, g, __fmul_rn(r, r))));
}

// image_shader.wgsl:148-155.  "top" is row +1 (texture v grows downward).  The sums are
// left-to-right; 2*v is exact so a*2+b has a single rounding whichever way it is issued.
__device__ __forceinline__ uint32_t sobel_mag(float tl, float t, float tr, float l, float r, float bl,
                                              float b, float br) {
    float gx = __fadd_rn(tl, __fmul_rn(2.0f, l));
    gx = __fadd_rn(gx, bl);
    gx = __fsub_rn(gx, tr);
    gx = __fsub_rn(gx, __fmul_rn(2.0f, r));
    gx = __fsub_rn(gx, br);
    float gy = __fsub_rn(-l, __fmul_rn(2.0f, t));
    gy = __fsub_rn(gy, tr);
    gy = __fadd_rn(gy, bl);
    gy = __fadd_rn(gy, __fmul_rn(2.0f, b));
    gy = __fadd_rn(gy, br);
    const float c = __fsqrt_rn(__fmaf_rn(gx, gx, __fmul_rn(gy, gy)));
    return quant_u8(c);
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

__device__ __forceinline__ const uint8_t *row_ptr(const ImgView &im, uint32_t frame, int y) {
    return im.ptr + (size_t)frame * im.frame_stride + (size_t)y * im.pitch;
}

// streaming 16-byte accesses (each byte of a frame is touched once per kernel)
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_stream(uint4 *p, const uint4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace normk
