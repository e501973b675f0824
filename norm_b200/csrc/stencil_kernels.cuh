// 3x3 stencils on a register window: Sobel line detection (image_shader.wgsl:132-156), sharpen
// (:120-129) and the box blur (:105-118), each optionally with the front stage (sRGB/gamma table,
// grayscale, invert) applied to the pixels as they are loaded, so a reference frame "compute chain ->
// one fragment effect" is one pass over HBM (4 B read + 4 B written per pixel).
//
// One warp owns a tile of 120 x 36 output pixels; lane i loads columns x0-4+4i .. +3 (lanes 0 and 31 only
// feed their neighbours' halo) and marches down the rows with a window of 3 rows x (1 + 4 + 1) elements
// in registers.  Per row a lane issues ONE 16-byte load, converts its four pixels once (Sobel: intensity)
// and gets the two halo elements from its neighbours with warp shuffles: no divergence away from the
// image border, no shared memory (except the 256-byte table), no barriers.  Arithmetic is the S-form of oracle/norm_oracle.c.
#pragma once
#include "chain_kernels.cuh"

namespace normk {

constexpr int kS3Rows = 36;   // rows per warp tile (38 rows loaded: 5.5 % halo)
constexpr int kS3Cols = 120;  // output columns per warp tile: lanes 1..30 x 4 pixels
enum { kOpSobel = 0, kOpSharpen = 1, kOpBox = 2 };

// FR bit 0: 256-entry table on r, g, b.  FR bit 1: grayscale.  inv_mask = 0x00ffffff inverts r, g, b.
template <int FR>
__device__ __forceinline__ uint32_t s3_front(uint32_t p, const uint8_t *slut, uint32_t inv_mask) {
    if (FR & 2) {
        float r, g, b;
        if (FR & 1) {
            r = (float)slut[p & 0xffu], g = (float)slut[(p >> 8) & 0xffu], b = (float)slut[(p >> 16) & 0xffu];
        } else {
            r = byte_f_alu(p, 0), g = byte_f_alu(p, 1), b = byte_f_alu(p, 2);
        }
        const uint32_t y = __float_as_uint(__fadd_rn(luma_fused(r, g, b), 12582912.0f)) & 0xffu;
        p = y * 0x010101u | (p & 0xff000000u);
    } else if (FR & 1) {
        p = pack_rgba(slut[p & 0xffu], slut[(p >> 8) & 0xffu], slut[(p >> 16) & 0xffu], p >> 24);
    }
    return p ^ inv_mask;
}

// window element: Sobel keeps the intensity, sharpen the pixel, the box blur the pixel split into its
// even and odd bytes (two 16-bit lanes per word, so byte sums add up without carries between channels) and
// summed with its left and right neighbours
template <int OP>
struct S3Elem {
    uint32_t a, b;
};

// sqrt(r^2 + g^2 + b^2) with an exact integer radicand (image_shader.wgsl:132-135 in raw scale).  A zero
// radicand is lifted to 1e-30: its root 1e-15 is absorbed by every sum that contains a non-zero
// intensity (those are >= 1) and gives a magnitude that rounds to 0 when all nine are zero, so the
// bytes written equal those of the exact 0 -- and the compare / select of the zero case is not needed.
__device__ __forceinline__ float s3_intensity(uint32_t p) {
    // r^2 + g^2 + b^2 is an integer < 2^18: one four-way byte dot product of the pixel with itself (alpha masked) and
    // one exact conversion give the same float as the three unpacks and three multiply-adds
    const uint32_t rgb = p & 0x00ffffffu;
    const float a = fmaxf((float)__dp4a(rgb, rgb, 0u), 1e-30f);
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(a));
    float s = __fmul_rn(a, y);
    const float h = __fmul_rn(y, 0.5f);
    const float e = __fmaf_rn(-s, s, a);
    return __fmaf_rn(e, h, s);
}

template <int OP>
__device__ __forceinline__ S3Elem<OP> s3_conv(uint32_t p) {
    S3Elem<OP> e;
    if (OP == kOpSobel) e.a = __float_as_uint(s3_intensity(p)), e.b = 0;
    else if (OP == kOpSharpen) e.a = p, e.b = 0;
    else e.a = p & 0x00ff00ffu, e.b = (p >> 8) & 0x00ff00ffu;
    return e;
}

template <int OP>
__device__ __forceinline__ S3Elem<OP> s3_shfl_up(const S3Elem<OP> &e) {
    S3Elem<OP> r;
    r.a = __shfl_up_sync(0xffffffffu, e.a, 1);
    r.b = OP == kOpBox ? __shfl_up_sync(0xffffffffu, e.b, 1) : 0u;
    return r;
}
template <int OP>
__device__ __forceinline__ S3Elem<OP> s3_shfl_down(const S3Elem<OP> &e) {
    S3Elem<OP> r;
    r.a = __shfl_down_sync(0xffffffffu, e.a, 1);
    r.b = OP == kOpBox ? __shfl_down_sync(0xffffffffu, e.b, 1) : 0u;
    return r;
}

// (sum + 4) / 9 for sum <= 9 * 255: round-to-nearest of sum / 9 (no ties: 9 is odd), which is what the
// S-form's single f32 division followed by round-half-even yields
__device__ __forceinline__ uint32_t div9_round(uint32_t s) { return ((s + 4u) * 7282u) >> 16; }

template <int OP>
__device__ __forceinline__ uint32_t s3_apply(const S3Elem<OP> (&top)[6], const S3Elem<OP> (&mid)[6], const S3Elem<OP> (&bot)[6], int j,
                                             int factor) {
    if (OP == kOpSobel) {
        const uint32_t cb = sobel_mag_bits(__uint_as_float(top[j].a), __uint_as_float(top[j + 1].a), __uint_as_float(top[j + 2].a),
                                           __uint_as_float(mid[j].a), __uint_as_float(mid[j + 2].a), __uint_as_float(bot[j].a),
                                           __uint_as_float(bot[j + 1].a), __uint_as_float(bot[j + 2].a));
        return __byte_perm(cb, 0xffu, 0x4000);
    } else if (OP == kOpSharpen) {
        const uint32_t c = mid[j + 1].a, u = top[j + 1].a, d = bot[j + 1].a, l = mid[j].a, r = mid[j + 2].a;
        // neighbour sums per channel in 16-bit lanes (<= 1020), then (1 + 4f) c - f s per channel
        const uint32_t se = (u & 0x00ff00ffu) + (d & 0x00ff00ffu) + (l & 0x00ff00ffu) + (r & 0x00ff00ffu);
        const uint32_t so = ((u >> 8) & 0x00ff00ffu) + ((d >> 8) & 0x00ff00ffu) + ((l >> 8) & 0x00ff00ffu) + ((r >> 8) & 0x00ff00ffu);
        const int k1 = 1 + 4 * factor;
        const int v0 = k1 * (int)(c & 0xffu) - factor * (int)(se & 0xffffu);
        const int v1 = k1 * (int)((c >> 8) & 0xffu) - factor * (int)(so & 0xffffu);
        const int v2 = k1 * (int)((c >> 16) & 0xffu) - factor * (int)(se >> 16);
        const int v3 = k1 * (int)(c >> 24) - factor * (int)(so >> 16);
        // max(min(v, 255), 0) is one instruction each (VIMNMX.RELU)
        return pack_rgba((uint32_t)__vimin_s32_relu(v0, 255), (uint32_t)__vimin_s32_relu(v1, 255),
                         (uint32_t)__vimin_s32_relu(v2, 255), (uint32_t)__vimin_s32_relu(v3, 255));
    } else {
        // window elements of the box blur are HORIZONTAL three-pixel sums (see convert): three adds per lane pair
        const uint32_t se = top[j + 1].a + mid[j + 1].a + bot[j + 1].a;
        const uint32_t so = top[j + 1].b + mid[j + 1].b + bot[j + 1].b;
        return pack_rgba(div9_round(se & 0xffffu), div9_round(so & 0xffffu), div9_round(se >> 16), div9_round(so >> 16));
    }
}

template <int OP, int FR>
__global__ void __launch_bounds__(256, 4) stencil3_kernel(ImgView in, ImgView out, const __grid_constant__ Lut256 lut,
                                                       uint32_t inv_mask, int factor, uint32_t tiles_x, uint32_t tiles_y) {
    __shared__ uint8_t slut[256];
    if (FR & 1) {
        slut[threadIdx.x & 255] = lut.v[threadIdx.x & 255];
        __syncthreads();
    }
    const int lane = threadIdx.x & 31;
    const int W = (int)in.w, H = (int)in.h;
    const uint32_t tiles_per_frame = tiles_x * tiles_y;
    const size_t total = (size_t)tiles_per_frame * in.frames;
    const size_t wstride = (size_t)gridDim.x * 8;
    for (size_t tile = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5); tile < total; tile += wstride) {
        const uint32_t f = (uint32_t)(tile / tiles_per_frame);
        const uint32_t tr = (uint32_t)(tile - (size_t)f * tiles_per_frame);
        const int x0 = (int)(tr % tiles_x) * kS3Cols, y0 = (int)(tr / tiles_x) * kS3Rows;
        // lanes 0 and 31 only feed their neighbours' halo: lane i loads columns x0 - 4 + 4i .. + 3
        const int col = x0 - 4 + 4 * lane;
        const bool vec = col >= 0 && col + 3 < W;                     // the whole 16-byte access lies inside the row
        const bool writes = lane >= 1 && lane <= 30 && col < W;
        const uint8_t *fbase = in.ptr + (size_t)f * in.frame_stride;
        const uint32_t ipitch = (uint32_t)in.pitch;
        // Every lane loads ONE aligned 16-byte chunk: its own, or - for lanes left / right of the image - the first /
        // last chunk of the row (rows are pitched to 16 bytes, so the chunk that holds pixel W-1 is readable to its
        // end).  ClampToEdge is then a register fix-up that only tiles on the image border execute (tile-uniform
        // branch); interior tiles pay one address computation and one load per row, nothing else.
        const int colc = clampi(col, 0, ((W - 1) >> 2) << 2);
        const bool border_tile = x0 == 0 || x0 + kS3Cols + 4 > W;
        const uint8_t *lbase = fbase + (size_t)colc * 4;

        // fetch(r): this lane's 16 bytes of row r (ClampToEdge).  Fetches run two rows ahead of the arithmetic
        // so that every lane keeps two loads in flight.
        struct Raw {
            uint32_t p[4];
        };
        auto fetch = [&](int r) {
            const uint4 v = ldg_stream(reinterpret_cast<const uint4 *>(lbase + (uint64_t)(uint32_t)clampi(r, 0, H - 1) * ipitch));
            Raw q;
            q.p[0] = v.x, q.p[1] = v.y, q.p[2] = v.z, q.p[3] = v.w;
            return q;
        };
        auto convert = [&](const Raw &qin, S3Elem<OP>(&w)[6]) {
            Raw q = qin;
            // (ClampToEdge fix-up here, where the data is consumed, not in fetch: touching the registers there would
            // make every load wait for its data at once and undo the two-rows-ahead prefetch)
            if (border_tile && !vec) {
                // element of the loaded chunk that holds the clamped pixel: column 0 left of the image, column W-1 otherwise
                const int le = (W - 1) & 3;
                const uint32_t last = le == 0 ? q.p[0] : le == 1 ? q.p[1] : le == 2 ? q.p[2] : q.p[3];
                const uint32_t edge = col < 0 ? q.p[0] : last;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (col + j < 0 || col + j > W - 1) q.p[j] = edge;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) w[1 + j] = s3_conv<OP>(s3_front<FR>(q.p[j], slut, inv_mask));
            w[0] = s3_shfl_up<OP>(w[4]);
            w[5] = s3_shfl_down<OP>(w[1]);
            if (OP == kOpBox) {
                // the nine bytes of a channel sum exactly in integers whatever the order: keep, per row, the sum of each
                // output column's three horizontal neighbours (two thirds of the window registers, fewer adds per pixel)
                S3Elem<OP> h[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) h[j].a = w[j].a + w[j + 1].a + w[j + 2].a, h[j].b = w[j].b + w[j + 1].b + w[j + 2].b;
#pragma unroll
                for (int j = 0; j < 4; ++j) w[1 + j] = h[j];
            }
        };

        uint8_t *obase = out.ptr + (size_t)f * out.frame_stride;
        const uint32_t opitch = (uint32_t)out.pitch;
        S3Elem<OP> wa[6], wb[6], wc[6];  // rows r-1, r, r+1 in rotating roles
        {
            const Raw qa = fetch(y0 - 1), qb = fetch(y0);
            convert(qa, wa);
            convert(qb, wb);
        }
        Raw q0 = fetch(y0 + 1), q1 = fetch(y0 + 2), q2 = fetch(y0 + 3);  // rows r+1 .. r+3 of the row being computed: three loads in flight
        // "top" is row r+1: texture v grows downward (image_shader.wgsl:136-146)
#define S3_ROW(R_, BOT, MID, TOP, QCUR, QNEXT)                                                            \
    if ((R_) < H) {                                                                                       \
        convert(QCUR, TOP);                                                                               \
        QCUR = fetch((R_) + 4);                                                                           \
        uint32_t o[4];                                                                                    \
        _Pragma("unroll") for (int j = 0; j < 4; ++j) o[j] = s3_apply<OP>(TOP, MID, BOT, j, factor);      \
        uint8_t *orow = obase + (uint64_t)(uint32_t)(R_) * opitch;                                        \
        if (writes) {                                                                                     \
            if (vec) {                                                                                    \
                stg_stream(reinterpret_cast<uint4 *>(orow + (size_t)col * 4), make_uint4(o[0], o[1], o[2], o[3])); \
            } else {                                                                                      \
                _Pragma("unroll") for (int j = 0; j < 4; ++j) if (col + j < W)                            \
                        reinterpret_cast<uint32_t *>(orow)[col + j] = o[j];                               \
            }                                                                                             \
        }                                                                                                 \
    }
        // 3 rows per trip: the window and the fetch queue both rotate with period 3
        static_assert(kS3Rows % 3 == 0, "row loop below is unrolled 3 x");
#pragma unroll 1
        for (int r = y0; r < y0 + kS3Rows; r += 3) {
            S3_ROW(r, wa, wb, wc, q0, q1)
            S3_ROW(r + 1, wb, wc, wa, q1, q2)
            S3_ROW(r + 2, wc, wa, wb, q2, q0)
        }
#undef S3_ROW
    }
}

}  // namespace normk
