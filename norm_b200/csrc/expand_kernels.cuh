// Pixel expansion: Png::rgba8() (src/png/grammar.rs:123-155) + the PNG-spec modes for every colour type / bit depth
// of PngSuite.  One kernel template over {colour type, bit depth, mode}: no run-time switch in the pixel path.
//
// Work item = (frame, row, 4 output pixels) = ONE 16-byte store; consecutive lanes take consecutive items, so a warp
// writes 512 contiguous bytes and reads the matching contiguous span of packed samples (48 .. 384 bytes for the byte
// and two-byte types, one to four bytes per lane below 8 bits per sample) with the widest aligned load the format
// allows.  The packed rows are kept at a 16-byte pitch by the host side, so every such load is aligned.  The item walk
// is incremental (no division in the loop), two items are in flight per thread, and the frames of a batch are part
// of the grid, so a batch is one launch.  Bytes are shuffled with PRMT; nothing is converted.
#pragma once
#include "common.cuh"

namespace normk {

struct ExpandParams {
    const uint8_t *src;       // packed scanlines (device), rows at src_pitch, frames at src_frame_stride
    size_t src_pitch;         // multiple of 16
    size_t src_frame_stride;
    uint32_t w, h, frames, frame0;  // frame f of the batch goes to frame frame0 + f of `out`
    int32_t n_plte, n_trns;
    uint32_t key[3];          // tRNS colour key at native depth (gray / RGB, spec mode)
    int32_t has_key;
    uint8_t plte[768];
    uint8_t trns[256];
};

// 16 -> 8 bits: the image crate's to_rgb8 convention (v + 128) / 257 (decoder.rs:267-278 compares against it)
__device__ __forceinline__ uint32_t to8_16(uint32_t v) { return (v + 128u) / 257u; }

// big-endian 16-bit sample s (0 or 1) of a 32-bit word
__device__ __forceinline__ uint32_t be16(uint32_t word, int s) { return __byte_perm(word, 0u, s ? 0x4423 : 0x4401); }

template <int CT, int BD, int MODE>
struct ExpandTraits {
    static constexpr int CH = CT == 0 ? 1 : CT == 2 ? 3 : CT == 3 ? 1 : CT == 4 ? 2 : 4;
    static constexpr int BITS4 = 4 * CH * BD;            // bits of four pixels
    static constexpr int WORDS = (BITS4 + 31) / 32;      // 32-bit words loaded per item (>= 8 bits per sample)
};

// four pixels of one item; `words` = the item's packed samples (for BD < 8: the item's bits, MSB-first, in bits
// [BITS4-1 .. 0] of words[0]); pal = RGBA palette in shared memory
template <int CT, int BD, int MODE>
__device__ __forceinline__ uint4 expand_item(const uint32_t *words, const ExpandParams &p, const uint32_t *pal) {
    constexpr uint32_t A = MODE == 0 ? 0u : 0xff000000u;  // opaque alpha: 0 in compat mode (grammar.rs:130,139), 255 by the spec
    uint32_t o[4];
    if constexpr (BD < 8) {  // gray or palette, 1 / 2 / 4 bits per sample
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t v = (words[0] >> ((3 - j) * BD)) & ((1u << BD) - 1u);
            if constexpr (CT == 3) {
                o[j] = pal[v];
            } else {
                const uint32_t y = v * (BD == 1 ? 255u : BD == 2 ? 85u : 17u);
                o[j] = y * 0x010101u | ((p.has_key && v == p.key[0]) ? 0u : A);
            }
        }
    } else if constexpr (BD == 8) {
        if constexpr (CT == 6) {
            o[0] = words[0], o[1] = words[1], o[2] = words[2], o[3] = words[3];
        } else if constexpr (CT == 2) {
            // r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
            const uint32_t c0 = words[0] & 0x00ffffffu, c1 = __byte_perm(words[0], words[1], 0x4543) & 0x00ffffffu,
                           c2 = __byte_perm(words[1], words[2], 0x4432) & 0x00ffffffu, c3 = words[2] >> 8;
            const uint32_t c[4] = {c0, c1, c2, c3};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint32_t a = A;
                if (MODE == 1 && p.has_key && c[j] == (p.key[0] | (p.key[1] << 8) | (p.key[2] << 16))) a = 0u;
                o[j] = c[j] | a;
            }
        } else if constexpr (CT == 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t y = (words[0] >> (8 * j)) & 0xffu;
                o[j] = __byte_perm(words[0], 0u, 0x4000 + 0x111 * j) | ((MODE == 1 && p.has_key && y == p.key[0]) ? 0u : A);
            }
        } else if constexpr (CT == 4) {  // y0 a0 y1 a1 | y2 a2 y3 a3
            o[0] = __byte_perm(words[0], 0u, 0x1000), o[1] = __byte_perm(words[0], 0u, 0x3222);
            o[2] = __byte_perm(words[1], 0u, 0x1000), o[3] = __byte_perm(words[1], 0u, 0x3222);
        } else {  // palette, one index per byte
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = pal[(words[0] >> (8 * j)) & 0xffu];
        }
    } else {  // 16 bits per sample, big-endian; spec mode only
        constexpr int CH = ExpandTraits<CT, BD, MODE>::CH;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint32_t s[4];
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                const int idx = j * CH + c;  // sample index within the item
                s[c] = be16(words[idx / 2], idx & 1);
            }
            if constexpr (CT == 0) {
                o[j] = to8_16(s[0]) * 0x010101u | ((p.has_key && s[0] == p.key[0]) ? 0u : A);
            } else if constexpr (CT == 2) {
                const bool hit = p.has_key && s[0] == p.key[0] && s[1] == p.key[1] && s[2] == p.key[2];
                o[j] = pack_rgba(to8_16(s[0]), to8_16(s[1]), to8_16(s[2]), 0u) | (hit ? 0u : A);
            } else if constexpr (CT == 4) {
                o[j] = to8_16(s[0]) * 0x010101u | (to8_16(s[1]) << 24);
            } else {
                o[j] = pack_rgba(to8_16(s[0]), to8_16(s[1]), to8_16(s[2]), to8_16(s[3]));
            }
        }
    }
    return make_uint4(o[0], o[1], o[2], o[3]);
}

// the item's packed samples, with the widest aligned load the format allows
template <int CT, int BD, int MODE>
__device__ __forceinline__ void load_item(const uint8_t *row, uint32_t cx, uint32_t *words) {
    using T = ExpandTraits<CT, BD, MODE>;
    if constexpr (BD < 8) {
        if constexpr (BD == 4) words[0] = __byte_perm((uint32_t) * reinterpret_cast<const uint16_t *>(row + 2 * cx), 0u, 0x4401);  // two bytes, MSB-first
        else if constexpr (BD == 2) words[0] = row[cx];
        else words[0] = (row[cx >> 1] >> ((cx & 1u) ? 0 : 4)) & 0xfu;
    } else if constexpr (T::WORDS == 1) {
        words[0] = *reinterpret_cast<const uint32_t *>(row + 4 * cx);
    } else if constexpr (T::WORDS == 2) {
        const uint2 v = *reinterpret_cast<const uint2 *>(row + 8 * cx);
        words[0] = v.x, words[1] = v.y;
    } else if constexpr (T::WORDS == 3) {  // RGB8: 12 bytes, 4-byte aligned
        const uint32_t *q = reinterpret_cast<const uint32_t *>(row + 12 * cx);
        words[0] = q[0], words[1] = q[1], words[2] = q[2];
    } else if constexpr (T::WORDS == 4) {
        const uint4 v = ldg_stream(reinterpret_cast<const uint4 *>(row + 16 * cx));
        words[0] = v.x, words[1] = v.y, words[2] = v.z, words[3] = v.w;
    } else if constexpr (T::WORDS == 6) {  // RGB16: 24 bytes, 8-byte aligned
        const uint2 *q = reinterpret_cast<const uint2 *>(row + 24 * cx);
        const uint2 a = q[0], b = q[1], c = q[2];
        words[0] = a.x, words[1] = a.y, words[2] = b.x, words[3] = b.y, words[4] = c.x, words[5] = c.y;
    } else {  // RGBA16: 32 bytes
        static_assert(T::WORDS == 8, "unexpected item size");
        const uint4 a = ldg_stream(reinterpret_cast<const uint4 *>(row + 32 * cx)),
                    b = ldg_stream(reinterpret_cast<const uint4 *>(row + 32 * cx) + 1);
        words[0] = a.x, words[1] = a.y, words[2] = a.z, words[3] = a.w;
        words[4] = b.x, words[5] = b.y, words[6] = b.z, words[7] = b.w;
    }
}

template <int CT, int BD, int MODE>
__global__ void __launch_bounds__(256) expand_kernel(const __grid_constant__ ExpandParams p, ImgView out) {
    using T = ExpandTraits<CT, BD, MODE>;
    __shared__ uint32_t pal[CT == 3 ? 256 : 1];
    if constexpr (CT == 3) {  // PLTE + tRNS merged into one RGBA table (indices past the palette: black, alpha per tRNS)
        const int i = threadIdx.x;
        uint32_t c = 0u;
        if (i < p.n_plte) c = pack_rgba(p.plte[3 * i], p.plte[3 * i + 1], p.plte[3 * i + 2], 0u);
        pal[i] = c | ((i < p.n_trns ? (uint32_t)p.trns[i] : 255u) << 24);
        __syncthreads();
    }
    const uint32_t chunks = (p.w + 3u) / 4u;
    const size_t per_frame = (size_t)chunks * p.h;
    const size_t total = per_frame * p.frames;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    // (frame, row, chunk) of the next item, advanced incrementally: the stride is split once, the loop has no division
    const uint32_t df = (uint32_t)(stride / per_frame);
    const uint32_t dy = (uint32_t)((stride - (size_t)df * per_frame) / chunks);
    const uint32_t dc = (uint32_t)(stride - (size_t)df * per_frame - (size_t)dy * chunks);
    size_t it = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t f = (uint32_t)(it / per_frame);
    uint32_t y = (uint32_t)((it - (size_t)f * per_frame) / chunks);
    uint32_t cx = (uint32_t)(it - (size_t)f * per_frame - (size_t)y * chunks);
    auto step = [&]() {
        it += stride, f += df, y += dy, cx += dc;
        if (cx >= chunks) cx -= chunks, ++y;
        if (y >= p.h) y -= p.h, ++f;
    };
    auto src_row = [&](uint32_t ff, uint32_t yy) { return p.src + (size_t)ff * p.src_frame_stride + (size_t)yy * p.src_pitch; };
    auto emit = [&](const uint32_t *words, uint32_t ff, uint32_t yy, uint32_t cc) {
        uint4 o = expand_item<CT, BD, MODE>(words, p, pal);
        if (4u * cc + 3u >= p.w) {  // ragged tail: pixels past the width are zero (they live in the pitch padding)
            if (4u * cc + 1u >= p.w) o.y = 0u;
            if (4u * cc + 2u >= p.w) o.z = 0u;
            o.w = 0u;
        }
        stg_stream(reinterpret_cast<uint4 *>(const_cast<uint8_t *>(row_ptr(out, p.frame0 + ff, (int)yy))) + cc, o);
    };
    constexpr int NW = T::WORDS < 1 ? 1 : T::WORDS;
    while (it + stride < total) {  // two items in flight
        const uint32_t f0 = f, y0 = y, c0 = cx;
        uint32_t w0[NW], w1[NW];
        load_item<CT, BD, MODE>(src_row(f0, y0), c0, w0);
        step();
        const uint32_t f1 = f, y1 = y, c1 = cx;
        load_item<CT, BD, MODE>(src_row(f1, y1), c1, w1);
        step();
        emit(w0, f0, y0, c0);
        emit(w1, f1, y1, c1);
    }
    if (it < total) {
        uint32_t w0[NW];
        load_item<CT, BD, MODE>(src_row(f, y), cx, w0);
        emit(w0, f, y, cx);
    }
}

}  // namespace normk
