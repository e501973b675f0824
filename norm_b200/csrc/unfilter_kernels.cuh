// PNG scanline unfilter + Adam7 de-interlace on the GPU (the step right before
// pixel expansion).  src/png/scanline_reader.rs:40-168 (filters None/Sub/Up/Average/Paeth),
// src/png/interlace.rs:35-83 (pass geometry) -- with the PNG-spec behaviour where the reference
// is incomplete (its Adam7 path handles None/Sub only and mis-indexes, scanline_reader.rs:201,220).
//
// A filtered byte depends on its left neighbour (bpp bytes back), the byte above and the byte
// above-left, so rows cannot be done independently and bytes within a row cannot either; what is
// parallel is the anti-diagonal wavefront: row y may process pixel k once row y-1 has processed
// pixel k.  One CTA handles one (sub-)image: thread r owns row r of the current block of rows and
// all threads step through time together, thread r being r pixels behind thread r-1.  Each step a
// thread reads the pixel its upper neighbour published in the previous step from a double-buffered
// shared-memory slot, reconstructs its own pixel (left / up-left live in registers), publishes it
// and writes it out.  Integer arithmetic, bit-exact against the oracle and the reference's blobs.
#pragma once
#include "common.cuh"

namespace normk {

constexpr int kUnfThreads = 512;
constexpr int kUnfMaxBpp = 8;

struct UnfilterPass {
    const uint8_t *src;   // filtered stream of this (sub-)image: rows of (1 + row_bytes) bytes
    uint8_t *dst;         // unfiltered rows, row_bytes each
    uint32_t rows, row_bytes, bpp;
};

struct UnfilterParams {
    UnfilterPass pass[7];
    int32_t npass;
    int32_t *status;  // set to 1 if a filter-type byte is > 4
};

__device__ __forceinline__ uint32_t paeth_predict(uint32_t a, uint32_t b, uint32_t c) {
    // scanline_reader.rs:149-167
    const int p = (int)a + (int)b - (int)c;
    const int pa = abs(p - (int)a), pb = abs(p - (int)b), pc = abs(p - (int)c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

__global__ void __launch_bounds__(kUnfThreads) unfilter_kernel(const __grid_constant__ UnfilterParams p) {
    __shared__ uint8_t slot[2][kUnfThreads][kUnfMaxBpp];
    const UnfilterPass ps = p.pass[blockIdx.x];
    if (ps.rows == 0 || ps.row_bytes == 0) return;
    const int t = threadIdx.x;
    const uint32_t bpp = ps.bpp, npx = (ps.row_bytes + bpp - 1) / bpp;  // "pixels" = groups of bpp bytes
    const size_t stride = (size_t)ps.row_bytes + 1;

    for (uint32_t y0 = 0; y0 < ps.rows; y0 += kUnfThreads) {
        const uint32_t y = y0 + t;
        const bool live = y < ps.rows;
        const uint8_t *row = ps.src + (size_t)(live ? y : 0) * stride;
        uint8_t *out = ps.dst + (size_t)(live ? y : 0) * ps.row_bytes;
        const uint8_t *up_row = (live && y > 0) ? ps.dst + (size_t)(y - 1) * ps.row_bytes : nullptr;  // thread 0 of a later block
        const uint32_t ft = live ? row[0] : 0u;
        if (live && ft > 4u) atomicExch(p.status, 1);
        uint8_t left[kUnfMaxBpp], upleft[kUnfMaxBpp];
#pragma unroll
        for (int i = 0; i < kUnfMaxBpp; ++i) left[i] = upleft[i] = 0;
        const uint32_t nsteps = npx + min((uint32_t)kUnfThreads, ps.rows - y0) - 1;
        // filtered bytes (and, for thread 0 of a later block, the row above) of the pixel handled in the
        // NEXT step are fetched one step ahead so the global-memory latency overlaps the barrier
        uint8_t nx[kUnfMaxBpp], nu[kUnfMaxBpp];
        auto fetch = [&](int k) {
#pragma unroll
            for (int i = 0; i < kUnfMaxBpp; ++i) {
                nx[i] = nu[i] = 0;
                const uint32_t bi = (uint32_t)k * bpp + i;
                if (live && k >= 0 && k < (int)npx && (uint32_t)i < bpp && bi < ps.row_bytes) {
                    nx[i] = row[1 + bi];
                    if (t == 0 && up_row) nu[i] = up_row[bi];
                }
            }
        };
        __syncthreads();  // the previous block of rows is completely written (thread 0 reads it as "up")
        fetch(0 - t);
        for (uint32_t s = 0; s < nsteps; ++s) {
            const int k = (int)s - t;  // pixel this thread handles in this step
            uint8_t cx[kUnfMaxBpp], cu[kUnfMaxBpp];
#pragma unroll
            for (int i = 0; i < kUnfMaxBpp; ++i) cx[i] = nx[i], cu[i] = nu[i];
            fetch(k + 1);
            if (live && k >= 0 && k < (int)npx) {
                const uint32_t b0 = (uint32_t)k * bpp, nb = min(bpp, ps.row_bytes - b0);
#pragma unroll
                for (int i = 0; i < kUnfMaxBpp; ++i) {
                    if ((uint32_t)i < nb) {
                        // "up": published by row y-1 in step s-1, or the previous block's last row
                        const uint32_t b = t > 0 ? slot[(s + 1) & 1][t - 1][i] : cu[i];
                        const uint32_t x = cx[i], a = left[i], c = upleft[i];
                        uint32_t pred = 0;
                        switch (ft) {
                        case 1: pred = a; break;
                        case 2: pred = b; break;
                        case 3: pred = (a + b) >> 1; break;
                        case 4: pred = paeth_predict(a, b, c); break;
                        default: break;
                        }
                        const uint8_t v = (uint8_t)(x + pred);
                        out[b0 + i] = v;
                        slot[s & 1][t][i] = v;
                        left[i] = v;
                        upleft[i] = (uint8_t)b;
                    }
                }
            }
            __syncthreads();
        }
    }
}

// Adam7: gathers every pixel of the raster image from its pass (interlace.rs:35-83 geometry).
struct DeinterlaceParams {
    const uint8_t *pass_px[7];  // unfiltered pass images
    uint32_t pass_row_bytes[7];
    uint8_t *dst;               // raster rows
    uint32_t w, h, row_bytes, bits;  // bits per pixel
};

// the Adam7 pass whose lattice contains pixel (x, y); the seven lattices partition the image
__device__ __forceinline__ int adam7_pass_of(uint32_t x, uint32_t y, uint32_t &px, uint32_t &py) {
    const uint32_t x0[7] = {0, 4, 0, 2, 0, 1, 0}, y0[7] = {0, 0, 4, 0, 2, 0, 1};
    const uint32_t dx[7] = {8, 8, 4, 4, 2, 2, 1}, dy[7] = {8, 8, 8, 4, 4, 2, 2};
    int ps = 6;
#pragma unroll
    for (int q = 0; q < 6; ++q)
        if (x >= x0[q] && y >= y0[q] && (x - x0[q]) % dx[q] == 0 && (y - y0[q]) % dy[q] == 0) {
            ps = q;
            break;
        }
    px = (x - x0[ps]) / dx[ps], py = (y - y0[ps]) / dy[ps];
    return ps;
}

__global__ void __launch_bounds__(256) deinterlace_kernel(const __grid_constant__ DeinterlaceParams p) {
    const size_t total = (size_t)p.row_bytes * p.h;
    for (size_t it = (size_t)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (size_t)gridDim.x * blockDim.x) {
        const uint32_t y = (uint32_t)(it / p.row_bytes), bx = (uint32_t)(it - (size_t)y * p.row_bytes);
        uint32_t outb = 0, px, py;
        if (p.bits >= 8) {
            const uint32_t bytes = p.bits / 8, x = bx / bytes, off = bx - x * bytes;
            const int ps = adam7_pass_of(x, y, px, py);
            outb = p.pass_px[ps][(size_t)py * p.pass_row_bytes[ps] + (size_t)px * bytes + off];
        } else {
            const uint32_t per = 8 / p.bits, mask = (1u << p.bits) - 1u;
            for (uint32_t i = 0; i < per; ++i) {
                const uint32_t x = bx * per + i;
                if (x >= p.w) break;
                const int ps = adam7_pass_of(x, y, px, py);
                const uint32_t bit = px * p.bits;
                const uint32_t v = (p.pass_px[ps][(size_t)py * p.pass_row_bytes[ps] + (bit >> 3)] >> (8 - p.bits - (bit & 7))) & mask;
                outb |= v << (8 - p.bits - i * p.bits);
            }
        }
        p.dst[it] = (uint8_t)outb;
    }
}

}  // namespace normk
