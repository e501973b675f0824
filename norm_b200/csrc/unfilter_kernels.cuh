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

struct UnfilterPass {
    const uint8_t *src;   // filtered stream of this (sub-)image: rows of (1 + row_bytes) bytes
    uint8_t *dst;         // rows at a 16-byte-aligned pitch; first filled with the filtered bytes, unfiltered in place
    uint8_t *ftype;       // one filter-type byte per row
    uint32_t rows, row_bytes, pitch, bpp;
};

// A batch holds `nimg` images of identical geometry: pass[] describes image 0 and image i sits
// src_stride / dst_stride / ft_stride bytes further in the three buffers.
struct UnfilterParams {
    UnfilterPass pass[7];
    int32_t npass;
    uint32_t nimg;
    size_t src_stride, dst_stride, ft_stride;
    int32_t *status;  // set to 1 if a filter-type byte is > 4
};

__device__ __forceinline__ UnfilterPass pass_of(const UnfilterParams &p, uint32_t pass, uint32_t img) {
    UnfilterPass ps = p.pass[pass];
    ps.src += (size_t)img * p.src_stride;
    ps.dst += (size_t)img * p.dst_stride;
    ps.ftype += (size_t)img * p.ft_stride;
    return ps;
}

// Step 1 (fully parallel): strip the filter-type bytes and re-align every row to a 16-byte pitch.
__global__ void __launch_bounds__(256) unfilter_align_kernel(const __grid_constant__ UnfilterParams p) {
    const UnfilterPass ps = pass_of(p, blockIdx.y, blockIdx.z);
    const uint32_t words = ps.pitch / 4;
    const size_t total = (size_t)words * ps.rows;
    for (size_t it = (size_t)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (size_t)gridDim.x * blockDim.x) {
        const uint32_t y = (uint32_t)(it / words), wx = (uint32_t)(it - (size_t)y * words);
        const uint8_t *row = ps.src + (size_t)y * ((size_t)ps.row_bytes + 1);
        uint32_t v = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t bx = 4 * wx + i;
            if (bx < ps.row_bytes) v |= (uint32_t)row[1 + bx] << (8 * i);
        }
        reinterpret_cast<uint32_t *>(ps.dst + (size_t)y * ps.pitch)[wx] = v;
        if (wx == 0) {
            ps.ftype[y] = row[0];
            if (row[0] > 4u) atomicExch(p.status, 1);
        }
    }
}

__device__ __forceinline__ uint32_t paeth_predict(uint32_t a, uint32_t b, uint32_t c) {
    // scanline_reader.rs:149-167
    const int p = (int)a + (int)b - (int)c;
    const int pa = abs(p - (int)a), pb = abs(p - (int)b), pc = abs(p - (int)c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

__device__ __forceinline__ uint32_t byte_of(const uint4 &v, int i) {
    const uint32_t w = i < 4 ? v.x : (i < 8 ? v.y : (i < 12 ? v.z : v.w));
    return (w >> (8 * (i & 3))) & 0xffu;
}

// Step 2: the wavefront.  Thread r owns row r of the current block of rows and works in 16-byte
// chunks: chunk c of row r may be reconstructed once row r-1 has finished chunk c (its bytes are the
// "up" values, and the last BPP bytes of its chunk c-1 the "up-left" values of the first pixel), so
// thread r runs one chunk behind thread r-1 and all threads step together.  The finished chunk is
// handed to the next row through a double-buffered shared-memory slot and stored in place with one
// 16-byte write; the next filtered chunk is fetched one step ahead.
template <int BPP>
__global__ void __launch_bounds__(kUnfThreads) unfilter_wave_kernel(const __grid_constant__ UnfilterParams p) {
    __shared__ uint4 slot[2][kUnfThreads];
    const UnfilterPass ps = pass_of(p, blockIdx.x, blockIdx.y);
    if (ps.rows == 0 || ps.row_bytes == 0) return;
    const int t = threadIdx.x;
    const uint32_t nchunks = ps.pitch / 16;

    for (uint32_t y0 = 0; y0 < ps.rows; y0 += kUnfThreads) {
        const uint32_t y = y0 + t;
        const bool live = y < ps.rows;
        uint4 *row = reinterpret_cast<uint4 *>(ps.dst + (size_t)(live ? y : 0) * ps.pitch);
        const uint4 *up_row = (live && y > 0) ? reinterpret_cast<const uint4 *>(ps.dst + (size_t)(y - 1) * ps.pitch) : nullptr;
        const uint32_t ft = live ? ps.ftype[y] : 0u;
        uint32_t left[BPP], upleft[BPP];  // last BPP reconstructed bytes of this row / of the row above
#pragma unroll
        for (int i = 0; i < BPP; ++i) left[i] = upleft[i] = 0;
        const uint32_t nsteps = nchunks + min((uint32_t)kUnfThreads, ps.rows - y0) - 1;
        __syncthreads();  // the previous block of rows is completely written (thread 0 reads it as "up")
        uint4 nx = make_uint4(0, 0, 0, 0), nu = make_uint4(0, 0, 0, 0);
        if (live && t == 0) {
            nx = row[0];
            if (up_row) nu = up_row[0];
        }
        for (uint32_t s = 0; s < nsteps; ++s) {
            const int c = (int)s - t;  // chunk this thread handles in this step
            const uint4 cx = nx, cu = nu;
            const int cn = c + 1;
            if (live && cn >= 0 && cn < (int)nchunks) {  // fetch the next chunk one step ahead
                nx = row[cn];
                if (t == 0 && up_row) nu = up_row[cn];
            }
            if (live && c >= 0 && c < (int)nchunks) {
                const uint4 up = t > 0 ? slot[(s + 1) & 1][t - 1] : cu;  // published by row y-1 in step s-1
                uint32_t o[4] = {0, 0, 0, 0};
                // one specialised 16-byte loop per filter type (rows of a warp may differ in type, but each
                // variant stays short); None and Up have no dependency along the row
                auto run = [&](auto predictor) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const uint32_t x = byte_of(cx, i), b = byte_of(up, i);
                        const uint32_t v = (x + predictor(left[i % BPP], b, upleft[i % BPP])) & 0xffu;
                        o[i / 4] |= v << (8 * (i & 3));
                        left[i % BPP] = v;
                        upleft[i % BPP] = b;
                    }
                };
                switch (ft) {
                case 1: run([](uint32_t a, uint32_t, uint32_t) { return a; }); break;
                case 2: run([](uint32_t, uint32_t b, uint32_t) { return b; }); break;
                case 3: run([](uint32_t a, uint32_t b, uint32_t) { return (a + b) >> 1; }); break;
                case 4: run([](uint32_t a, uint32_t b, uint32_t c) { return paeth_predict(a, b, c); }); break;
                default: run([](uint32_t, uint32_t, uint32_t) { return 0u; }); break;
                }
                // left[j] / upleft[j] are indexed by (byte position in chunk) % BPP; re-base them for the
                // next chunk, whose byte 0 sits 16 positions further (a static rotation, no-op if BPP | 16)
                if (16 % BPP != 0) {
                    uint32_t tl[BPP], tu[BPP];
#pragma unroll
                    for (int j = 0; j < BPP; ++j) tl[j] = left[(j + 16) % BPP], tu[j] = upleft[(j + 16) % BPP];
#pragma unroll
                    for (int j = 0; j < BPP; ++j) left[j] = tl[j], upleft[j] = tu[j];
                }
                const uint4 res = make_uint4(o[0], o[1], o[2], o[3]);
                row[c] = res;  // bytes past row_bytes inside the pitch are scratch
                slot[s & 1][t] = res;
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
    uint32_t dst_pitch;              // bytes between raster rows of dst (16-byte multiple: pixel expansion reads it aligned)
    size_t pass_stride, dst_stride;  // image blockIdx.y of a batch sits this far into the two buffers
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
    const size_t poff = (size_t)blockIdx.y * p.pass_stride;
    uint8_t *dst = p.dst + (size_t)blockIdx.y * p.dst_stride;
    for (size_t it = (size_t)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (size_t)gridDim.x * blockDim.x) {
        const uint32_t y = (uint32_t)(it / p.row_bytes), bx = (uint32_t)(it - (size_t)y * p.row_bytes);
        uint32_t outb = 0, px, py;
        if (p.bits >= 8) {
            const uint32_t bytes = p.bits / 8, x = bx / bytes, off = bx - x * bytes;
            const int ps = adam7_pass_of(x, y, px, py);
            outb = p.pass_px[ps][poff + (size_t)py * p.pass_row_bytes[ps] + (size_t)px * bytes + off];
        } else {
            const uint32_t per = 8 / p.bits, mask = (1u << p.bits) - 1u;
            for (uint32_t i = 0; i < per; ++i) {
                const uint32_t x = bx * per + i;
                if (x >= p.w) break;
                const int ps = adam7_pass_of(x, y, px, py);
                const uint32_t bit = px * p.bits;
                const uint32_t v = (p.pass_px[ps][poff + (size_t)py * p.pass_row_bytes[ps] + (bit >> 3)] >> (8 - p.bits - (bit & 7))) & mask;
                outb |= v << (8 - p.bits - i * p.bits);
            }
        }
        dst[(size_t)y * p.dst_pitch + bx] = (uint8_t)outb;
    }
}

}  // namespace normk
