// PNG scanline filtering for the encoder (the step right after the path, so results can
// leave as PNG).  src/png/scanline_writer.rs:22-61: for every row compute the None / Sub / Up /
// Average / Paeth candidates and keep the one with the smallest sum of output bytes (first minimum in
// that order); emit the filter-type byte followed by the filtered row.  Unlike unfiltering this is
// row-parallel: every candidate depends on ORIGINAL neighbour bytes only.
//
// paeth_compat = 0: the PNG-spec Paeth predictor (p = a + b - c), what any decoder undoes.
// paeth_compat = 1: bug-for-bug with the reference, whose encoder computes p = a + b + c
//                   (scanline_writer.rs:8); rows it filters with type 4 do not decode correctly,
//                   so this mode exists only to check parity with the reference's arithmetic.
// One warp per row, two sweeps: candidate sums (u32, warp-shuffle reduction), then the chosen row.
#pragma once
#include "common.cuh"

namespace normk {

struct FilterParams {
    const uint8_t *src;  // pixel rows (device), `src_pitch` apart
    size_t src_pitch;
    uint8_t *dst;        // filtered stream: rows of 1 + row_bytes bytes
    uint8_t *ftypes;     // optional: chosen filter per row
    uint32_t rows, row_bytes, bpp;
    int32_t paeth_compat;
};

__device__ __forceinline__ uint32_t paeth_enc(uint32_t a, uint32_t b, uint32_t c, bool compat) {
    const int p = compat ? (int)a + (int)b + (int)c : (int)a + (int)b - (int)c;
    const int pa = abs(p - (int)a), pb = abs(p - (int)b), pc = abs(p - (int)c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

__global__ void __launch_bounds__(256) filter_rows_kernel(const __grid_constant__ FilterParams p) {
    const int lane = threadIdx.x & 31;
    const uint32_t warps_per_grid = gridDim.x * (blockDim.x >> 5);
    const bool compat = p.paeth_compat != 0;
    for (uint32_t y = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); y < p.rows; y += warps_per_grid) {
        const uint8_t *cur = p.src + (size_t)y * p.src_pitch;
        const uint8_t *prev = y ? cur - p.src_pitch : nullptr;  // the row above the first one is all zero
        uint32_t sum[5] = {0, 0, 0, 0, 0};
        for (uint32_t i = lane; i < p.row_bytes; i += 32) {
            const uint32_t x = cur[i], a = i >= p.bpp ? cur[i - p.bpp] : 0u, b = prev ? prev[i] : 0u,
                           c = (prev && i >= p.bpp) ? prev[i - p.bpp] : 0u;
            sum[0] += x;
            sum[1] += (x - a) & 0xffu;
            sum[2] += (x - b) & 0xffu;
            sum[3] += (x - ((a + b) >> 1)) & 0xffu;
            sum[4] += (x - paeth_enc(a, b, c, compat)) & 0xffu;
        }
        uint32_t best = 0, best_sum = 0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const uint32_t s = warp_sum(sum[k]);
            const uint32_t tot = __shfl_sync(0xffffffffu, s, 0);
            if (k == 0 || tot < best_sum) best = k, best_sum = tot;  // first minimum wins (min_by_key)
        }
        uint8_t *out = p.dst + (size_t)y * ((size_t)p.row_bytes + 1);
        if (lane == 0) {
            out[0] = (uint8_t)best;
            if (p.ftypes) p.ftypes[y] = (uint8_t)best;
        }
        for (uint32_t i = lane; i < p.row_bytes; i += 32) {
            const uint32_t x = cur[i], a = i >= p.bpp ? cur[i - p.bpp] : 0u, b = prev ? prev[i] : 0u,
                           c = (prev && i >= p.bpp) ? prev[i - p.bpp] : 0u;
            uint32_t pred = 0;
            switch (best) {
            case 1: pred = a; break;
            case 2: pred = b; break;
            case 3: pred = (a + b) >> 1; break;
            case 4: pred = paeth_enc(a, b, c, compat); break;
            default: break;
            }
            out[1 + i] = (uint8_t)(x - pred);
        }
    }
}

}  // namespace normk
