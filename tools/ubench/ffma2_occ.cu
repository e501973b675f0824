// FFMA2 issue rate against warps per SM sub-partition (round 2): does an FFMA2 stream in the forms the fused chain
// uses (scalar-broadcast operand, tap pair in a uniform register, accumulator pair) fill the FMA pipe from ONE warp,
// or does it need several?  The chain runs 3 - 6 warps per sub-partition, about two of them in an FFMA2 phase.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_occ.bin ffma2_occ.cu
// MODE 0: the H item: 30 source values x 5 output pairs x 3 channels, tap k = column - 2 x pair (group A, r = 21, T = 10)
// MODE 1: the shift-form V step: pair m-1 = fma2(tB, h2, fma2(tA, h1, pair m)), 10 pairs x 3 channels (group B, r = 21)
// MODE 2: FFMA (scalar) x 30 independent accumulators, taps as constant-bank operands
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

struct Taps { float2 p[24]; float t[24]; };
#define ITERS 1024
template <int MODE>
__global__ void __launch_bounds__(512) k(float *out, const __grid_constant__ Taps t) {
    __shared__ float sm[64 * 33];
    for (int i = threadIdx.x; i < 64 * 33; i += blockDim.x) sm[i] = 1.0f + (float)(i % 7);
    __syncthreads();
    const float *sp = sm + (threadIdx.x & 31);
    float keep = 0.0f;
    float2 va[3][10];
    float sacc[30];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int m = 0; m < 10; ++m) va[c][m] = make_float2(0.0f, 0.0f);
#pragma unroll
    for (int i = 0; i < 30; ++i) sacc[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; ++it) {
        if (MODE == 0) {
            float2 acc2[5][3];
#pragma unroll
            for (int jj = 0; jj < 30; ++jj) {
                float cv[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) cv[c] = sp[(jj * 3 + c) % 60 * 33];
#pragma unroll
                for (int i = 0; i < 5; ++i) {
                    const int kk = jj - 2 * i;
                    if (kk == 0) {
#pragma unroll
                        for (int c = 0; c < 3; ++c) acc2[i][c] = __fmul2_rn(t.p[0], make_float2(cv[c], cv[c]));
                    } else if (kk > 0 && kk <= 21) {
#pragma unroll
                        for (int c = 0; c < 3; ++c) acc2[i][c] = __ffma2_rn(t.p[kk], make_float2(cv[c], cv[c]), acc2[i][c]);
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 5; ++i)
#pragma unroll
                for (int c = 0; c < 3; ++c) keep += acc2[i][c].x + acc2[i][c].y;
        } else if (MODE == 1) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float ha = sp[(c * 2) * 33], hb = sp[(c * 2 + 1) * 33];
                const float2 h1 = make_float2(ha, ha), h2 = make_float2(hb, hb);
                const float2 done = __ffma2_rn(t.p[21], h2, __ffma2_rn(t.p[20], h1, va[c][0]));
#pragma unroll
                for (int m = 1; m < 10; ++m) va[c][m - 1] = __ffma2_rn(t.p[21 - 2 * m], h2, __ffma2_rn(t.p[20 - 2 * m], h1, va[c][m]));
                va[c][9] = __ffma2_rn(t.p[1], h2, __fmul2_rn(t.p[0], h1));
                keep += done.x + done.y;
            }
        } else {
            const float h = sp[0];
#pragma unroll
            for (int k2 = 0; k2 < 2; ++k2)
#pragma unroll
                for (int i = 0; i < 30; ++i) sacc[i] = __fmaf_rn(t.t[(i + k2) % 22], h, sacc[i]);
        }
    }
    float s = keep;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int m = 0; m < 10; ++m) s += va[c][m].x + va[c][m].y;
#pragma unroll
    for (int i = 0; i < 30; ++i) s += sacc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name, int ffma_per_iter, float *out) {
    Taps t;
    for (int i = 0; i < 24; ++i) t.p[i] = make_float2(0.01f * i, 0.02f * i), t.t[i] = 0.01f * i;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    for (int warps = 1; warps <= 4; ++warps) {  // warps per sub-partition: 4 * warps per CTA, one CTA per SM
        const int threads = 128 * warps;
        k<MODE><<<148, threads>>>(out, t);
        cudaEventRecord(e0);
        k<MODE><<<148, threads>>>(out, t);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        const double instr_per_smsp = (double)ITERS * ffma_per_iter * warps;
        printf("%-34s %d warp(s)/sub-partition  %.3f ms  %.3f FMA-pipe warp-instr/clk/SMSP (1.965 GHz)\n", name, warps, ms,
               instr_per_smsp / (ms * 1e-3 * 1.965e9));
    }
}

int main() {
    float *out;
    cudaMalloc(&out, 148 * 512 * 4);
    run<0>("H item (315 FFMA2/FMUL2)", 315, out);
    run<1>("shift-form V step (66 FFMA2/FMUL2)", 66, out);
    run<2>("FFMA x60 scalar", 60, out);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
