// Issue-rate microbenchmarks for planning the blur inner loops on sm_100a:
// FFMA vs FFMA2 (fma.rn.f32x2), IDP4A, I2F.U8, FADD, mixed FFMA+I2F.
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, float a, int ia) {
    float acc[16];
    float2 acc2[8];
    int iacc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x + i, iacc[i] = threadIdx.x + i;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc2[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
    const float2 a2 = make_float2(a, a * 0.5f);
    unsigned w = threadIdx.x * 2654435761u;
    for (int it = 0; it < ITERS; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = __fmaf_rn(acc[i], a, 1.0f);
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) acc2[i] = __ffma2_rn(acc2[i], a2, a2);
        } else if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < 16; ++i) iacc[i] = __dp4a(iacc[i], ia, iacc[i]);
        } else if (MODE == 3) {
#pragma unroll
            for (int i = 0; i < 16; ++i) { acc[i] += (float)((w >> (8 * (i & 3))) & 0xff); }
            w = w * 1664525u + 1013904223u;
        } else if (MODE == 4) {
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = __fadd_rn(acc[i], a);
        } else if (MODE == 5) {  // 16 FFMA + 16 IADD (different pipes)
#pragma unroll
            for (int i = 0; i < 16; ++i) { acc[i] = __fmaf_rn(acc[i], a, 1.0f); iacc[i] = iacc[i] * 3 + ia; }
        } else if (MODE == 6) {  // 8 FFMA2 + 16 IMAD-ish
#pragma unroll
            for (int i = 0; i < 8; ++i) acc2[i] = __ffma2_rn(acc2[i], a2, a2);
#pragma unroll
            for (int i = 0; i < 16; ++i) iacc[i] = (iacc[i] ^ ia) + (iacc[i] >> 3);
        } else if (MODE == 7) {  // 16 FFMA + 16 LOP/SHF alu
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = __fmaf_rn(acc[i], a, 1.0f);
#pragma unroll
            for (int i = 0; i < 16; ++i) iacc[i] = (iacc[i] ^ ia) + (iacc[i] >> 3);
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i] + iacc[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc2[i].x + acc2[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name, double ops_per_iter, float *d) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    const int grid = 148 * 8;
    k<MODE><<<grid, 256>>>(d, 1.0001f, 0x01020304);
    cudaEventRecord(e0);
    k<MODE><<<grid, 256>>>(d, 1.0001f, 0x01020304);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)grid * 256 * ITERS * ops_per_iter;
    printf("%-28s %8.3f ms  %8.2f Tops/s  (%.1f lane-ops/clk/SM at 1.965 GHz)\n", name, ms, ops / ms / 1e9,
           ops / (ms * 1e-3) / 148 / 1.965e9);
}

int main() {
    float *d;
    cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("FFMA x16", 16, d);
    run<1>("FFMA2 x8 (=16 fma)", 16, d);
    run<2>("IDP4A x16 (=64 mac)", 16, d);
    run<3>("I2F.U8+FADD x16", 16, d);
    run<4>("FADD x16", 16, d);
    run<5>("FFMA x16 + IMAD x16", 32, d);
    run<6>("FFMA2 x8 + ALU x32", 8 + 32, d);
    run<7>("FFMA x16 + ALU x32", 16 + 32, d);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
