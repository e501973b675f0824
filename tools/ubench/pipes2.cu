// Issue-rate microbenchmarks, second set (round 2): which pipe takes the candidates for the fused chain's
// non-blur work, and what co-issues with an FFMA2 stream on sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes2.bin pipes2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, float a, int ia, const unsigned *wsrc) {
    float acc[16];
    float2 acc2[8];
    unsigned iacc[16];
    double dacc[8];
    unsigned long long lacc[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x + i, iacc[i] = threadIdx.x + i;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        acc2[i] = make_float2(threadIdx.x + i, threadIdx.x - i), dacc[i] = threadIdx.x + i, lacc[i] = ((unsigned long long)wsrc[i] << 32) + threadIdx.x * 77 + i;
    const float2 a2 = make_float2(a, a * 0.5f);
    const double da = a;
    unsigned w = wsrc[threadIdx.x & 31];
    for (int it = 0; it < ITERS; ++it) {
        if (MODE == 0) {  // DFMA x8
#pragma unroll
            for (int i = 0; i < 8; ++i) dacc[i] = fma(dacc[i], da, 1.0);
        } else if (MODE == 1) {  // FFMA2 x8 + DFMA x8
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                acc2[i] = __ffma2_rn(acc2[i], a2, a2);
                dacc[i] = fma(dacc[i], da, 1.0);
            }
        } else if (MODE == 2) {  // IMAD.WIDE.U32 x8 (64-bit accumulate)
#pragma unroll
            for (int i = 0; i < 8; ++i) lacc[i] = (unsigned long long)(unsigned)lacc[i] * (unsigned)(lacc[i] >> 32) + lacc[i];
        } else if (MODE == 3) {  // FFMA2 x8 + IDP4A x8
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                acc2[i] = __ffma2_rn(acc2[i], a2, a2);
                iacc[i] = __dp4a(iacc[i], (unsigned)ia, iacc[i]);
            }
        } else if (MODE == 4) {  // I2FP.F32.U32 of a masked byte: LOP + convert + FADD
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] += (float)((w + i) & 0x1ffu);
            w = w * 1664525u + 1013904223u;
        } else if (MODE == 5) {  // PRMT + FADD(-2^23) unpack + FADD
#pragma unroll
            for (int i = 0; i < 16; ++i)
                acc[i] += __uint_as_float(__byte_perm(w + i, 0x4b000000u, 0x7540 + (0))) - 8388608.0f;
            w = w * 1664525u + 1013904223u;
        } else if (MODE == 6) {  // FFMA2 x8 + I2F.U8 x8 (XU)
            float c[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) c[i] = (float)((w >> (8 * (i & 3))) & 0xff);
#pragma unroll
            for (int i = 0; i < 8; ++i) acc2[i] = __ffma2_rn(acc2[i], a2, make_float2(c[i], c[i]));
            w = w * 1664525u + 1013904223u;
        } else if (MODE == 7) {  // FFMA2 x8 + ALU x8 (LOP3/SHF)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                acc2[i] = __ffma2_rn(acc2[i], a2, a2);
                iacc[i] = (iacc[i] ^ ia) + (iacc[i] >> 3);
            }
        } else if (MODE == 8) {  // FFMA2 x16 alone (two independent sets -> more ILP)
#pragma unroll
            for (int i = 0; i < 8; ++i) acc2[i] = __ffma2_rn(acc2[i], a2, a2);
#pragma unroll
            for (int i = 0; i < 8; ++i) acc2[i] = __ffma2_rn(acc2[i], a2, a2);
        } else if (MODE == 9) {  // IDP4A x16 alone
#pragma unroll
            for (int i = 0; i < 16; ++i) iacc[i] = __dp4a(iacc[i], (unsigned)ia, iacc[i]);
        } else if (MODE == 10) {  // FRND (rintf) x16
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = rintf(acc[i] * a);
        } else if (MODE == 11) {  // FFMA2 x8 + IMAD.WIDE x4
#pragma unroll
            for (int i = 0; i < 8; ++i) acc2[i] = __ffma2_rn(acc2[i], a2, a2);
#pragma unroll
            for (int i = 0; i < 4; ++i) lacc[i] = (unsigned long long)(unsigned)lacc[i] * (unsigned)(lacc[i] >> 32) + lacc[i];
        } else if (MODE == 12) {  // FFMA2 x8 + MUFU.RSQ x2
#pragma unroll
            for (int i = 0; i < 8; ++i) acc2[i] = __ffma2_rn(acc2[i], a2, a2);
#pragma unroll
            for (int i = 0; i < 2; ++i) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(acc[i]));
        } else if (MODE == 13) {  // FFMA2 x8 + F2F.F64.F32 x2 + DADD x2
#pragma unroll
            for (int i = 0; i < 8; ++i) acc2[i] = __ffma2_rn(acc2[i], a2, a2);
#pragma unroll
            for (int i = 0; i < 2; ++i) dacc[i] += (double)acc2[i].x;
        } else if (MODE == 14) {  // FMUL2 x8
#pragma unroll
            for (int i = 0; i < 8; ++i) acc2[i] = __fmul2_rn(acc2[i], a2);
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i] + iacc[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc2[i].x + acc2[i].y + (float)dacc[i] + (float)lacc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name, double ops_per_iter, float *d, unsigned *w) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    const int grid = 148 * 8;
    k<MODE><<<grid, 256>>>(d, 1.0001f, 0x01020304, w);
    cudaEventRecord(e0);
    k<MODE><<<grid, 256>>>(d, 1.0001f, 0x01020304, w);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)grid * 256 * ITERS * ops_per_iter;
    // warp-instructions per clock per SM sub-partition (4 per SM) at 1.965 GHz
    printf("%-34s %8.3f ms  %7.2f T lane-instr/s  %.3f warp-instr/clk/SMSP\n", name, ms, ops / ms / 1e9,
           ops / 32 / (ms * 1e-3) / 148 / 4 / 1.965e9);
}

int main() {
    float *d;
    unsigned *w;
    cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaMalloc(&w, 128);
    cudaMemset(w, 0x5a, 128);
    run<0>("DFMA x8", 8, d, w);
    run<1>("FFMA2 x8 + DFMA x8", 16, d, w);
    run<2>("IMAD.WIDE.U32 x8", 8, d, w);
    run<3>("FFMA2 x8 + IDP4A x8", 16, d, w);
    run<4>("LOP+I2FP+FADD x16 (3 instr)", 48, d, w);
    run<5>("PRMT+FADD+FADD x16 (3 instr)", 48, d, w);
    run<6>("FFMA2 x8 + I2F.U8 x8", 16, d, w);
    run<7>("FFMA2 x8 + ALU x16", 24, d, w);
    run<8>("FFMA2 x16", 16, d, w);
    run<9>("IDP4A x16", 16, d, w);
    run<10>("FMUL+FRND x16", 32, d, w);
    run<11>("FFMA2 x8 + IMAD.WIDE x4", 12, d, w);
    run<12>("FFMA2 x8 + MUFU.RSQ x2", 10, d, w);
    run<13>("FFMA2 x8 + F2F+DADD x2", 12, d, w);
    run<14>("FMUL2 x8", 8, d, w);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
