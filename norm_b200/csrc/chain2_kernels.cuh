// Fused chain kernel, second generation: three warp groups in a software pipeline.
//
// Same arithmetic as chain_fused_kernel (S-form of oracle/norm_oracle.c, bit-exact), different
// decomposition.  Work item = (frame, row band, 120-column strip); the CTA (12 warps) walks
// down the strip in batches of RB rows with three groups working on different batches at once:
//
//   group A   cp.async staging of RB source rows (ClampToEdge), optional in-place front stage,
//             H pass: thread = (row, T-column item), T x CH outputs in registers, every tap one
//             FFMA2, 8- or 16-byte smem stores.
//   group B   thread = one blurred column for ALL channels: V pass with the running outputs of
//             each channel in registers across batches, quantise, intensity sqrt(r^2+g^2+b^2)
//             -> one f32 / pixel.  Two forms of the V pass: the rotation (row loop unrolled over
//             RB rows, RB % R == 0, output slot s takes tap (row - s) mod R) for small radii, and
//             the shift form (two rows per step, the outputs move down one register pair per
//             step through the three-operand FFMA2; RB free) for the rest.
//   group C   thread = one output column: 3x3 Sobel on the intensity rows (window kept in
//             registers, rotated by a 3-row static unroll), RGBA8 store, f64 SSIM moments
//             against the source pixel (prefetched three rows ahead).
//   (roles -> warps: see the permutation at the top of the kernel)
//
// A->B hand over f32 H planes, B->C intensity rows, each through a ring of slots with a
// full/empty pair of named barriers per slot, so no group ever waits on more than its
// neighbour.  Register budgets follow the roles (setmaxnreg, two CTAs per SM: 3 x 80 per
// thread trio): B takes 112, A keeps 72 (no spills in the H pass), C gives its share up (56).
// RGB chains from r = 33 and the RGBA blur from r = 37 hold more running outputs than that
// allows and run one CTA per SM without re-division (Fused2Cfg::BIG).
#pragma once
#include "chain_kernels.cuh"

namespace normk {

constexpr int kC2Group = 128, kC2Threads = 3 * kC2Group;
// register budgets per group (setmaxnreg); they must add up to 3 x the launch allocation (80)
#ifndef NORM_C2_ROLE_PERM
#define NORM_C2_ROLE_PERM 1   // measured: +2.4 % at r = 21 with group A on the highest warp ids
#endif
#ifndef NORM_C2_SQRT_NOSEL
#define NORM_C2_SQRT_NOSEL 1
#endif
#ifndef NORM_C2_CPAIR
#define NORM_C2_CPAIR 0   // experimental group C: two pixels per thread (fewer instructions, measured slower: see DESIGN.md)
#endif
#ifndef NORM_C2_T
#define NORM_C2_T 10
#endif
#ifndef NORM_C2_VSHIFT
#define NORM_C2_VSHIFT 1      // group B (MODE 0): shift-form V pass, two rows per step (see the group's comment)
#define NORM_C2_VSHIFT_RB 20  // its rows per batch (even; no longer tied to R)
#define NORM_C2_VUNROLL 1     // two-row steps per loop trip
#endif
#ifndef NORM_C2_VSHIFT_RBG
#define NORM_C2_VSHIFT_RBG 32 // ... in the chain after grayscale (one H plane: more rows fit; r = 21: 156 -> 162 Gpixel/s)
#endif
#ifndef NORM_C2_VSHIFT_MINR0G
#define NORM_C2_VSHIFT_MINR0G 17  // chain after grayscale: the rotation (RB = 2 x R up to r = 15) wins below
#endif
#ifndef NORM_C2_BIG_MINR
#define NORM_C2_BIG_MINR 33   // RGB chain from this radius on: one CTA per SM, see Fused2Cfg::BIG
#endif
#ifndef NORM_C2_BIG_MINR1
#define NORM_C2_BIG_MINR1 37  // ... and the RGBA blur (2 x 2 x (R - 1) running outputs in groups B and C)
#endif
#ifndef NORM_C2_BIG_T1
#define NORM_C2_BIG_T1 8      // (8-column items: the unrolled H item of four channels stays below ~900 FFMA2, beyond which
#define NORM_C2_BIG_RB1 34    //  nvcc leaves the column loop rolled, with dynamic tap indices); 15 items per row: 510 for four rounds
#endif
#ifndef NORM_C2_BIG_T
#define NORM_C2_BIG_T 12      // its H-pass outputs per thread and rows per batch (11 items per row: 374 for three rounds of 128)
#define NORM_C2_BIG_RB 34
#endif
#ifndef NORM_C2_VSHIFT_MINR0
// smallest radius that takes the shift form (measured, profiles/r02_radius_sweep.txt): below it the rotation's longer
// batches (RB = 21 or R) win; above it the rotation's RB >= R no longer fits half an SM's shared memory
#define NORM_C2_VSHIFT_MINR0 11   // chain on RGB
#define NORM_C2_VSHIFT_MINR1 9    // blur alone
#endif
#ifndef NORM_C2_ROT21
#define NORM_C2_ROT21 1       // the RGB chain at r = 21 keeps the rotation: its 21-row batch gives 252 H items = 98 % of group A's lanes (20 rows: 94 %); measured 134.3 vs 131.8 Gpixel/s
#endif
#ifndef NORM_C2_HG2_MAXR
#define NORM_C2_HG2_MAXR 27   // RGBA blur: two H groups + one V group up to this radius (0: never); see Fused2Cfg::HG2
#endif
#ifndef NORM_C2_NOTAIL
#define NORM_C2_NOTAIL 0      // 1: whole H items only (the last one computes a few columns in vain) instead of the short item
#endif
#ifndef NORM_C2_NSH
#define NORM_C2_NSH 2         // H-plane slots (A -> B ring) and intensity slots (B -> C ring) of the shift-form chain;
#define NORM_C2_NSI 2         // 2 x NSH + 2 x NSI + 2 named barriers, at most 15
#endif
#ifndef NORM_C2_REGS_H4
#define NORM_C2_REGS_H4 64  // RGBA blur: group A (H pass of four channels); groups B and C share the rest
#define NORM_C2_T4 10       // and its H-pass outputs per thread above r = 9 (10-column items: 252 per 21-row batch, 98 % of the lanes; 8 gave 315 = 82 %)
#endif
#ifndef NORM_C2_REGS_A
#define NORM_C2_REGS_A 72   // measured best of 56/128/56, 64/112/64, 72/104/64, 72/112/56, 80/96/64, 80/104/56, 80/112/48:
#define NORM_C2_REGS_B 112  // group A stops spilling at 72, group B's 3 x R running outputs + temporaries fit 112
#define NORM_C2_REGS_C 56
#endif
static_assert(NORM_C2_REGS_A + NORM_C2_REGS_B + NORM_C2_REGS_C == 240, "register budgets must sum to 3 x 80");

// MODE 0: front -> blur -> Sobel (+ SSIM)   one blurred channel triple -> intensity -> group C
// MODE 1: front -> blur                     all four channels are blurred (image_shader.wgsl:88-103 blurs
//         alpha too); after grayscale the three colour channels coincide, so two distinct channels (y, a)
//         remain: groups A and C both run the H pass (one channel each, staging shared) and group B the V
//         pass of both.  On RGBA up to r = 27 likewise: groups A and C run the H pass of two channels each, group B the V
//         pass of all four (Fused2Cfg::HG2); above, group A runs the H pass of all four channels, B and C the V pass of two each.
template <int R, bool GRAY, int MODE = 0>
struct Fused2Cfg {
    static_assert(R % 2 == 1, "odd radii");
    static constexpr int M = (R - 1) / 2;
    static constexpr int HALO = MODE == 0 ? 1 : 0;      // Sobel needs one blurred pixel around each output
    static constexpr int SW = 120;                      // output columns per strip
    static constexpr int BW = SW + 2 * HALO;            // blurred / intensity columns
    // H-pass outputs per thread.  The 4-channel H pass of the RGBA blur takes 10 where that fills the rounds of 128
    // threads (r = 21: 21 rows x 12 items = 252) and 8 where 8 does (r = 15, 17: 15 items per row give 225 / 255;
    // measured 146 vs 141 / 139 Gpixel/s).
    static constexpr bool VSHIFT = NORM_C2_VSHIFT && !NORM_C2_CPAIR && R >= (MODE == 0 ? (GRAY ? NORM_C2_VSHIFT_MINR0G : NORM_C2_VSHIFT_MINR0) : NORM_C2_VSHIFT_MINR1) &&
                                   !(NORM_C2_ROT21 && MODE == 0 && !GRAY && R == 21);
    // RGB chain at the largest radii: 3 x (R - 1) running outputs (114 at r = 39) exceed what a thread of group B can
    // hold when two CTAs share an SM (3 x 80 registers per thread trio).  Those radii run ONE CTA per SM with up to 168
    // registers per thread and no re-division, and spend the freed shared memory and registers on longer batches and
    // 16-column H items (fewer unpacks per pixel: (T + R - 1) / T).
    static constexpr bool BIG = VSHIFT && !GRAY && R >= (MODE == 0 ? NORM_C2_BIG_MINR : NORM_C2_BIG_MINR1);
    static constexpr int CTAS = BIG ? 1 : 2;            // resident CTAs per SM the kernel is built for
    static constexpr int REGS_LAUNCH = BIG ? 168 : 80;  // registers per thread at launch (__launch_bounds__)
    static constexpr int T = BIG ? (MODE == 0 ? NORM_C2_BIG_T : NORM_C2_BIG_T1) : (MODE == 1 && !GRAY && R > 9) ? ((!VSHIFT && (R == 15 || R == 17)) ? 8 : NORM_C2_T4) : NORM_C2_T;
    // H-pass items per row: SEGS full items of T columns + one short item of TAIL columns.  T = 10 with the 21-row batch
    // gives 252 items for the group's 128 threads (two rounds, 98 % of the lanes busy) and no column is computed in vain.
    static constexpr int SEGS = (T % 4 == 0 || NORM_C2_NOTAIL) ? (BW + T - 1) / T : BW / T;
    static constexpr int TAIL = (T % 4 == 0 || NORM_C2_NOTAIL) ? 0 : BW - SEGS * T;
    static_assert(TAIL % 2 == 0 && TAIL >= 0 && TAIL <= 4, "short H item: an even number of columns");
    static constexpr int HS = ((TAIL ? BW : SEGS * T) + 3) / 4 * 4;  // 124: H plane row stride = columns actually stored
    static constexpr int HL = ((M + HALO) + 3) / 4 * 4; // staged left halo, 16-byte chunks
    static constexpr int OFF0 = HL - HALO - M;          // staged column of blurred column 0's first tap
    static constexpr int SRCW = (OFF0 + HS + R - 1 + 3) / 4 * 4;
    static constexpr int STG_STRIDE = SRCW;             // words, multiple of 4
    // rows per batch: a multiple of R, about 16.  The RGBA blur (four H planes) takes R itself from R = 13 on:
    // two times 13 or 15 rows would push its shared memory past half an SM and cost the second resident CTA
    // (measured: +11 % / +19 % at r = 13 / 15; the chain variants are faster with the longer batch).
    static constexpr int RB = BIG ? (MODE == 0 ? NORM_C2_BIG_RB : NORM_C2_BIG_RB1) : VSHIFT ? ((MODE == 0 && GRAY) ? NORM_C2_VSHIFT_RBG : NORM_C2_VSHIFT_RB) : (R >= 13 && MODE == 1 && !GRAY) ? R : (R == 3 ? 21 : R * ((16 + R - 1) / R));
    static_assert(!VSHIFT || RB % 2 == 0, "shift-form V pass: two rows per step");
    // slots of the two hand-over rings and their named barriers (full / empty per slot); barrier 0 is __syncthreads
    static constexpr int NSH = VSHIFT ? NORM_C2_NSH : 2, NSI = VSHIFT ? NORM_C2_NSI : 2;
    static constexpr int BFULL = 1, BEMPTY = BFULL + NSH, BIFULL = BEMPTY + NSH, BIEMPTY = BIFULL + NSI, BA = BIEMPTY + NSI, BC = BA + 1;
    static_assert(BC <= 15, "16 named barriers per CTA");
    static constexpr int CH = MODE == 0 ? (GRAY ? 1 : 3) : (GRAY ? 2 : 4);
    // MODE 1 after grayscale (channels y, alpha): groups A and C both run the H pass (one channel each, staging
    // shared) and group B runs the V pass of both and stores RGBA8.  MODE 1 on RGBA: group A runs the H pass
    // of all four channels and groups B and C the V pass of two channels each (2 x R running outputs), because
    // 4 x R running outputs do not fit one thread's registers.
    // RGBA blur up to r = NORM_C2_HG2_MAXR: the H pass of four channels is twice the V pass of two, so with "A = H, B and
    // C = V of two channels each" groups B and C idle half the time.  Where 4 x (R - 1) running outputs fit one thread
    // (112 registers), groups A and C run the H pass of two channels each and group B the V pass of all four.
    static constexpr bool HG2 = NORM_C2_HG2_MAXR > 0 && MODE == 1 && !GRAY && VSHIFT && !BIG && R <= NORM_C2_HG2_MAXR;
    static constexpr int HGROUPS = (MODE == 1 && (GRAY || HG2)) ? 2 : 1;
    static constexpr int VGROUPS = (MODE == 1 && !GRAY && !HG2) ? 2 : 1;
    static constexpr int CH_A = (MODE == 1 && GRAY) ? 1 : HG2 ? 2 : CH;   // channels of group A's H pass (from 0)
    static constexpr int CH_C = HG2 ? 2 : 1;                              // channels of group C's H pass (the last ones)
    static constexpr int HTHREADS = HGROUPS * 128;
    // register budgets (setmaxnreg), summing to 3 x 80
    static constexpr int REGS_H = BIG ? 168 : HG2 ? (R > 23 ? 56 : 64) : MODE == 1 ? (GRAY ? 56 : NORM_C2_REGS_H4) : NORM_C2_REGS_A;              // MODE 1 + GRAY: two H groups
    static constexpr int REGS_V = BIG ? 168 : HG2 ? (R > 23 ? 128 : 112) : MODE == 1 ? (GRAY ? 128 : (240 - NORM_C2_REGS_H4) / 2) : NORM_C2_REGS_B;
    static constexpr int REGS_C = BIG ? 168 : MODE == 0 ? 240 - REGS_H - REGS_V : 0;
    // source byte of channel c: MODE 1 + GRAY blurs (y, alpha) = bytes (0, 3)
    __host__ __device__ static constexpr int chan_byte(int c) { return (MODE == 1 && GRAY && c == 1) ? 3 : c; }
    static constexpr int IS = 128;                      // intensity row stride
    static constexpr size_t kStgBytes = (size_t)RB * STG_STRIDE * 4;
    static constexpr size_t kHbBytes = (size_t)CH * RB * HS * 4;
    // intensity rows exist only when Sobel follows.  Pair version of group C: two extra rows in front of every slot
    // hold the previous batch's last two rows (group B re-stores them), so a slot is self-contained for the 3x3 window.
    static constexpr int ICARRY = NORM_C2_CPAIR ? 2 : 0;
    static constexpr size_t kIstBytes = MODE == 0 ? (size_t)(RB + ICARRY) * IS * 4 : 0;
    static constexpr size_t kLutDBytes = (NORM_C2_CPAIR && MODE == 0) ? 256 * 8 : 0;  // luma of (c,c,c) as f64
    // + 64 words of slack: the last H segment's unused outputs read a few words past a staged row
    static constexpr size_t kSmem = 2 * kStgBytes + 256 + NSH * kHbBytes + NSI * kIstBytes + 256 * 4 * 2 + 256 + kLutDBytes;
    static_assert(T % 2 == 0 && HS % 4 == 0, "8- or 16-byte H stores");
};

enum {
    kB2Full0 = 1, kB2Full1 = 2, kB2Empty0 = 3, kB2Empty1 = 4,      // A <-> B: H planes
    kB2IFull0 = 5, kB2IFull1 = 6, kB2IEmpty0 = 7, kB2IEmpty1 = 8,  // B <-> C: intensity rows
    kB2A = 9, kB2C = 10
};

// stores executed under a predicate instead of a branch
__device__ __forceinline__ void st_global_pred_u32(void *ptr, uint32_t v, bool pred) {
    asm volatile("{ .reg .pred p; setp.ne.u32 p, %0, 0; @p st.global.u32 [%1], %2; }" ::"r"((uint32_t)pred), "l"(ptr), "r"(v) : "memory");
}
__device__ __forceinline__ void st_global_pred_u16(void *ptr, uint32_t v, bool pred) {
    asm volatile("{ .reg .pred p; setp.ne.u32 p, %0, 0; @p st.global.u16 [%1], %2; }" ::"r"((uint32_t)pred), "l"(ptr), "h"((uint16_t)v) : "memory");
}

// Which frame, strip and output rows [ya, yb) work item `item` stands for.  Static partition: item = (frame, band,
// strip), bands of band_rows rows.  Balanced partition (one giant frame): CTA k = item % gridDim.x owns strip-rows
// [k*per, (k+1)*per) of the strips laid end to end; its j-th item (j = item / gridDim.x) is the part of that range inside
// strip (first strip + j); parts that do not exist are `empty` (every warp group skips them alike).
__device__ __forceinline__ void c2_item_rows(const FusedParams &p, uint32_t item, uint32_t items_per_frame, uint32_t &f, int &strip,
                                             int &ya, int &yb, bool &empty) {
    if (p.balanced) {
        const uint32_t rows = p.out_y1 - p.out_y0, k = item % gridDim.x, j = item / gridDim.x;
        const uint32_t total = p.strips * rows, beg = k * p.per, end = min(total, beg + p.per);
        const uint32_t col = beg / rows + j;
        const uint32_t start = j == 0 ? beg : col * rows, stop = min(end, (col + 1) * rows);
        f = 0, strip = (int)col;
        empty = start >= stop || beg >= total;
        ya = (int)(p.out_y0 + (empty ? 0u : start - col * rows));
        yb = (int)(p.out_y0 + (empty ? 0u : stop - col * rows));
    } else {
        f = item / items_per_frame;
        const uint32_t irem = item - f * items_per_frame;
        const int band = (int)(irem / p.strips);
        strip = (int)(irem - (uint32_t)band * p.strips);
        ya = (int)p.out_y0 + band * (int)p.band_rows, yb = min(ya + (int)p.band_rows, (int)p.out_y1);
        empty = false;
    }
}

// V pass of one batch for channels [C0, C0+NCH) of one blurred column (MODE 1): NCH x R running outputs in
// registers (static rotation: the row loop is fully unrolled and RB % R == 0), quantise, store those bytes
// of the RGBA8 result.  hp = H planes of this batch's slot at this thread's column.
template <typename C, int R, bool GRAY, int C0, int NCH>
__device__ __forceinline__ void v_pass_batch(const FusedParams &p, const float *hp, int b, int NB, bool in_image, uint8_t *orow,
                                             float2 (*vacc2)[R / 2], float *vaccl) {
    if constexpr (C::VSHIFT) {
        // shift form (see group B of MODE 0): pair m of vacc2 = running outputs (first unfinished + 2m, + 2m + 1)
        constexpr int NP = (R - 1) / 2;
        for (int j = 0; j < C::RB; j += 2) {
            uint32_t q0[NCH], q1[NCH];
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const float ha = hp[((C0 + c) * C::RB + j) * C::HS], hb = hp[((C0 + c) * C::RB + j + 1) * C::HS];
                const float2 h1 = make_float2(ha, ha), h2 = make_float2(hb, hb);
                const float2 done = __ffma2_rn(p.tp[R], h2, __ffma2_rn(p.tp[R - 1], h1, vacc2[c][0]));
#pragma unroll
                for (int m = 1; m < NP; ++m)
                    vacc2[c][m - 1] = __ffma2_rn(p.tp[R - 2 * m], h2, __ffma2_rn(p.tp[R - 1 - 2 * m], h1, vacc2[c][m]));
                vacc2[c][NP - 1] = __ffma2_rn(p.tp[1], h2, __fmul2_rn(p.tp[0], h1));
                q0[c] = __float_as_uint(__fadd_rn(done.x, 12582912.0f)) & 0xffu;
                q1[c] = __float_as_uint(__fadd_rn(done.y, 12582912.0f)) & 0xffu;
            }
            const int n = b * C::RB + j - (R - 1);  // output row (relative to the band's first) completed by row j
            uint8_t *px = orow + (ptrdiff_t)n * (ptrdiff_t)p.out.pitch;
            const bool v0 = n >= 0 && n < NB && in_image, v1 = n + 1 >= 0 && n + 1 < NB && in_image;
            if constexpr (GRAY) {
                st_global_pred_u32(px, q0[0] * 0x010101u | (q0[NCH - 1] << 24), v0);
                st_global_pred_u32(px + p.out.pitch, q1[0] * 0x010101u | (q1[NCH - 1] << 24), v1);
            } else if constexpr (NCH == 4) {  // all four bytes of the pixel
                st_global_pred_u32(px, q0[0] | (q0[1] << 8) | (q0[2] << 16) | (q0[3] << 24), v0);
                st_global_pred_u32(px + p.out.pitch, q1[0] | (q1[1] << 8) | (q1[2] << 16) | (q1[3] << 24), v1);
            } else {
                static_assert(NCH == 2, "RGBA: two or four channels per V-pass thread");
                st_global_pred_u16(px + C0, q0[0] | (q0[1] << 8), v0);
                st_global_pred_u16(px + p.out.pitch + C0, q1[0] | (q1[1] << 8), v1);
            }
        }
    } else {
#pragma unroll
    for (int j = 0; j < C::RB; ++j) {
        uint32_t q[NCH];
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            const float h = hp[((C0 + c) * C::RB + j) * C::HS];
            const float2 h2 = make_float2(h, h);
#pragma unroll
            for (int i = 0; i < R / 2; ++i) {
                const int k = ((j - 2 * i) % R + R) % R;  // tap of slot 2i; slot 2i+1 takes k-1 (mod R)
                if (k >= 1)
                    vacc2[c][i] = __ffma2_rn(p.tp[k], h2, vacc2[c][i]);
                else
                    vacc2[c][i] = __ffma2_rn(make_float2(p.taps.t[0], p.taps.t[R - 1]), h2, vacc2[c][i]);
            }
            vaccl[c] = __fmaf_rn(p.taps.t[(j + 1) % R], h, vaccl[c]);
            const int done = (j + 1) % R;  // the output that just took its last tap
            float o;
            if (done == R - 1) {
                o = vaccl[c], vaccl[c] = 0.0f;
            } else if (done & 1) {
                o = vacc2[c][done / 2].y, vacc2[c][done / 2].y = 0.0f;
            } else {
                o = vacc2[c][done / 2].x, vacc2[c][done / 2].x = 0.0f;
            }
            // blurred values lie in [0, 255.0001]: the 1.5*2^23 trick alone is round-half-even + clamp
            q[c] = __float_as_uint(__fadd_rn(o, 12582912.0f)) & 0xffu;
        }
        const int n = b * C::RB + j - (R - 1);  // output row relative to the band's first row
        // predicated stores, no branch: a branch per row inside this unrolled body keeps the compiler from
        // holding the tap pairs in uniform registers (they end up in 2 x R vector registers and spill)
        const bool valid = n >= 0 && n < NB && in_image;
        uint8_t *px = orow + (size_t)n * p.out.pitch;
        if constexpr (GRAY) {  // channels (y, alpha) -> (y, y, y, alpha)
            st_global_pred_u32(px, q[0] * 0x010101u | (q[NCH - 1] << 24), valid);
        } else {               // two of the four bytes
            static_assert(NCH == 2, "RGBA: two channels per V-pass thread");
            st_global_pred_u16(px + C0, q[0] | (q[1] << 8), valid);
        }
    }
    }  // rotation
}

// Blur-only consumer (MODE 1, group B): V pass for channels [C0, C0+NCH) of one blurred column per thread.
// kParties = all threads that meet at the H-plane barriers (producing groups and this one).
template <typename C, int R, bool GRAY, int C0, int NCH, int kParties>
__device__ __forceinline__ void blur_only_consumer(const FusedParams &p, const float *Hb_base, int bcol, uint32_t items_per_frame,
                                                   uint32_t total) {
    constexpr int kHbFloats = C::CH * C::RB * C::HS;
    const int W = (int)p.in.w, Himg = (int)p.in.h;
    const bool active = bcol < C::BW;
    uint32_t g = 0;
    bar_arrive(kB2Empty0, kParties);  // both H-plane slots start out empty
    bar_arrive(kB2Empty1, kParties);
    for (uint32_t item = blockIdx.x; item < total; item += gridDim.x) {
        uint32_t f;
        int strip, ya, yb;
        bool empty = false;
        c2_item_rows(p, item, items_per_frame, f, strip, ya, yb, empty);
        if (empty) continue;
        // The last strip of a row is shifted left so that it ends at the image's right edge (to a multiple of four columns:
        // the staging copies are 16-byte aligned).  A ragged strip with a few valid columns would stage and blur 120
        // columns through the slow clamped path and, with one item per CTA, set the kernel's time.  The `skip` columns the
        // shifted strip shares with its neighbour are computed twice and stored once.
        const int x0 = (W >= C::SW && strip * C::SW + C::SW > W) ? ((W - C::SW + 3) & ~3) : strip * C::SW;
        const int skip = strip * C::SW - x0;
        const int NB = yb - ya;  // MODE 1: blurred rows == output rows
        const int nbatch = (NB + R - 1 + C::RB - 1) / C::RB;
        const int bx = x0 + bcol;
        uint8_t *orow = p.out.ptr + (size_t)f * p.out.frame_stride + (size_t)ya * p.out.pitch + (size_t)min(bx, W - 1) * 4;
        float2 vacc2[NCH][R / 2];
        float vaccl[NCH];
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            vaccl[c] = 0.0f;
#pragma unroll
            for (int i = 0; i < R / 2; ++i) vacc2[c][i] = make_float2(0.0f, 0.0f);
        }
        for (int b = 0; b < nbatch; ++b, ++g) {
            const int slot = (int)(g & 1u);
            bar_sync(kB2Full0 + slot, kParties);
            if (active)
                v_pass_batch<C, R, GRAY, C0, NCH>(p, Hb_base + slot * kHbFloats + bcol, b, NB, bx < W && bcol >= skip, orow, vacc2, vaccl);
            __threadfence_block();
            bar_arrive(kB2Empty0 + slot, kParties);
        }
    }
}

template <int R, bool GRAY, bool FRONT, int MODE = 0>
__global__ void __launch_bounds__(kC2Threads, (Fused2Cfg<R, GRAY, MODE>::CTAS)) chain_fused2_kernel(const __grid_constant__ FusedParams p) {
    using C = Fused2Cfg<R, GRAY, MODE>;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ double red[5][4];
    uint32_t *stg_base = reinterpret_cast<uint32_t *>(smem_raw);
    float *Hb_base = reinterpret_cast<float *>(smem_raw + 2 * C::kStgBytes + 256);
    float *Ist_base = reinterpret_cast<float *>(smem_raw + 2 * C::kStgBytes + 256 + C::NSH * C::kHbBytes);
    float *lumaLut = reinterpret_cast<float *>(smem_raw + 2 * C::kStgBytes + 256 + C::NSH * C::kHbBytes + C::NSI * C::kIstBytes);
    float *intLut = lumaLut + 256;
    uint8_t *flut = reinterpret_cast<uint8_t *>(intLut + 256);
    double *lumaLutD = reinterpret_cast<double *>(flut + 256);
    constexpr int kHbFloats = C::CH * C::RB * C::HS;
    constexpr int kIstFloats = (C::RB + C::ICARRY) * C::IS;

    // role index: which of the three groups a warp belongs to.  The warp scheduler prefers higher warp ids, so the
    // mapping of roles to physical warps decides which group wins an issue slot when several are ready.
    const int tid = NORM_C2_ROLE_PERM == 1   ? (int)((threadIdx.x + 128u) % 384u)                                        /* A = warps 8-11, C = 4-7, B = 0-3 */
                    : NORM_C2_ROLE_PERM == 2 ? (int)(threadIdx.x >= 256u ? threadIdx.x - 256u : (threadIdx.x >= 128u ? threadIdx.x : threadIdx.x + 256u))  /* A = 8-11, B = 4-7, C = 0-3 */
                                             : (int)threadIdx.x;
    const int W = (int)p.in.w, Himg = (int)p.in.h;

    if (tid < 256) {
        const float c = (float)tid;
        lumaLut[tid] = __fadd_rn(__fadd_rn(__fmul_rn(c, kLumaR), __fmul_rn(c, kLumaG)), __fmul_rn(c, kLumaB));
        intLut[tid] = __fsqrt_rn(__fmaf_rn(c, c, __fmaf_rn(c, c, __fmul_rn(c, c))));
        flut[tid] = p.lut.v[tid];
        if (C::kLutDBytes) lumaLutD[tid] = (double)lumaLut[tid];
    }
    __syncthreads();

    const uint32_t items_per_frame = p.strips * p.bands;
#define C2_TOTAL_ITEMS() (p.balanced ? gridDim.x * p.maxi : items_per_frame * p.in.frames)  /* < 2^31: checked by the host */
    uint32_t g = 0;  // running batch counter: ring slot = g mod (slots of the ring), identical sequence in all groups
    // threads meeting at the H-plane barriers: A + B, plus C when it shares the blur-only V pass (RGBA)
    constexpr int kABParties = (C::HGROUPS + C::VGROUPS) * kC2Group;  // producing + consuming groups

// per-item geometry, identical in the three groups
#define C2_ITEM_GEOMETRY()                                                                              \
    uint32_t f;                                                                                         \
    int strip, ya, yb;                                                                                  \
    bool c2_empty = false;                                                                              \
    c2_item_rows(p, item, items_per_frame, f, strip, ya, yb, c2_empty);                                 \
    /* last strip shifted left to end at the right image edge; its first `skip` columns belong to its neighbour */ \
    const int x0 = (W >= C::SW && strip * C::SW + C::SW > W) ? ((W - C::SW + 3) & ~3) : strip * C::SW;               \
    const int skip = strip * C::SW - x0;                                                                \
    const int o_lo = max(ya - C::HALO, 0), o_hi = min(yb - 1 + C::HALO, Himg - 1); /* blurred rows needed */ \
    const int NB = o_hi - o_lo + 1;                                                                     \
    const int Q = NB + R - 1; /* rows of the clamped source sequence s(q) = clamp(o_lo - M + q) */      \
    const int nbatch = (Q + C::RB - 1) / C::RB;                                                         \
    (void)x0, (void)NB, (void)skip, (void)c2_empty

    // =========================== group A (and C in MODE 1): staging + H pass ===========================
    // ROLE 0 = group A: channels [0, CH_A).  ROLE 1 = group C in MODE 1 after grayscale: the alpha channel.
    auto h_group = [&](auto role_tag) {
        constexpr int ROLE = decltype(role_tag)::value;
        constexpr int c0 = ROLE == 0 ? 0 : C::CH - C::CH_C;
        constexpr int NCHH = ROLE == 0 ? C::CH_A : C::CH_C;
        if constexpr (!C::BIG) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(C::REGS_H));
        constexpr int kHT = C::HTHREADS;
        const int hid = tid < kC2Group ? tid : tid - kC2Group;  // index within the H-pass threads
        constexpr int NC4 = C::SRCW / 4;           // 16-byte chunks per staged row
        constexpr int NRG = kHT / NC4;             // staging row groups (threads beyond NRG*NC4 idle)
        const int st_c4 = hid % NC4, st_rg = hid / NC4;

        for (uint32_t item = blockIdx.x, total = C2_TOTAL_ITEMS(); item < total; item += gridDim.x) {
            C2_ITEM_GEOMETRY();
            if (c2_empty) continue;  // (balanced partition: this CTA's range ended in an earlier strip)
            const uint8_t *in_frame = p.in.ptr + (size_t)f * p.in.frame_stride;
            const uint32_t pitch32 = (uint32_t)p.in.pitch;
            // cp.async one batch of source rows (ClampToEdge in x and y) into staging buffer `buf`.
            // Everything but the frame base is recomputed here so that nothing stays live (or gets
            // spilled) across the H pass; the row loop is unrolled and costs a clamp, one 32x32+64
            // multiply-add and the copy per 16-byte chunk.
            auto stage_issue = [&](int bb, int buf) {
                if (st_rg < NRG) {
                    constexpr int KMAX = (C::RB + NRG - 1) / NRG;
                    const int st_x = x0 - C::HL + 4 * st_c4;
                    const uint32_t d0 = (uint32_t)__cvta_generic_to_shared(stg_base + buf * (C::RB * C::STG_STRIDE) + 4 * st_c4) +
                                        (uint32_t)st_rg * (C::STG_STRIDE * 4);
                    const int r0 = o_lo - C::M + bb * C::RB + st_rg;
                    if (st_x >= 0 && st_x + 3 <= W - 1) {
                        const uint8_t *src0 = in_frame + (size_t)st_x * 4;
#pragma unroll
                        for (int k = 0; k < KMAX; ++k) {
                            if (st_rg + k * NRG < C::RB) {
                                const uint32_t sy = (uint32_t)clampi(r0 + k * NRG, 0, Himg - 1);
                                const uint8_t *src = src0 + (uint64_t)sy * pitch32;
                                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + k * NRG * (C::STG_STRIDE * 4)), "l"(src)
                                             : "memory");
                            }
                        }
                    } else {
                        const int xc0 = clampi(st_x, 0, W - 1), xc1 = clampi(st_x + 1, 0, W - 1), xc2 = clampi(st_x + 2, 0, W - 1),
                                  xc3 = clampi(st_x + 3, 0, W - 1);
                        for (int k = 0; st_rg + k * NRG < C::RB; ++k) {
                            const uint32_t sy = (uint32_t)clampi(r0 + k * NRG, 0, Himg - 1);
                            const uint32_t *row = reinterpret_cast<const uint32_t *>(in_frame + (uint64_t)sy * pitch32);
                            const uint32_t d = d0 + k * NRG * (C::STG_STRIDE * 4);
                            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(row + xc0) : "memory");
                            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 4), "l"(row + xc1) : "memory");
                            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 8), "l"(row + xc2) : "memory");
                            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 12), "l"(row + xc3) : "memory");
                        }
                    }
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
            };

            bar_sync(C::BA, kHT);  // previous item's last H pass is done with the staging buffers
            stage_issue(0, 0);

            for (int b = 0; b < nbatch; ++b, ++g) {
                uint32_t *stg = stg_base + (b & 1) * (C::RB * C::STG_STRIDE);
                asm volatile("cp.async.wait_group 0;" ::: "memory");
                bar_sync(C::BA, kHT);
                if (b + 1 < nbatch) stage_issue(b + 1, (b + 1) & 1);
                if (FRONT) {
                    if (st_rg < NRG) {
                        for (int j = st_rg; j < C::RB; j += NRG) {
                            uint4 *cell = reinterpret_cast<uint4 *>(stg + j * C::STG_STRIDE + 4 * st_c4);
                            uint4 v = *cell;
                            uint32_t *pv = &v.x;
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                uint32_t r = pv[e] & 0xffu, gg = (pv[e] >> 8) & 0xffu, bb = (pv[e] >> 16) & 0xffu;
                                if (!p.lut_identity) r = flut[r], gg = flut[gg], bb = flut[bb];  // (uniform branch)
                                if (GRAY) {
                                    r = quant_u8(luma_fused((float)r, (float)gg, (float)bb));
                                    gg = bb = r;
                                }
                                if (p.invert) r = 255u - r, gg = 255u - gg, bb = 255u - bb;
                                pv[e] = pack_rgba(r, gg, bb, pv[e] >> 24);
                            }
                            *cell = v;
                        }
                    }
                    bar_sync(C::BA, kHT);
                }

                const int slot = (int)(g % (uint32_t)C::NSH);
                bar_sync(C::BEMPTY + slot, kABParties);  // the consuming group(s) released this H-plane slot
                // one thread = (row, TT consecutive blurred columns) of the channels [c0, c0 + NCHH): the TT + R - 1 source
                // pixels are read once and scattered into TT x NCHH register accumulators, outputs paired so that every
                // tap is one FFMA2.  The first tap pair {t[0], 0} initialises both halves by a multiply.
                auto h_item = [&](auto tt_tag, int j, int col0) {
                    constexpr int TT = decltype(tt_tag)::value, TP = TT / 2;
                    const uint32_t *sp = stg + j * C::STG_STRIDE + C::OFF0 + col0;
                    float2 acc2[TP][NCHH];
#pragma unroll
                    for (int jj = 0; jj < TT + R - 1; ++jj) {
                        const uint32_t px = sp[jj];
                        float cv[NCHH];
#pragma unroll
                        for (int c = 0; c < NCHH; ++c) cv[c] = byte_f(px, C::chan_byte(c0 + c));
#pragma unroll
                        for (int i = 0; i < TP; ++i) {
                            const int k = jj - 2 * i;  // tap of output 2i; output 2i+1 takes k-1 (a zero tap adds exactly 0)
                            if (k == 0) {              // t[0]*v rounds like fma(t[0], v, 0); 0*v = +0 (v >= 0, finite)
#pragma unroll
                                for (int c = 0; c < NCHH; ++c) acc2[i][c] = __fmul2_rn(p.tp[0], make_float2(cv[c], cv[c]));
                            } else if (k > 0 && k <= R) {
#pragma unroll
                                for (int c = 0; c < NCHH; ++c)
                                    acc2[i][c] = __ffma2_rn(p.tp[k], make_float2(cv[c], cv[c]), acc2[i][c]);
                            }
                        }
                    }
                    float *hp = Hb_base + slot * kHbFloats + j * C::HS + col0;
#pragma unroll
                    for (int c = 0; c < NCHH; ++c) {
                        if constexpr (TT % 4 == 0) {
#pragma unroll
                            for (int i = 0; i < TP; i += 2)
                                *reinterpret_cast<float4 *>(hp + (c0 + c) * (C::RB * C::HS) + 2 * i) =
                                        make_float4(acc2[i][c].x, acc2[i][c].y, acc2[i + 1][c].x, acc2[i + 1][c].y);
                        } else {
#pragma unroll
                            for (int i = 0; i < TP; ++i)
                                *reinterpret_cast<float2 *>(hp + (c0 + c) * (C::RB * C::HS) + 2 * i) = acc2[i][c];
                        }
                    }
                };
                for (int it = hid & (kC2Group - 1); it < C::RB * C::SEGS; it += kC2Group)
                    h_item(std::integral_constant<int, C::T>{}, it / C::SEGS, (it % C::SEGS) * C::T);
                if constexpr (C::TAIL > 0) {
                    // the columns left over (the Sobel halo makes BW = 122): one thread per (row, column, channel) of them,
                    // spread over the whole group, so the leftovers lengthen every warp's batch by one short pass of R
                    // taps instead of one warp's by a whole item (measured +0.6 % at r = 21; scalar FMAs, taps as constant-bank operands: a second
                    // user of the tap PAIRS would push them out of the uniform registers)
                    constexpr int UNITS = C::RB * C::TAIL * NCHH;
                    for (int u = hid & (kC2Group - 1); u < UNITS; u += kC2Group) {
                        const int j = u / (C::TAIL * NCHH), rem = u - j * (C::TAIL * NCHH), o = rem / NCHH, c = rem - o * NCHH;
                        const uint32_t *sp = stg + j * C::STG_STRIDE + C::OFF0 + C::SEGS * C::T + o;
                        const uint32_t sel = 0x4440u | (uint32_t)C::chan_byte(c0 + c);  // byte -> low byte, zeros above
                        float acc = __fmul_rn(p.taps.t[0], __uint2float_rn(__byte_perm(sp[0], 0u, sel)));
#pragma unroll
                        for (int k = 1; k < R; ++k) acc = __fmaf_rn(p.taps.t[k], __uint2float_rn(__byte_perm(sp[k], 0u, sel)), acc);
                        Hb_base[slot * kHbFloats + (c0 + c) * (C::RB * C::HS) + j * C::HS + C::SEGS * C::T + o] = acc;
                    }
                }
                __threadfence_block();
                bar_arrive(C::BFULL + slot, kABParties);
            }
        }
    };
    if (tid < kC2Group) {
        h_group(std::integral_constant<int, 0>{});
    } else if (C::HGROUPS == 2 && tid >= 2 * kC2Group) {
        h_group(std::integral_constant<int, 1>{});
    } else if (tid < 2 * kC2Group) {
        // =========================== group B: V pass + quantise + intensity ===========================
        if constexpr (MODE == 1) {
            if constexpr (!C::BIG) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(C::REGS_V));
            blur_only_consumer<C, R, GRAY, 0, (C::HG2 ? 4 : 2), kABParties>(p, Hb_base, tid - kC2Group, items_per_frame, C2_TOTAL_ITEMS());
            return;
        }
        if constexpr (!C::BIG) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(C::REGS_V));
        const int bcol = tid - kC2Group;  // blurred column of the strip owned by this thread
        const bool active = bcol < C::BW;

        for (int sl = 0; sl < C::NSH; ++sl) bar_arrive(C::BEMPTY + sl, 2 * kC2Group);  // every H-plane slot starts out empty

        for (uint32_t item = blockIdx.x, total = C2_TOTAL_ITEMS(); item < total; item += gridDim.x) {
            C2_ITEM_GEOMETRY();
            if (c2_empty) continue;  // (balanced partition: this CTA's range ended in an earlier strip)
            if constexpr (C::VSHIFT) {
                // Shift form.  After the step for H rows (j, j+1) register pair m holds the running outputs (o, o+1) =
                // first unfinished output + (2m, 2m+1); output o + i has taken taps 0 .. R-2-i.  One step feeds row j with
                // tap pair tp[R-1-2m] and row j+1 with tp[R-2m] into pair m and writes the result to pair m-1: the
                // three-operand FFMA2 shifts the window by two outputs for free, the taps are the same for every step, so
                // the row loop is a real loop (a tenth of the unrolled rotation's code) and RB is not tied to R.
                // Pair 0's result is the two outputs that completed with rows j and j+1; the top pair starts from
                // nothing: tp[0] = {t[0], 0} by a multiply, then tp[1] = {t[1], t[0]}; tp[R] = {0, t[R-1]} adds an
                // exact zero to the output that was complete after row j.  Same fma chain per output as the oracle's.
                constexpr int NP = (R - 1) / 2;  // pairs kept between steps
                constexpr int kVU = NORM_C2_VUNROLL;
                float2 va[C::CH][NP];
#pragma unroll
                for (int c = 0; c < C::CH; ++c)
#pragma unroll
                    for (int m = 0; m < NP; ++m) va[c][m] = make_float2(0.0f, 0.0f);
                for (int b = 0; b < nbatch; ++b, ++g) {
                    const int slot = (int)(g % (uint32_t)C::NSH), islot = (int)(g % (uint32_t)C::NSI);
                    bar_sync(C::BFULL + slot, 2 * kC2Group);      // H planes of this batch are ready
                    bar_sync(C::BIEMPTY + islot, 2 * kC2Group);   // group C has released this intensity slot
                    if (active) {
                        const float *hp = Hb_base + slot * kHbFloats + bcol;
                        float *ist = Ist_base + islot * kIstFloats + bcol;
#pragma unroll kVU
                        for (int j = 0; j < C::RB; j += 2) {
                            float v0[C::CH], v1[C::CH];
#pragma unroll
                            for (int c = 0; c < C::CH; ++c) {
                                const float ha = hp[(c * C::RB + j) * C::HS], hb = hp[(c * C::RB + j + 1) * C::HS];
                                const float2 h1 = make_float2(ha, ha), h2 = make_float2(hb, hb);
                                const float2 done = __ffma2_rn(p.tp[R], h2, __ffma2_rn(p.tp[R - 1], h1, va[c][0]));
#pragma unroll
                                for (int m = 1; m < NP; ++m)
                                    va[c][m - 1] = __ffma2_rn(p.tp[R - 2 * m], h2, __ffma2_rn(p.tp[R - 1 - 2 * m], h1, va[c][m]));
                                va[c][NP - 1] = __ffma2_rn(p.tp[1], h2, __fmul2_rn(p.tp[0], h1));
                                // round-half-even through the 1.5*2^23 trick: the quantised byte sits in the low
                                // mantissa bits of (o + 1.5*2^23)
                                v0[c] = __fadd_rn(done.x, 12582912.0f), v1[c] = __fadd_rn(done.y, 12582912.0f);
                            }
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const float *v = e ? v1 : v0;
                                float I;
                                if (GRAY) {
                                    I = intLut[__float_as_uint(v[0]) & 0xffu];
                                } else {
                                    // r^2 + g^2 + b^2 of the three quantised bytes: pack them (two PRMT) and take one byte
                                    // dot product; the radicand is an integer < 2^18, so its conversion is exact and
                                    // equals the float sum of squares the oracle forms
                                    const uint32_t rg = __byte_perm(__float_as_uint(v[0]), __float_as_uint(v[C::CH > 1 ? 1 : 0]), 0x7740);
                                    const uint32_t rgb = __byte_perm(rg, __float_as_uint(v[C::CH - 1]), 0x5410);
                                    I = sqrt_rn_inrange0((float)__dp4a(rgb, rgb, 0u));
                                }
                                ist[(j + e) * C::IS] = I;
                            }
                        }
                    }
                    __threadfence_block();
                    bar_arrive(C::BEMPTY + slot, 2 * kC2Group);   // done reading the H planes
                    bar_arrive(C::BIFULL + islot, 2 * kC2Group);   // intensity rows of this batch are ready
                }
                continue;
            }
            float2 vacc2[C::CH][R / 2];  // running outputs: slots (2i, 2i+1) packed, slot R-1 alone
            float vaccl[C::CH];
#if NORM_C2_CPAIR
            float icarry0 = 0.0f, icarry1 = 0.0f;
#endif
#pragma unroll
            for (int c = 0; c < C::CH; ++c) {
                vaccl[c] = 0.0f;
#pragma unroll
                for (int i = 0; i < R / 2; ++i) vacc2[c][i] = make_float2(0.0f, 0.0f);
            }
            for (int b = 0; b < nbatch; ++b, ++g) {
                const int slot = (int)(g % (uint32_t)C::NSH), islot = (int)(g % (uint32_t)C::NSI);
                bar_sync(C::BFULL + slot, 2 * kC2Group);      // H planes of this batch are ready
                bar_sync(C::BIEMPTY + islot, 2 * kC2Group);   // group C has released this intensity slot
                if (active) {
                    const float *hp = Hb_base + slot * kHbFloats + bcol;
#if NORM_C2_CPAIR
                    // blurred column bcol lives at index bcol + 1 (keeps group C's pixel pairs 8-byte aligned);
                    // rows -2, -1 of the slot = the previous batch's last two rows
                    float *ist = Ist_base + islot * kIstFloats + C::ICARRY * C::IS + bcol + 1;
                    ist[-2 * C::IS] = icarry0, ist[-1 * C::IS] = icarry1;
#else
                    float *ist = Ist_base + islot * kIstFloats + bcol;
#endif
#pragma unroll
                    for (int j = 0; j < C::RB; ++j) {
                        float v[C::CH];
#pragma unroll
                        for (int c = 0; c < C::CH; ++c) {
                            const float h = hp[(c * C::RB + j) * C::HS];
                            const float2 h2 = make_float2(h, h);
                            // slot s takes tap (j - s) mod R; one FFMA2 per slot pair
#pragma unroll
                            for (int i = 0; i < R / 2; ++i) {
                                const int k = ((j - 2 * i) % R + R) % R;  // tap of slot 2i; slot 2i+1 takes k-1 (mod R)
                                if (k >= 1)
                                    vacc2[c][i] = __ffma2_rn(p.tp[k], h2, vacc2[c][i]);
                                else
                                    vacc2[c][i] = __ffma2_rn(make_float2(p.taps.t[0], p.taps.t[R - 1]), h2, vacc2[c][i]);
                            }
                            vaccl[c] = __fmaf_rn(p.taps.t[(j + 1) % R], h, vaccl[c]);  // slot R-1
                            const int done = (j + 1) % R;  // the output that just took its last tap
                            float o;
                            if (done == R - 1) {
                                o = vaccl[c], vaccl[c] = 0.0f;
                            } else if (done & 1) {
                                o = vacc2[c][done / 2].y, vacc2[c][done / 2].y = 0.0f;
                            } else {
                                o = vacc2[c][done / 2].x, vacc2[c][done / 2].x = 0.0f;
                            }
                            // round-half-even through the 1.5*2^23 trick; after grayscale only the byte (the low
                            // mantissa bits of the biased sum) is needed, so the subtraction is skipped
                            // the quantised byte sits in the low mantissa bits of (o + 1.5*2^23)
                            v[c] = __fadd_rn(o, 12582912.0f);
                        }
                        {
                            float I;
                            if (GRAY) {
                                I = intLut[__float_as_uint(v[0]) & 0xffu];
                            } else {
                                // r^2 + g^2 + b^2 of the three quantised bytes: pack them (two PRMT) and take one byte dot
                                // product; the radicand is an integer < 2^18, so its conversion is exact and equals the
                                // float sum of squares the oracle forms
                                const uint32_t rg = __byte_perm(__float_as_uint(v[0]), __float_as_uint(v[C::CH > 1 ? 1 : 0]), 0x7740);
                                // (byte 1 of the biased float's bits 0x4B4000qq is zero: it fills the fourth lane)
                                const uint32_t rgb = __byte_perm(rg, __float_as_uint(v[C::CH - 1]), 0x5410);
                                I = sqrt_rn_inrange0((float)__dp4a(rgb, rgb, 0u));
                            }
                            ist[j * C::IS] = I;
#if NORM_C2_CPAIR
                            if (j == C::RB - 2) icarry0 = I;
                            if (j == C::RB - 1) icarry1 = I;
#endif
                        }
                    }
                }
                __threadfence_block();
                bar_arrive(C::BEMPTY + slot, 2 * kC2Group);   // done reading the H planes
                bar_arrive(C::BIFULL + islot, 2 * kC2Group);   // intensity rows of this batch are ready
            }
        }
    } else {
        // =========================== group C: Sobel + store + SSIM ===========================
        if constexpr (MODE == 1) {  // RGBA blur: the V pass of channels 2 and 3 (after grayscale group C never gets here)
            if constexpr (!C::BIG) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(C::REGS_V));
            blur_only_consumer<C, R, GRAY, 2, 2, kABParties>(p, Hb_base, tid - 2 * kC2Group, items_per_frame, C2_TOTAL_ITEMS());
            return;
        }
        if constexpr (!C::BIG) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(C::REGS_C));
#if NORM_C2_CPAIR
        // Pair version.  A thread owns TWO adjacent output columns (one 8-byte store, one 8-byte source load per row);
        // warps 0/1 of the group take the column halves of batch rows [0, JH), warps 2/3 those of rows [JH, RB), so all
        // four SM sub-partitions carry the same load.  Per-row overhead (pointer steps, loop control) is paid once per two
        // pixels, and batches that lie wholly inside the band run a fully unrolled body with immediate shared-memory
        // offsets; the first two and the last batch of a strip go through a generic per-row path.
        const int tc = tid - 2 * kC2Group;
        const int lane = tc & 31, warp = tc >> 5;
        const int cp = (warp & 1) * 32 + lane;           // column pair of the strip: columns 2cp, 2cp + 1
        const bool act = cp < C::SW / 2;
        const int cpa = act ? cp : C::SW / 2 - 1;        // idle lanes shadow the last pair: valid addresses, results dropped
        constexpr int JH = 3 * ((C::RB + 5) / 6);          // the longer share of a batch's rows, a multiple of three
        // shared-memory address of lumaLutD minus 8 * (bits of 1.5*2^23), modulo 2^32: lutd + 8*bits = &lumaLutD[byte]
        uint32_t lutd = (uint32_t)__cvta_generic_to_shared(lumaLutD) - (kMagicBits << 3);
        asm volatile("" : "+r"(lutd));

        for (int sl = 0; sl < C::NSI; ++sl) bar_arrive(C::BIEMPTY + sl, 2 * kC2Group);  // every intensity slot starts out empty

        for (uint32_t item = blockIdx.x, total = C2_TOTAL_ITEMS(); item < total; item += gridDim.x) {
            C2_ITEM_GEOMETRY();
            if (c2_empty) {  // (balanced partition) nothing to do: the finalize pass still sums this slot
                if (p.do_ssim && tc < 5) p.partials[(size_t)item * 5 + tc] = 0.0;
                continue;
            }
            const int x = x0 + 2 * cpa;                   // first column of the pair
            const bool full_strip = x0 + C::SW <= W && skip == 0;  // every pair of the strip lies inside the image and is this strip's own
            // intensity columns (index = blurred column + 1): the pair at ic, ic + 1; ClampToEdge on the blurred image
            const int ic = 2 * cpa + 2;
            const int il = (x - 1 < 0) ? ic : ic - 1;
            const int ir = (x + 2 > W - 1) ? ic + 1 : ic + 2;
            const int xs = min(x, W - 1);                 // ragged strips: keep the source / output addresses inside the row
            const uint8_t *in_col = p.in.ptr + (size_t)f * p.in.frame_stride + (size_t)xs * 4;
            uint8_t *out_col = p.out.ptr + (size_t)f * p.out.frame_stride + (size_t)xs * 4;
            double m5[5] = {0, 0, 0, 0, 0};

            // SSIM moments of one pixel: x = luma of the (c,c,c) result (f64 table), y = f32 luma of the source pixel
            auto ssim_px = [&](uint32_t cb, uint32_t spx) {
                double xv;
                asm("ld.shared.f64 %0, [%1];" : "=d"(xv) : "r"(lutd + (cb << 3)));
                const double yv = (double)luma_ssim(spx);
                m5[0] += xv;
                m5[1] += yv;
                m5[2] = fma(xv, xv, m5[2]);
                m5[3] = fma(yv, yv, m5[3]);
                m5[4] = fma(xv, yv, m5[4]);
            };

            for (int b = 0; b < nbatch; ++b, ++g) {
                const int islot = (int)(g % (uint32_t)C::NSI);
                bar_sync(C::BIFULL + islot, 2 * kC2Group);
                const float *ist = Ist_base + islot * kIstFloats + C::ICARRY * C::IS;  // batch row j at ist + j*IS
                const int n0 = b * C::RB - (R - 1);   // blurred row (relative to o_lo) that batch row 0 completes
                const int yfirst = o_lo + n0 - 1;     // output row emitted when batch row 0 arrives
                // this warp's rows [JA, JB) of the batch; the two warp pairs swap the longer share from batch to batch
                const bool first_half = ((warp >> 1) ^ (b & 1)) == 0;
                const int JA = first_half ? 0 : JH, JB = first_half ? JH : C::RB;
                // Rows [FA, FB) of the share run the fast body: their three window rows exist (n >= 2), nothing clamps,
                // the row is not the band's last (n < NB - 1) and the source rows fetched ahead (three below) exist;
                // every row of the batch lies in the SSIM row range.  The rows before and after go through the generic code.
                const bool fast_ok = full_strip && (!p.do_ssim || (yfirst >= (int)p.ssim_y0 && yfirst + C::RB <= (int)p.ssim_y1));
                const int FA = fast_ok ? min(max(JA, 2 - n0), JB) : JB, FB = fast_ok ? max(min(JB, NB - 4 - n0), FA) : JB;
                const bool fast = FA < FB;
                if (fast) {
                    // ---- interior batch: three rows per trip (the 3x3 window rotates statically), window rows JA-2, JA-1
                    // primed from shared memory, shared-memory offsets immediate, source pixels three rows ahead
                    auto body = [&](auto ssim_tag) {
                        constexpr bool SSIM = decltype(ssim_tag)::value;
                        const float *pl = ist + FA * C::IS + il, *pc = ist + FA * C::IS + ic, *pr = ist + FA * C::IS + ir;
                        float wl[3], wa[3], wb[3], wr[3];  // window rows: (left, pair, right)
#pragma unroll
                        for (int q = 0; q < 2; ++q) {
                            const float2 cc = *reinterpret_cast<const float2 *>(pc + (q - 2) * C::IS);
                            wl[q + 1] = pl[(q - 2) * C::IS], wa[q + 1] = cc.x, wb[q + 1] = cc.y, wr[q + 1] = pr[(q - 2) * C::IS];
                        }
                        const int y = yfirst + FA;
                        const uint8_t *ip = in_col + (size_t)y * p.in.pitch;
                        uint8_t *op = out_col + (size_t)y * p.out.pitch;
                        uint2 sp[3];
                        if (SSIM) {
#pragma unroll
                            for (int q = 0; q < 3; ++q) {
                                sp[q] = *reinterpret_cast<const uint2 *>(ip);
                                ip += p.in.pitch;
                            }
                        }
                        int j = FA;
                        for (; j + 3 <= FB; j += 3) {
#pragma unroll
                            for (int jj = 0; jj < 3; ++jj) {
                                // rows: new = slot jj, previous = slot (jj+2)%3, the one before = slot (jj+1)%3
                                constexpr int dummy = 0;
                                (void)dummy;
                                const int t = jj, m = (jj + 2) % 3, bt = (jj + 1) % 3;
                                const float2 cc = *reinterpret_cast<const float2 *>(pc + jj * C::IS);
                                wl[t] = pl[jj * C::IS], wa[t] = cc.x, wb[t] = cc.y, wr[t] = pr[jj * C::IS];
                                const uint32_t cb0 = sobel_mag_bits(wl[t], wa[t], wb[t], wl[m], wb[m], wl[bt], wa[bt], wb[bt]);
                                const uint32_t cb1 = sobel_mag_bits(wa[t], wb[t], wr[t], wa[m], wr[m], wa[bt], wb[bt], wr[bt]);
                                if (act) *reinterpret_cast<uint2 *>(op) = make_uint2(__byte_perm(cb0, 0xffu, 0x4000), __byte_perm(cb1, 0xffu, 0x4000));
                                op += p.out.pitch;
                                if (SSIM) {
                                    const uint2 s2 = sp[jj];
                                    sp[jj] = *reinterpret_cast<const uint2 *>(ip);
                                    ip += p.in.pitch;
                                    ssim_px(cb0, s2.x);
                                    ssim_px(cb1, s2.y);
                                }
                            }
                            pl += 3 * C::IS, pc += 3 * C::IS, pr += 3 * C::IS;
                        }
                        // rows left over when the share is not a multiple of three: all three window rows from shared memory
                        for (; j < FB; ++j) {
                            const float *rt = ist + j * C::IS, *rm = rt - C::IS, *rb = rm - C::IS;
                            const int yy = yfirst + j;
                            const float2 ct = *reinterpret_cast<const float2 *>(rt + ic), cm = *reinterpret_cast<const float2 *>(rm + ic),
                                         cbm = *reinterpret_cast<const float2 *>(rb + ic);
                            const uint32_t cb0 = sobel_mag_bits(rt[il], ct.x, ct.y, rm[il], cm.y, rb[il], cbm.x, cbm.y);
                            const uint32_t cb1 = sobel_mag_bits(ct.x, ct.y, rt[ir], cm.x, rm[ir], cbm.x, cbm.y, rb[ir]);
                            if (act) *reinterpret_cast<uint2 *>(out_col + (size_t)yy * p.out.pitch) =
                                    make_uint2(__byte_perm(cb0, 0xffu, 0x4000), __byte_perm(cb1, 0xffu, 0x4000));
                            if (SSIM) {
                                const uint2 s2 = *reinterpret_cast<const uint2 *>(in_col + (size_t)yy * p.in.pitch);
                                ssim_px(cb0, s2.x);
                                ssim_px(cb1, s2.y);
                            }
                        }
                    };
                    if (p.do_ssim) body(std::true_type{});
                    else body(std::false_type{});
                }
                if (FA > JA || FB < JB) {
                    // ---- generic path: every row checks for itself what it completes (band / image edges, SSIM range)
                    for (int j = JA; j < JB; ++j) {
                        if (j == FA) j = FB;  // (FA == FB == JB when nothing ran fast)
                        if (j >= JB) break;
                        const int n = n0 + j;
                        if (n < 0 || n >= NB) continue;
                        const float *rt = ist + j * C::IS, *rm = rt - C::IS, *rb = (o_lo + n - 2 < 0) ? rm : rm - C::IS;
                        // one or two emissions: row y = o_lo + n - 1, and at the bottom of the image also row y + 1,
                        // whose "row below" clamps to itself
                        for (int e = 0; e < 2; ++e) {
                            const int y = o_lo + n - 1 + e;
                            if (e == 1) {
                                if (!(yb == Himg && n == NB - 1)) break;
                                rb = (o_lo + n - 1 < 0) ? rt : rm, rm = rt;  // a one-row image: the row above clamps too
                            }
                            if (y < ya || (e == 0 && n == 0)) continue;
                            const bool ss_row = p.do_ssim && (uint32_t)(y - (int)p.ssim_y0) < p.ssim_y1 - p.ssim_y0;
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
                                const int xx = x + h;
                                if (!act || xx >= W || 2 * cpa + h < skip) continue;
                                const int c0 = ic + h;
                                const int cl = (xx - 1 < 0) ? c0 : c0 - 1, cr = (xx + 1 > W - 1) ? c0 : c0 + 1;
                                const uint32_t cb = sobel_mag_bits(rt[cl], rt[c0], rt[cr], rm[cl], rm[cr], rb[cl], rb[c0], rb[cr]);
                                *reinterpret_cast<uint32_t *>(out_col + (size_t)y * p.out.pitch + 4 * h) = __byte_perm(cb, 0xffu, 0x4000);
                                if (ss_row) ssim_px(cb, *reinterpret_cast<const uint32_t *>(in_col + (size_t)y * p.in.pitch + 4 * h));
                            }
                        }
                    }
                }
                __threadfence_block();
                bar_arrive(C::BIEMPTY + islot, 2 * kC2Group);
            }
            if (p.do_ssim) {
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const double sm = warp_sum(act ? m5[k] : 0.0);
                    if (lane == 0) red[k][warp] = sm;
                }
                bar_sync(C::BC, kC2Group);
                if (tc < 5) p.partials[(size_t)item * 5 + tc] = ((red[tc][0] + red[tc][1]) + red[tc][2]) + red[tc][3];
                bar_sync(C::BC, kC2Group);
            }
        }
#else
        const int tc = tid - 2 * kC2Group;  // output column of the strip
        const int lane = tc & 31, warp = tc >> 5;
        // shared-memory address of lumaLut minus 4 * (bits of 1.5*2^23), modulo 2^32: lut_biased + 4*bits = &lumaLut[byte]
        uint32_t lut_biased = (uint32_t)__cvta_generic_to_shared(lumaLut) - (kMagicBits << 2);
        asm volatile("" : "+r"(lut_biased));  // keep it in a register: one LEA per lookup instead of LEA + add

        for (int sl = 0; sl < C::NSI; ++sl) bar_arrive(C::BIEMPTY + sl, 2 * kC2Group);  // every intensity slot starts out empty

        for (uint32_t item = blockIdx.x, total = C2_TOTAL_ITEMS(); item < total; item += gridDim.x) {
            C2_ITEM_GEOMETRY();
            if (c2_empty) {  // (balanced partition) nothing to do: the finalize pass still sums this slot
                if (p.do_ssim && tc < 5) p.partials[(size_t)item * 5 + tc] = 0.0;
                continue;
            }
            const int x = x0 + tc;
            const bool writes = tc < C::SW && x < W && tc >= skip;
            const int xl = min(x, W - 1);  // keeps addresses valid on idle lanes
            // intensity column of x is tc+1; ClampToEdge: column -1 reads column 0, column W reads column W-1
            const int il = (x - 1 < 0) ? tc + 1 : tc, ic = min(tc + 1, C::IS - 1), ir = (x + 1 > W - 1) ? ic : min(tc + 2, C::IS - 1);
            // source pixels for the SSIM, three rows in flight (L2-resident: group A staged those rows a few
            // batches ago).  ipn walks down column xl one row per emitted row and stops at the last image row.
            const uint8_t *ipn = p.in.ptr + (size_t)f * p.in.frame_stride + (size_t)ya * p.in.pitch + (size_t)xl * 4;
            uint32_t *op = reinterpret_cast<uint32_t *>(p.out.ptr + (size_t)f * p.out.frame_stride + (size_t)ya * p.out.pitch) + xl;
            const bool ssim_on = p.do_ssim && writes;
            uint32_t sp0, sp1, sp2;
            sp0 = *reinterpret_cast<const uint32_t *>(ipn);
            if (ya + 1 < Himg) ipn += p.in.pitch;
            sp1 = *reinterpret_cast<const uint32_t *>(ipn);
            if (ya + 2 < Himg) ipn += p.in.pitch;
            sp2 = *reinterpret_cast<const uint32_t *>(ipn);
            if (ya + 3 < Himg) ipn += p.in.pitch;
// next source pixel; y already counts the row just emitted, the pixel fetched belongs to row y + 2
#define C2_NEXT_SRC(DST)                                   \
    DST = *reinterpret_cast<const uint32_t *>(ipn);        \
    if (y + 3 < Himg) ipn += p.in.pitch;
            double m5[5] = {0, 0, 0, 0, 0};
            // 3x3 window rows as (left, centre, right) triples; w0 = row o-2, w1 = row o-1 between steps
            float w0l = 0, w0c = 0, w0r = 0, w1l = 0, w1c = 0, w1r = 0, w2l, w2c, w2r;
            int y = ya;  // next output row

// one output row: Sobel(top, mid, bottom), store, SSIM against source pixel SPX.
// The quantised magnitude stays inside the bits of (c + 1.5*2^23): one PRMT builds the (c,c,c,255) pixel
// and the luma table is addressed as lut_biased + 4*bits (the constant prefix is folded into the base).
#define C2_EMIT(TL, TC, TR, ML, MR, BL, BC, BR, SPX, SSCOND)                                                \
    {                                                                                                       \
        const uint32_t spx = SPX;                                                                           \
        const uint32_t cb = sobel_mag_bits(TL, TC, TR, ML, MR, BL, BC, BR);                                 \
        if (writes) *op = __byte_perm(cb, 0xffu, 0x4000);                                                   \
        op = reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(op) + p.out.pitch);                   \
        if (SSCOND) {                                                                                       \
            float lx;                                                                                       \
            asm("ld.shared.f32 %0, [%1];" : "=f"(lx) : "r"(lut_biased + (cb << 2)));                        \
            const double xv = (double)lx, yv = (double)luma_ssim(spx);                                      \
            m5[0] += xv;                                                                                    \
            m5[1] += yv;                                                                                    \
            m5[2] = fma(xv, xv, m5[2]);                                                                     \
            m5[3] = fma(yv, yv, m5[3]);                                                                     \
            m5[4] = fma(xv, yv, m5[4]);                                                                     \
        }                                                                                                   \
        ++y;                                                                                                \
    }
#define C2_ROWTEST (ssim_on && (uint32_t)(y - (int)p.ssim_y0) < p.ssim_y1 - p.ssim_y0)

            for (int b = 0; b < nbatch; ++b, ++g) {
                const int islot = (int)(g % (uint32_t)C::NSI);
                // rows of this batch whose blurred output exists: n = b*RB + j - (R-1) in [0, NB)
                const int j_lo = max(0, (R - 1) - b * C::RB), j_hi = min(C::RB, NB + (R - 1) - b * C::RB);
                // first row of this batch that completes an OWNED output row (y = o-1 >= ya)
                const int j_e = max(j_lo, (ya + 1 - o_lo) + (R - 1) - b * C::RB);
                bar_sync(C::BIFULL + islot, 2 * kC2Group);
                const float *ist = Ist_base + islot * kIstFloats;
                int j = j_lo;
                // rows that only fill the window (at most two per item)
                for (; j < min(j_e, j_hi); ++j) {
                    const float *rp = ist + j * C::IS;
                    w2l = rp[il], w2c = rp[ic], w2r = rp[ir];
                    if (b * C::RB + j == R - 1) w1l = w2l, w1c = w2c, w1r = w2r;  // first blurred row: row above = itself
                    w0l = w1l, w0c = w1c, w0r = w1r;
                    w1l = w2l, w1c = w2c, w1r = w2r;
                }
                // main: three rows per trip, window registers rotate statically.  When every row this batch
                // emits lies inside the SSIM row range (always, outside band mode) the per-row test is dropped.
                if (y >= (int)p.ssim_y0 && y + (j_hi - j) <= (int)p.ssim_y1) {
                    for (; j + 3 <= j_hi; j += 3) {
                        const float *rp = ist + j * C::IS;
                        w2l = rp[il], w2c = rp[ic], w2r = rp[ir];
                        C2_EMIT(w2l, w2c, w2r, w1l, w1r, w0l, w0c, w0r, sp0, ssim_on);
                        C2_NEXT_SRC(sp0)
                        w0l = rp[C::IS + il], w0c = rp[C::IS + ic], w0r = rp[C::IS + ir];
                        C2_EMIT(w0l, w0c, w0r, w2l, w2r, w1l, w1c, w1r, sp1, ssim_on);
                        C2_NEXT_SRC(sp1)
                        w1l = rp[2 * C::IS + il], w1c = rp[2 * C::IS + ic], w1r = rp[2 * C::IS + ir];
                        C2_EMIT(w1l, w1c, w1r, w0l, w0r, w2l, w2c, w2r, sp2, ssim_on);
                        C2_NEXT_SRC(sp2)
                    }
                } else {
                    for (; j + 3 <= j_hi; j += 3) {
                        const float *rp = ist + j * C::IS;
                        w2l = rp[il], w2c = rp[ic], w2r = rp[ir];
                        C2_EMIT(w2l, w2c, w2r, w1l, w1r, w0l, w0c, w0r, sp0, C2_ROWTEST);
                        C2_NEXT_SRC(sp0)
                        w0l = rp[C::IS + il], w0c = rp[C::IS + ic], w0r = rp[C::IS + ir];
                        C2_EMIT(w0l, w0c, w0r, w2l, w2r, w1l, w1c, w1r, sp1, C2_ROWTEST);
                        C2_NEXT_SRC(sp1)
                        w1l = rp[2 * C::IS + il], w1c = rp[2 * C::IS + ic], w1r = rp[2 * C::IS + ir];
                        C2_EMIT(w1l, w1c, w1r, w0l, w0r, w2l, w2c, w2r, sp2, C2_ROWTEST);
                        C2_NEXT_SRC(sp2)
                    }
                }
                // remainder rows
                for (; j < j_hi; ++j) {
                    const float *rp = ist + j * C::IS;
                    w2l = rp[il], w2c = rp[ic], w2r = rp[ir];
                    C2_EMIT(w2l, w2c, w2r, w1l, w1r, w0l, w0c, w0r, sp0, C2_ROWTEST);
                    sp0 = sp1, sp1 = sp2;
                    C2_NEXT_SRC(sp2)
                    w0l = w1l, w0c = w1c, w0r = w1r;
                    w1l = w2l, w1c = w2c, w1r = w2r;
                }
                __threadfence_block();
                bar_arrive(C::BIEMPTY + islot, 2 * kC2Group);
            }
            // bottom edge of the image: row H-1 reads row H (clamped) = row H-1
            if (yb == Himg) C2_EMIT(w1l, w1c, w1r, w1l, w1r, w0l, w0c, w0r, sp0, C2_ROWTEST);
#undef C2_EMIT
#undef C2_ROWTEST
#undef C2_NEXT_SRC
            if (p.do_ssim) {
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const double s = warp_sum(m5[k]);
                    if (lane == 0) red[k][warp] = s;
                }
                bar_sync(C::BC, kC2Group);
                if (tc < 5) p.partials[(size_t)item * 5 + tc] = ((red[tc][0] + red[tc][1]) + red[tc][2]) + red[tc][3];
                bar_sync(C::BC, kC2Group);
            }
        }
#endif  // NORM_C2_CPAIR
    }
#undef C2_ITEM_GEOMETRY
#undef C2_TOTAL_ITEMS
}

}  // namespace normk
