/*
 * cc_dev.cuh — device-side helpers and constant tables shared by the kernels of libceracuda (sm_100a).
 * Integer pipes only: byte-SIMD (__vsadu4, __vavgu4), funnel shifts for unaligned windows, warp shuffles.
 * Tables restate ITU-T H.264 (clause at each table).
 */
#ifndef CC_DEV_CUH
#define CC_DEV_CUH
#include "cc_internal.h"
#ifndef CC_EMU
#include <cuda_runtime.h>
#define CC_LAUNCH(kern, grid, block, stream, ...) kern<<<grid, block, 0, stream>>>(__VA_ARGS__)
#define CC_LAUNCH_SMEM(kern, grid, block, smem, stream, ...) kern<<<grid, block, smem, stream>>>(__VA_ARGS__)
#define CC_DYN_SMEM(name) extern __shared__ __align__(16) uint8_t name[]
#ifndef CC_SPIN_NS
#define CC_SPIN_NS 20
#endif
#define CC_SPIN_PAUSE() __nanosleep(CC_SPIN_NS)
#else
#define CC_LAUNCH(kern, grid, block, stream, ...) emu::launch(grid, block, 0, [&] { kern(__VA_ARGS__); })
#define CC_LAUNCH_SMEM(kern, grid, block, smem, stream, ...) emu::launch(grid, block, smem, [&] { kern(__VA_ARGS__); })
#define CC_DYN_SMEM(name) uint8_t *name = (uint8_t *)(((uintptr_t)emu::dyn_smem() + 15) & ~(uintptr_t)15)
#define CC_SPIN_PAUSE() __nanosleep(20)
#endif

#define FULL 0xffffffffu

/* ---- tables ------------------------------------------------------------------------------------- */
/* zig-zag scan (H.264 Table 8-2, frame): scan position -> raster index, and its inverse */
__constant__ uint8_t c_zigzag[16] = {0, 1, 4, 8, 5, 2, 3, 6, 9, 12, 13, 10, 7, 11, 14, 15};
__constant__ uint8_t c_unzigzag[16] = {0, 1, 5, 6, 2, 4, 7, 12, 3, 8, 11, 13, 9, 10, 14, 15};
/* luma 4x4 block index (decoding order, Figure 6-10) -> x,y in units of 4 pixels */
__constant__ uint8_t c_blk_x[16] = {0, 1, 0, 1, 2, 3, 2, 3, 0, 1, 0, 1, 2, 3, 2, 3};
__constant__ uint8_t c_blk_y[16] = {0, 0, 1, 1, 0, 0, 1, 1, 2, 2, 3, 3, 2, 2, 3, 3};
/* (by,bx) -> block index */
__constant__ uint8_t c_blk_of[16] = {0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15};
/* normAdjust4x4 (8.5.9): [qp%6][class], class 0 = (even,even), 1 = (odd,odd), 2 = mixed */
__constant__ uint8_t c_dequant_v[6][3] = {{10, 16, 13}, {11, 18, 14}, {13, 20, 16}, {14, 23, 18}, {16, 25, 20}, {18, 29, 23}};
/* chroma qp (Table 8-15) */
__constant__ uint8_t c_chroma_qp[52] = {0,  1,  2,  3,  4,  5,  6,  7,  8,  9,  10, 11, 12, 13, 14, 15, 16, 17,
                                        18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 29, 30, 31, 32, 32, 33,
                                        34, 34, 35, 35, 36, 36, 37, 37, 37, 38, 38, 38, 39, 39, 39, 39};
/* SATD-domain lambda per qp: round(2^((qp-12)/6)), at least 1 */
__constant__ uint8_t c_lambda[52] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1,  1,  1,  1,  1,  1,  1,  2,  2,
                                     2, 2, 3, 3, 3, 4, 4, 4, 5, 6,  6,  7,  8,  9,  10, 11, 13, 14,
                                     16, 18, 20, 23, 25, 29, 32, 36, 40, 45, 51, 57, 64, 72, 81, 91};
/* deblocking thresholds (Tables 8-16, 8-17) */
__constant__ uint8_t c_alpha[52] = {0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,  0,   0,   0,   0,   4,   4,
                                    5,  6,  7,  8,  9,  10, 12, 13, 15, 17, 20, 22, 25,  28,  32,  36,  40,  45,
                                    50, 56, 63, 71, 80, 90, 101, 113, 127, 144, 162, 182, 203, 226, 255, 255};
__constant__ uint8_t c_beta[52] = {0, 0, 0, 0, 0, 0, 0, 0, 0,  0,  0,  0,  0,  0,  0,  0,  2,  2,
                                   2, 3, 3, 3, 3, 4, 4, 4, 6,  6,  7,  7,  8,  8,  9,  9,  10, 10,
                                   11, 11, 12, 12, 13, 13, 14, 14, 15, 15, 16, 16, 17, 17, 18, 18};
__constant__ uint8_t c_tc0[52][3] = {
    {0, 0, 0},  {0, 0, 0},  {0, 0, 0},  {0, 0, 0},   {0, 0, 0},   {0, 0, 0},   {0, 0, 0},   {0, 0, 0},   {0, 0, 0},
    {0, 0, 0},  {0, 0, 0},  {0, 0, 0},  {0, 0, 0},   {0, 0, 0},   {0, 0, 0},   {0, 0, 0},   {0, 0, 0},   {0, 0, 1},
    {0, 0, 1},  {0, 0, 1},  {0, 0, 1},  {0, 1, 1},   {0, 1, 1},   {1, 1, 1},   {1, 1, 1},   {1, 1, 1},   {1, 1, 1},
    {1, 1, 2},  {1, 1, 2},  {1, 1, 2},  {1, 1, 2},   {1, 2, 3},   {1, 2, 3},   {2, 2, 3},   {2, 2, 4},   {2, 3, 4},
    {2, 3, 4},  {3, 3, 5},  {3, 4, 6},  {3, 4, 6},   {4, 5, 7},   {4, 5, 8},   {4, 6, 9},   {5, 7, 10},  {6, 8, 11},
    {6, 8, 13}, {7, 10, 14}, {8, 11, 16}, {9, 12, 18}, {10, 13, 20}, {11, 15, 23}, {13, 17, 25}};
/* decimation weights by run length [EXT: x264_decimate_table4] */
__constant__ uint8_t c_decimate[16] = {3, 2, 2, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};

/* Quarter-pel interpolation as the average of two samples taken from the four planes
 * G (integer), b (horizontal half), h (vertical half), j (centre) — H.264 8.4.2.2.1, Figure 8-4.
 * entry [fy*4+fx] = planeA | dxA<<2 | dyA<<3 | planeB<<4 | dxB<<6 | dyB<<7 ; plane 0 G, 1 b, 2 h, 3 j */
#define QP_E(pa, xa, ya, pb, xb, yb) ((pa) | ((xa) << 2) | ((ya) << 3) | ((pb) << 4) | ((xb) << 6) | ((yb) << 7))
__constant__ uint8_t c_qpel[16] = {
    QP_E(0, 0, 0, 0, 0, 0), QP_E(0, 0, 0, 1, 0, 0), QP_E(1, 0, 0, 1, 0, 0), QP_E(1, 0, 0, 0, 1, 0),
    QP_E(0, 0, 0, 2, 0, 0), QP_E(1, 0, 0, 2, 0, 0), QP_E(1, 0, 0, 3, 0, 0), QP_E(1, 0, 0, 2, 1, 0),
    QP_E(2, 0, 0, 2, 0, 0), QP_E(2, 0, 0, 3, 0, 0), QP_E(3, 0, 0, 3, 0, 0), QP_E(3, 0, 0, 2, 1, 0),
    QP_E(2, 0, 0, 0, 0, 1), QP_E(2, 0, 0, 1, 0, 1), QP_E(3, 0, 0, 1, 0, 1), QP_E(2, 1, 0, 1, 0, 1)};

/* ---- small integer helpers ---------------------------------------------------------------------- */
__device__ __forceinline__ int cc_iabs(int x) { return x < 0 ? -x : x; }
__device__ __forceinline__ int cc_clip255(int x) { return x < 0 ? 0 : (x > 255 ? 255 : x); }
__device__ __forceinline__ int cc_clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }
__device__ __forceinline__ int cc_median3(int a, int b, int c) {
    int mn = a < b ? a : b, mx = a < b ? b : a;
    return c < mn ? mn : (c > mx ? mx : c);
}
/* bit length of se(v), H.264 9.1 */
__device__ __forceinline__ int cc_se_bits(int d) {
    unsigned k = d > 0 ? 2u * (unsigned)d : (d < 0 ? 2u * (unsigned)(-d) + 1u : 1u);
    return 2 * (31 - __clz((int)k)) + 1;
}
__device__ __forceinline__ int cc_pos_class(int i) { /* raster index in a 4x4 -> quant class */
    int x = i & 1, y = (i >> 2) & 1;
    return (x & y) ? 1 : ((x | y) ? 2 : 0);
}
/* 32-bit load from an arbitrary byte address: two aligned words + funnel shift */
__device__ __forceinline__ uint32_t cc_ld32u(const uint8_t *p) {
    uintptr_t a = (uintptr_t)p;
    const uint32_t *q = (const uint32_t *)(a & ~(uintptr_t)3);
    return __funnelshift_r(q[0], q[1], (unsigned)(a & 3) * 8u);
}
/* 8 bytes from an arbitrary byte address: three aligned words */
__device__ __forceinline__ void cc_ld64u(const uint8_t *p, uint32_t &lo, uint32_t &hi) {
    uintptr_t a = (uintptr_t)p;
    const uint32_t *q = (const uint32_t *)(a & ~(uintptr_t)3);
    unsigned sh = (unsigned)(a & 3) * 8u;
    uint32_t w0 = q[0], w1 = q[1], w2 = q[2];
    lo = __funnelshift_r(w0, w1, sh);
    hi = __funnelshift_r(w1, w2, sh);
}
/* acc + sum of |a.b[i] - b.b[i]| over the four bytes: one VABSDIFF4.U8.ACC (the __vsadu4 intrinsic leaves the add separate) */
__device__ __forceinline__ uint32_t cc_sad4_acc(uint32_t a, uint32_t b, uint32_t acc) {
#ifdef CC_EMU_BUILD
    return acc + __vsadu4(a, b);
#else
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(acc));
    return d;
#endif
}
__device__ __forceinline__ int cc_byte(uint32_t w, int i) { return (int)((w >> (8 * i)) & 255u); }

/* c + sum over the four bytes of (unsigned byte of a) * (signed byte of b): one IDP.4A.U8.S8 */
__device__ __forceinline__ int cc_dp4a_us(uint32_t a, uint32_t b, int c) {
#ifdef CC_EMU_BUILD
    for (int i = 0; i < 4; i++) c += (int)((a >> (8 * i)) & 255u) * (int)(int8_t)((b >> (8 * i)) & 255u);
    return c;
#else
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#endif
}
/* c + (signed low half of a) * (signed byte 0 of b) + (signed high half of a) * (signed byte 1 of b): IDP.2A.LO; _hi: bytes 2 and 3 of b */
__device__ __forceinline__ int cc_dp2a_lo(uint32_t a, uint32_t b, int c) {
#ifdef CC_EMU_BUILD
    return c + (int)(int16_t)(a & 0xffffu) * (int)(int8_t)(b & 255u) + (int)(int16_t)(a >> 16) * (int)(int8_t)((b >> 8) & 255u);
#else
    int d;
    asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#endif
}
__device__ __forceinline__ int cc_dp2a_hi(uint32_t a, uint32_t b, int c) {
#ifdef CC_EMU_BUILD
    return c + (int)(int16_t)(a & 0xffffu) * (int)(int8_t)((b >> 16) & 255u) + (int)(int16_t)(a >> 16) * (int)(int8_t)(b >> 24);
#else
    int d;
    asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#endif
}
/* rows of the 4-point Hadamard matrix and their negatives as packed signed bytes (pixel 0 = low byte) */
#define CC_H0 0x01010101u /* + + + + */
#define CC_H1 0xffff0101u /* + + - - */
#define CC_H2 0x01ffff01u /* + - - + */
#define CC_H3 0xff01ff01u /* + - + - */
#define CC_NEG4(h) ((~(h) & 0xfefefefeu) | 0x01010101u) /* bytes are +1 (0x01) or -1 (0xff): negation swaps them */

/* one 4x4 SATD (Hadamard of the difference, sum|.| >> 1) from packed rows: c[r], p[r] = 4 pixels each.  The horizontal
 * transform of the difference comes straight off the packed bytes as dot products with +-1 rows (2 IDP.4A per coefficient). */
__device__ __forceinline__ int cc_satd4x4_words(const uint32_t c[4], const uint32_t p[4]) {
    int t[16];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        t[r * 4 + 0] = cc_dp4a_us(c[r], CC_H0, cc_dp4a_us(p[r], CC_NEG4(CC_H0), 0));
        t[r * 4 + 1] = cc_dp4a_us(c[r], CC_H1, cc_dp4a_us(p[r], CC_NEG4(CC_H1), 0));
        t[r * 4 + 2] = cc_dp4a_us(c[r], CC_H2, cc_dp4a_us(p[r], CC_NEG4(CC_H2), 0));
        t[r * 4 + 3] = cc_dp4a_us(c[r], CC_H3, cc_dp4a_us(p[r], CC_NEG4(CC_H3), 0));
    }
    /* second butterfly stage folded into the sum: |x + y| + |x - y| = 2 max(|x|, |y|), and the 2 is the SATD's final halving */
    int s = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int s01 = t[i] + t[4 + i], m01 = t[i] - t[4 + i], s23 = t[8 + i] + t[12 + i], m23 = t[8 + i] - t[12 + i];
        s += max(cc_iabs(s01), cc_iabs(s23)) + max(cc_iabs(m01), cc_iabs(m23));
    }
    return s;
}
/* The same with the source block's half of the horizontal transform taken out of the loop over candidates: the Hadamard transform is
 * linear, so H(c - p) = H(c) + H(-p).  tc[r * 4 + k] = row r of the source block times Hadamard row k (cc_satd4x4_src), and every
 * candidate costs 16 dot products instead of 32. */
__device__ __forceinline__ void cc_satd4x4_src(const uint32_t c[4], int tc[16]) {
#pragma unroll
    for (int r = 0; r < 4; r++) {
        tc[r * 4 + 0] = cc_dp4a_us(c[r], CC_H0, 0); tc[r * 4 + 1] = cc_dp4a_us(c[r], CC_H1, 0);
        tc[r * 4 + 2] = cc_dp4a_us(c[r], CC_H2, 0); tc[r * 4 + 3] = cc_dp4a_us(c[r], CC_H3, 0);
    }
}
__device__ __forceinline__ int cc_satd4x4_pre(const int tc[16], const uint32_t p[4]) {
    int t[16];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        t[r * 4 + 0] = cc_dp4a_us(p[r], CC_NEG4(CC_H0), tc[r * 4 + 0]);
        t[r * 4 + 1] = cc_dp4a_us(p[r], CC_NEG4(CC_H1), tc[r * 4 + 1]);
        t[r * 4 + 2] = cc_dp4a_us(p[r], CC_NEG4(CC_H2), tc[r * 4 + 2]);
        t[r * 4 + 3] = cc_dp4a_us(p[r], CC_NEG4(CC_H3), tc[r * 4 + 3]);
    }
    /* second butterfly stage folded into the sum: |x + y| + |x - y| = 2 max(|x|, |y|), and the 2 is the SATD's final halving */
    int s = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int s01 = t[i] + t[4 + i], m01 = t[i] - t[4 + i], s23 = t[8 + i] + t[12 + i], m23 = t[8 + i] - t[12 + i];
        s += max(cc_iabs(s01), cc_iabs(s23)) + max(cc_iabs(m01), cc_iabs(m23));
    }
    return s;
}
#endif
