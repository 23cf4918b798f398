/*
 * cc_xform.cuh — the 4x4 transform chain as one lane computes it: residual -> forward core transform ->
 * quantisation (dead-zone form) -> level scaling -> inverse transform -> reconstruction, plus the DC
 * Hadamards and the decimation score.  Everything lives in one thread's registers (16 ints per block);
 * kernels choose which block a lane owns.  Restates H.264 8.5.10-8.5.14 and the libx264 C forms of
 * sub4x4_dct / quant_4x4 / dequant_4x4 / add4x4_idct / decimate_score [EXT]; oracle: oracle/ref_prims.c.
 */
#ifndef CC_XFORM_CUH
#define CC_XFORM_CUH
#include "cc_dev.cuh"

/* ---- transform chain (one 4x4 block per lane) ------------------------------------------------------- */
/* rows of the forward core transform Cf = [1 1 1 1; 2 1 -1 -2; 1 -1 -1 1; 1 -2 2 -1] as packed signed bytes, and their negatives */
#define CC_CF0 0x01010101u
#define CC_CF1 0xfeff0102u
#define CC_CF2 0x01ffff01u
#define CC_CF3 0xff02fe01u
#define CC_CF0N 0xffffffffu
#define CC_CF1N 0x0201fffeu
#define CC_CF2N 0xff0101ffu
#define CC_CF3N 0x01fe02ffu
__device__ __forceinline__ void dct4x4_res(const uint32_t c[4], const uint32_t p[4], int d[16]) {
    int t[16];
#pragma unroll
    for (int r = 0; r < 4; r++) { /* horizontal pass on the packed bytes: Cf_k . cur - Cf_k . pred */
        t[r * 4 + 0] = cc_dp4a_us(c[r], CC_CF0, cc_dp4a_us(p[r], CC_CF0N, 0));
        t[r * 4 + 1] = cc_dp4a_us(c[r], CC_CF1, cc_dp4a_us(p[r], CC_CF1N, 0));
        t[r * 4 + 2] = cc_dp4a_us(c[r], CC_CF2, cc_dp4a_us(p[r], CC_CF2N, 0));
        t[r * 4 + 3] = cc_dp4a_us(c[r], CC_CF3, cc_dp4a_us(p[r], CC_CF3N, 0));
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int s03 = t[i] + t[12 + i], s12 = t[4 + i] + t[8 + i], m03 = t[i] - t[12 + i], m12 = t[4 + i] - t[8 + i];
        d[i] = s03 + s12; d[4 + i] = 2 * m03 + m12; d[8 + i] = s03 - s12; d[12 + i] = m03 - 2 * m12;
    }
}
/* [EXT] x264 quant: sign(c) * min(((f + |c|) * mf) >> 16, 2047) — branch-free: magnitude once, sign restored by xor / subtract */
__device__ __forceinline__ int quant_one(int c, int mf, int f) {
    const int s = c >> 31;                       /* 0 or -1 */
    const int l = min(((f + cc_iabs(c)) * mf) >> 16, 2047);
    return (l ^ s) - s;
}
/* quantise raster coefficients in place; returns number of non-zero levels (skipping position 0 if ac_only) */
__device__ __forceinline__ int quant4x4(int d[16], const uint16_t mf[3], const uint16_t bias[3], bool ac_only) {
    int n = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const int cls = ((i & 1) && ((i >> 2) & 1)) ? 1 : (((i & 1) || ((i >> 2) & 1)) ? 2 : 0);
        if (ac_only && i == 0) { d[0] = 0; continue; }
        d[i] = quant_one(d[i], mf[cls], bias[cls]);
        n += d[i] != 0;
    }
    return n;
}
/* H.264 8.5.12.1: level scaling with a flat matrix, in place */
__device__ __forceinline__ void dequant4x4(int d[16], int qp) {
    const int qb = qp / 6 - 4, r = qp % 6;
    int v0 = c_dequant_v[r][0] * 16, v1 = c_dequant_v[r][1] * 16, v2 = c_dequant_v[r][2] * 16;
    if (qb >= 0) { /* qp >= 24 (uniform): the shift folds into the scale, one multiply per coefficient */
        v0 <<= qb; v1 <<= qb; v2 <<= qb;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const int cls = ((i & 1) && ((i >> 2) & 1)) ? 1 : (((i & 1) || ((i >> 2) & 1)) ? 2 : 0);
            d[i] *= cls == 0 ? v0 : (cls == 1 ? v1 : v2);
        }
    } else {
        const int sh = -qb, rnd = 1 << (sh - 1);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const int cls = ((i & 1) && ((i >> 2) & 1)) ? 1 : (((i & 1) || ((i >> 2) & 1)) ? 2 : 0);
            d[i] = (d[i] * (cls == 0 ? v0 : (cls == 1 ? v1 : v2)) + rnd) >> sh;
        }
    }
}
/* H.264 8.5.12.2 + 8.5.14: inverse transform, add prediction, clip */
__device__ __forceinline__ void idct4x4_add(const int d[16], const uint32_t p[4], uint32_t out[4]) {
    int t[16];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        int e0 = d[r * 4] + d[r * 4 + 2], e1 = d[r * 4] - d[r * 4 + 2];
        int e2 = (d[r * 4 + 1] >> 1) - d[r * 4 + 3], e3 = d[r * 4 + 1] + (d[r * 4 + 3] >> 1);
        t[r * 4] = e0 + e3; t[r * 4 + 1] = e1 + e2; t[r * 4 + 2] = e1 - e2; t[r * 4 + 3] = e0 - e3;
    }
    int o[16];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int e0 = t[i] + t[8 + i], e1 = t[i] - t[8 + i];
        int e2 = (t[4 + i] >> 1) - t[12 + i], e3 = t[4 + i] + (t[12 + i] >> 1);
        o[i] = e0 + e3; o[4 + i] = e1 + e2; o[8 + i] = e1 - e2; o[12 + i] = e0 - e3;
    }
#pragma unroll
    for (int r = 0; r < 4; r++) {
        uint32_t w = 0;
#pragma unroll
        for (int x = 0; x < 4; x++) w |= (uint32_t)cc_clip255(cc_byte(p[r], x) + ((o[r * 4 + x] + 32) >> 6)) << (8 * x);
        out[r] = w;
    }
}
/* raster -> scan order.  ac_only: z[0..14] = scan positions 1..15, z[15] = 0 */
__device__ __forceinline__ void to_scan(const int d[16], int z[16], bool ac_only) {
    int s[16];
    s[0] = d[0]; s[1] = d[1]; s[2] = d[4]; s[3] = d[8]; s[4] = d[5]; s[5] = d[2]; s[6] = d[3]; s[7] = d[6];
    s[8] = d[9]; s[9] = d[12]; s[10] = d[13]; s[11] = d[10]; s[12] = d[7]; s[13] = d[11]; s[14] = d[14]; s[15] = d[15];
    if (ac_only) {
#pragma unroll
        for (int i = 0; i < 15; i++) z[i] = s[i + 1];
        z[15] = 0;
    } else {
#pragma unroll
        for (int i = 0; i < 16; i++) z[i] = s[i];
    }
}
__device__ __forceinline__ void store_levels(int16_t *dst, const int z[16]) {
    uint32_t *w = (uint32_t *)dst;
#pragma unroll
    for (int i = 0; i < 8; i++) w[i] = ((uint32_t)z[2 * i] & 0xffffu) | ((uint32_t)z[2 * i + 1] << 16);
}
/* What a P_Skip probe needs of the quantiser: which scan positions hold a non-zero level (nzm) and which a level of magnitude
 * above 1 (big) — two compares per coefficient against the thresholds of cc_kparams.qthr instead of the multiply / shift / clamp /
 * sign of quant_one.  d: raster coefficients; bit k of the masks = scan position k. */
__device__ __forceinline__ void quant_masks(const int d[16], const uint16_t thr[3][2], unsigned &nzm, unsigned &big) {
    const int t1[3] = {thr[0][0], thr[1][0], thr[2][0]}, t2[3] = {thr[0][1], thr[1][1], thr[2][1]};
    nzm = 0; big = 0;
#pragma unroll
    for (int kk = 0; kk < 16; kk++) {
        const int i = (int)((0xfeb7adc963258410ull >> (4 * kk)) & 15u); /* raster index of scan position kk (zigzag) */
        const int cls = ((i & 1) && ((i >> 2) & 1)) ? 1 : (((i & 1) || ((i >> 2) & 1)) ? 2 : 0);
        const int a = cc_iabs(d[i]);
        nzm |= (unsigned)(a >= t1[cls]) << kk;
        big |= (unsigned)(a >= t2[cls]) << kk;
    }
}
/* decimate_score from the masks of quant_masks */
__device__ __forceinline__ int decimate_masks(unsigned nzm, unsigned big) {
    if (big) return 9;
    int score = 0;
    while (nzm) {
        int hi = 31 - __clz((int)nzm);
        nzm &= ~(1u << hi);
        int run = nzm ? hi - (31 - __clz((int)nzm)) - 1 : hi;
        score += run == 0 ? 3 : (run <= 2 ? 2 : (run <= 5 ? 1 : 0));
    }
    return score;
}
/* [EXT] x264_decimate_score over scan-ordered levels */
__device__ __forceinline__ int decimate_score(const int z[16]) {
    unsigned nzm = 0, big = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) { nzm |= (unsigned)(z[i] != 0) << i; big |= (unsigned)(cc_iabs(z[i]) > 1) << i; }
    if (big) return 9;
    int score = 0;
    while (nzm) {
        int hi = 31 - __clz((int)nzm);
        nzm &= ~(1u << hi);
        int run = nzm ? hi - (31 - __clz((int)nzm)) - 1 : hi;
        score += run == 0 ? 3 : (run <= 2 ? 2 : (run <= 5 ? 1 : 0));
    }
    return score;
}
__device__ __forceinline__ void hadamard4x4(int d[16], int round_half) {
    int t[16];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        int s01 = d[r * 4] + d[r * 4 + 1], m01 = d[r * 4] - d[r * 4 + 1], s23 = d[r * 4 + 2] + d[r * 4 + 3], m23 = d[r * 4 + 2] - d[r * 4 + 3];
        t[r * 4] = s01 + s23; t[r * 4 + 1] = s01 - s23; t[r * 4 + 2] = m01 - m23; t[r * 4 + 3] = m01 + m23;
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int s01 = t[i] + t[4 + i], m01 = t[i] - t[4 + i], s23 = t[8 + i] + t[12 + i], m23 = t[8 + i] - t[12 + i];
        d[i] = (s01 + s23 + round_half) >> round_half; d[4 + i] = (s01 - s23 + round_half) >> round_half;
        d[8 + i] = (m01 - m23 + round_half) >> round_half; d[12 + i] = (m01 + m23 + round_half) >> round_half;
    }
}

#endif
