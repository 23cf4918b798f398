/*
 * cc_xform8.cuh — the 8x8 transform path of inter macroblocks (High profile, transform_size_8x8_flag = 1;
 * sub8x8_dct8 / quant_8x8 / dequant_8x8 / add8x8_idct8 [EXT] and pixel_sa8d for the
 * transform-size decision).
 *
 * Lane mapping: the warp holds the four 8x8 luma blocks of a macroblock, lane = blk8 * 8 + line.  A lane owns one
 * 8-sample line of its block in registers; the 1-D butterflies run in-lane, the row<->column transposes go through a
 * small shared-memory scratch (int32 [4][8][9], padded against bank conflicts).  Separable passes with intermediate
 * rounding are order-sensitive: the inverse runs rows first, then columns, as H.264 8.5.13 prescribes; the forward
 * transform (non-normative) runs rows then columns like the oracle (oracle/ref_prims.c ref_sub8x8_dct8).
 * CAVLC carries an 8x8 block as four interleaved 4x4 blocks (scan position k -> block k&3, index k>>2), so the levels
 * land in WarpMem::coef exactly where the 4x4 path puts them and symbol emission / host CAVLC are unchanged.
 */
#ifndef CC_XFORM8_CUH
#define CC_XFORM8_CUH
#include "cc_xform.cuh"

/* 8x8 frame zig-zag (H.264 Figure 8-9): scan position -> raster index */
__constant__ uint8_t c_zigzag8[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                      41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                      30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
/* the same table in global memory, for loads with a lane-dependent index (a constant-bank load would serialise them) */
__device__ const uint8_t g_zigzag8[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                          41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                          30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
/* normAdjust8x8 (8.5.9): [qp%6][class] */
__constant__ uint8_t c_dequant8_v[6][6] = {{20, 18, 32, 19, 25, 24}, {22, 19, 35, 21, 28, 26}, {26, 23, 42, 24, 33, 31},
                                           {28, 25, 45, 26, 35, 33}, {32, 28, 51, 30, 40, 38}, {36, 32, 58, 34, 46, 43}};
/* class of position (y,x): table[(y&3)*4 + (x&3)] = {0,3,4,3, 3,1,5,1, 4,5,2,5, 3,1,5,1}, packed 4 bits each */
#define CC_CLS8_PACKED 0x1513525415133430ull
__device__ __forceinline__ int cc_cls8(int y, int x) { return (int)((CC_CLS8_PACKED >> (4 * (((y & 3) << 2) | (x & 3)))) & 15ull); }

struct X8Mem {
    int32_t s[4][8][9];   /* transpose scratch */
    int16_t q[4][64];     /* quantised levels, raster */
    uint8_t zz[64];       /* c_zigzag8 copied to shared memory (lane-dependent lookups) */
};

__device__ __forceinline__ void dct8_1d(int v[8]) {
    const int a0 = v[0] + v[7], a1 = v[1] + v[6], a2 = v[2] + v[5], a3 = v[3] + v[4];
    const int a4 = v[0] - v[7], a5 = v[1] - v[6], a6 = v[2] - v[5], a7 = v[3] - v[4];
    const int b0 = a0 + a3, b1 = a1 + a2, b2 = a0 - a3, b3 = a1 - a2;
    const int b4 = a5 + a6 + ((a4 >> 1) + a4), b5 = a4 - a7 - ((a6 >> 1) + a6);
    const int b6 = a4 + a7 - ((a5 >> 1) + a5), b7 = a5 - a6 + ((a7 >> 1) + a7);
    v[0] = b0 + b1; v[1] = b4 + (b7 >> 2); v[2] = b2 + (b3 >> 1); v[3] = b5 + (b6 >> 2);
    v[4] = b0 - b1; v[5] = b6 - (b5 >> 2); v[6] = (b2 >> 1) - b3; v[7] = (b4 >> 2) - b7;
}
__device__ __forceinline__ void idct8_1d(int v[8]) { /* H.264 8.5.13 */
    const int a0 = v[0] + v[4], a2 = v[0] - v[4], a4 = (v[2] >> 1) - v[6], a6 = v[2] + (v[6] >> 1);
    const int a1 = -v[3] + v[5] - v[7] - (v[7] >> 1), a3 = v[1] + v[7] - v[3] - (v[3] >> 1);
    const int a5 = -v[1] + v[7] + v[5] + (v[5] >> 1), a7 = v[3] + v[5] + v[1] + (v[1] >> 1);
    const int b0 = a0 + a6, b2 = a2 + a4, b4 = a2 - a4, b6 = a0 - a6;
    const int b1 = a1 + (a7 >> 2), b3 = a3 + (a5 >> 2), b5 = (a3 >> 2) - a5, b7 = a7 - (a1 >> 2);
    v[0] = b0 + b7; v[1] = b2 + b5; v[2] = b4 + b3; v[3] = b6 + b1;
    v[4] = b6 - b1; v[5] = b4 - b3; v[6] = b2 - b5; v[7] = b0 - b7;
}
__device__ __forceinline__ void had8_1d(int v[8]) {
    const int a0 = v[0] + v[1], a1 = v[0] - v[1], a2 = v[2] + v[3], a3 = v[2] - v[3];
    const int a4 = v[4] + v[5], a5 = v[4] - v[5], a6 = v[6] + v[7], a7 = v[6] - v[7];
    const int b0 = a0 + a2, b1 = a1 + a3, b2 = a0 - a2, b3 = a1 - a3, b4 = a4 + a6, b5 = a5 + a7, b6 = a4 - a6, b7 = a5 - a7;
    v[0] = b0 + b4; v[1] = b1 + b5; v[2] = b2 + b6; v[3] = b3 + b7; v[4] = b0 - b4; v[5] = b1 - b5; v[6] = b2 - b6; v[7] = b3 - b7;
}
/* rows held by the lanes -> columns held by the lanes (and back): lane (b, l) writes line l, reads line-index l of every line */
__device__ __forceinline__ void x8_transpose(X8Mem *x, int b, int l, int v[8]) {
#pragma unroll
    for (int i = 0; i < 8; i++) x->s[b][l][i] = v[i];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = x->s[b][i][l];
    __syncwarp();
}
/* residual line of this lane: cur - pred for row (oy + l), columns ox..ox+7, from the 16-pixel-wide word arrays */
__device__ __forceinline__ void x8_residual(const uint32_t *curY, const uint32_t *predY, int b, int l, int d[8]) {
    const int row = (b >> 1) * 8 + l, w0 = row * 4 + (b & 1) * 2;
    const uint32_t c0 = curY[w0], c1 = curY[w0 + 1], p0 = predY[w0], p1 = predY[w0 + 1];
#pragma unroll
    for (int i = 0; i < 4; i++) { d[i] = cc_byte(c0, i) - cc_byte(p0, i); d[4 + i] = cc_byte(c1, i) - cc_byte(p1, i); }
}

/* [EXT] x264 pixel_sa8d over the macroblock's residual: (sum over the four 8x8 blocks of sum|Hadamard8x8| + 2) >> 2 */
__device__ __forceinline__ int sa8d_mb(X8Mem *x, const uint32_t *curY, const uint32_t *predY, int lane) {
    const int b = lane >> 3, l = lane & 7;
    int d[8];
    x8_residual(curY, predY, b, l, d);
    had8_1d(d);
    x8_transpose(x, b, l, d);
    had8_1d(d);
    int s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += cc_iabs(d[i]);
    s = __reduce_add_sync(FULL, s);
    return (s + 2) >> 2;
}

/* [EXT] x264_decimate_score64 from a 64-bit significance mask in scan order (bit k = scan position k non-zero) */
__device__ __forceinline__ int decimate_score64(unsigned long long nzm) {
    int score = 0;
    while (nzm) {
        const int hi = 63 - __clzll((long long)nzm);
        nzm &= ~(1ull << hi);
        const int run = nzm ? hi - (63 - __clzll((long long)nzm)) - 1 : hi;
        score += run < 4 ? 3 : (run < 12 ? 2 : (run < 32 ? 1 : 0));
    }
    return score;
}

/* Codes the luma of an inter macroblock with the 8x8 transform: levels -> coef[16][16] (interleaved 4x4 layout),
 * reconstruction -> recY (64 words).  Returns the luma cbp in all lanes; *nzmask gets all four 4x4 bits of every coded
 * 8x8 block (what the boundary-strength rule needs for 8x8-transform macroblocks, H.264 8.7.2.1). */
__device__ __forceinline__ int inter_luma8(const cc_kparams &k, X8Mem *x, const uint32_t *curY, const uint32_t *predY, uint32_t *recY,
                                           int16_t (*coef)[16], int qp, int lane, unsigned *nzmask, bool decimate = true) {
    const int b = lane >> 3, l = lane & 7;
    int d[8];
    x8_residual(curY, predY, b, l, d);
    dct8_1d(d);                 /* rows */
    x8_transpose(x, b, l, d);
    dct8_1d(d);                 /* columns: d[y] = coefficient (y, x = l) */
#pragma unroll
    for (int y = 0; y < 8; y++) {
        const int cls = cc_cls8(y, l);
        x->q[b][y * 8 + l] = (int16_t)quant_one(d[y], k.q8mf[cls], k.q8bias[cls]);
    }
    __syncwarp();
    /* scan order: this lane looks at scan positions 8l .. 8l+7 of its block */
    int z[8];
    unsigned m8 = 0, big = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        z[j] = x->q[b][x->zz[8 * l + j]];
        m8 |= (unsigned)(z[j] != 0) << j;
        big |= (unsigned)(cc_iabs(z[j]) > 1);
    }
    unsigned long long nzm = (unsigned long long)m8 << (8 * l);
#pragma unroll
    for (int s = 1; s < 8; s <<= 1) {
        nzm |= __shfl_xor_sync(FULL, nzm, s);
        big |= __shfl_xor_sync(FULL, big, s);
    }
    const int score = big ? 9 : decimate_score64(nzm);          /* identical in the 8 lanes of a block */
    int tot = score + __shfl_xor_sync(FULL, score, 8);
    tot += __shfl_xor_sync(FULL, tot, 16);
    const bool coded = nzm != 0ull && !(decimate && (tot < 6 || score < 4));  /* [EXT] x264 luma decimation thresholds */
    /* interleaved 4x4 layout: scan position p -> block (p & 3), index p >> 2 */
#pragma unroll
    for (int j = 0; j < 8; j++) coef[b * 4 + (j & 3)][2 * l + (j >> 2)] = coded ? (int16_t)z[j] : (int16_t)0;
    /* reconstruction (dead blocks ride along with zero levels: the transposes are warp-wide) */
    const int row = (b >> 1) * 8 + l, w0 = row * 4 + (b & 1) * 2;
    int v[8];
    const int qb = qp / 6 - 6, r6 = qp % 6;
#pragma unroll
    for (int i = 0; i < 8; i++) { /* row l of the block: coefficient (l, i) */
        const int ls = c_dequant8_v[r6][cc_cls8(l, i)] * 16, q = coded ? (int)x->q[b][l * 8 + i] : 0;
        v[i] = qb >= 0 ? (q * ls) << qb : (q * ls + (1 << (-qb - 1))) >> -qb;
    }
    idct8_1d(v);                /* horizontal first (8.5.13) */
    x8_transpose(x, b, l, v);
    idct8_1d(v);                /* v[y] = residual (y, x = l) before the final shift */
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = (v[i] + 32) >> 6;
    x8_transpose(x, b, l, v);   /* back to rows: v[i] = residual (y = l, x = i) */
    const uint32_t p0 = predY[w0], p1 = predY[w0 + 1];
    uint32_t o0 = 0, o1 = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        o0 |= (uint32_t)cc_clip255(cc_byte(p0, i) + v[i]) << (8 * i);
        o1 |= (uint32_t)cc_clip255(cc_byte(p1, i) + v[4 + i]) << (8 * i);
    }
    recY[w0] = o0; recY[w0 + 1] = o1;
    const unsigned bal = __ballot_sync(FULL, coded);
    int cbp = 0;
    unsigned nzb = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) if ((bal >> (8 * i)) & 1u) { cbp |= 1 << i; nzb |= 0xfu << (4 * i); }
    *nzmask = nzb;
    __syncwarp();
    return cbp;
}
#endif
