/*
 * cc_k_mb.cu — macroblock analysis, mode decision, transform/quant/reconstruction and in-loop deblocking
 * (SURVEY.md §8a rows K4, K5, K6, K7, K8).
 *
 * Execution model.  A slice is a band of `rps` macroblock rows and shares nothing with other slices
 * (intra/MV prediction stop at slice edges; deblocking uses disable_deblocking_filter_idc=2), so
 * one CTA = one slice, one warp = one macroblock row.  Rows of a slice advance as a wavefront with a
 * 2-MB lag (top-right neighbour), synchronised through shared-memory progress counters — all waiting is
 * inside a CTA, whose warps are co-resident by construction.  After a CTA-wide barrier the same warps
 * run the deblocking filter over the slice in the same wavefront order (the filter is normatively
 * raster-ordered: H.264 8.7).  Many slices x many concurrent streams fill the 148 SMs.
 *
 * Within a macroblock the warp uses whichever lane mapping keeps the arithmetic scalar-simple:
 * lane = 4x4 block for transforms (the whole DCT/quant/dequant/IDCT chain of one block lives in one
 * lane's registers), lane = candidate mode for intra 4x4, half-warp = candidate for intra 16x16.
 * Integer pipes only; no tensor cores (this is not a contraction).
 *
 * Normative parts follow ITU-T H.264 8.3 (intra), 8.4.1.3 / 8.4.2.2 (MV prediction, interpolation),
 * 8.5 (transform), 8.7 (deblocking); decision rules follow the oracle (oracle/ref_mb.c), which is pinned
 * by libavcodec's decoder.
 */
#include "cc_xform.cuh"
#include "cc_deblock.cuh"

#define MB_MAX_RPS 8

struct __align__(16) WarpMem {
    uint32_t curY[64];     /* 16 rows x 4 words */
    uint32_t curC[2][16];  /* 8 rows x 2 words */
    uint32_t predY[64];
    uint32_t predC[2][16];
    uint32_t recY[64];
    uint32_t recC[2][16];
    uint8_t tile[17][24];  /* row 0: [3]=top-left, [4..19]=top, [20..23]=top-right; rows 1..16: [3]=left, [4..19]=MB */
    uint8_t edgeC[2][20];  /* [0]=top-left, [1..8]=top, [9..16]=left */
    int16_t coef[27][16];  /* final levels in scan order: 0-15 luma, 16-23 chroma AC, 24 luma DC, 25/26 chroma DC */
    int16_t c4[16][16];    /* intra-4x4 candidate levels */
    int16_t dc[16];
    uint8_t E[16];         /* intra-4x4 edge of the current block: left3..left0, top-left, top0..7, DC */
    uint8_t m4[16];        /* intra-4x4 candidate modes */
    uint8_t m4code[16];    /* syntax nibble per block: 8 = predicted mode, else rem_intra4x4_pred_mode */
    uint8_t n4[16];
};

/* intra 4x4 predictors as 3-tap filters over the edge array E (H.264 8.3.1.2):
 * pred = (E[t0] + 2*E[t1] + E[t2] + 2) >> 2, entry = t0 | t1<<4 | t2<<8.  E index: 0..3 = left[3..0], 4 = top-left,
 * 5..12 = top[0..7], 13 = DC.  Generated from the standard's formulas by tools/gen_i4_table.py. */
#include "cc_i4_table.h"

/* ---- per-row context --------------------------------------------------------------------------------- */
struct MbCtx {
    int lane, mbx, mby, mb, px, py;
    int L, T, TR;      /* neighbour availability inside the slice */
    int qp, qpc, lambda;
    WarpMem *m;
};

/* ---- motion compensation ----------------------------------------------------------------------------- */
/* 16x16 luma predictor at quarter-pel mv into m->predY; lane = (row, half) */
__device__ __forceinline__ void mc_luma16(const cc_kparams &k, const MbCtx &c, int mvx, int mvy) {
    const int e = c_qpel[(mvy & 3) * 4 + (mvx & 3)];
    const int row = c.lane >> 1, half = c.lane & 1;
    const ptrdiff_t base = (ptrdiff_t)(c.py + row + (mvy >> 2)) * k.sy + c.px + half * 8 + (mvx >> 2);
    const int pa = e & 3, pb = (e >> 4) & 3;
    const uint8_t *A = (pa == 0 ? k.ref[0] : (pa == 1 ? k.hb : (pa == 2 ? k.hh : k.hj))) + base + ((e >> 2) & 1) + (ptrdiff_t)((e >> 3) & 1) * k.sy;
    const uint8_t *B = (pb == 0 ? k.ref[0] : (pb == 1 ? k.hb : (pb == 2 ? k.hh : k.hj))) + base + ((e >> 6) & 1) + (ptrdiff_t)((e >> 7) & 1) * k.sy;
    uint32_t a0, a1, b0, b1;
    cc_ld64u(A, a0, a1);
    cc_ld64u(B, b0, b1);
    c.m->predY[row * 4 + half * 2] = __vavgu4(a0, b0);
    c.m->predY[row * 4 + half * 2 + 1] = __vavgu4(a1, b1);
}
/* 8x8 chroma predictors (both planes), H.264 8.4.2.2.2; lane = (plane, row, half) */
__device__ __forceinline__ void mc_chroma8(const cc_kparams &k, const MbCtx &c, int mvx, int mvy) {
    const int pl = c.lane >> 4, row = (c.lane >> 1) & 7, half = c.lane & 1, fx = mvx & 7, fy = mvy & 7;
    const uint8_t *p = k.ref[1 + pl] + (ptrdiff_t)(c.mby * 8 + row + (mvy >> 3)) * k.sc + c.mbx * 8 + half * 4 + (mvx >> 3);
    uint32_t w = 0;
    int a = p[0], b = p[k.sc];
#pragma unroll
    for (int x = 0; x < 4; x++) {
        int a1 = p[x + 1], b1 = p[k.sc + x + 1];
        int v = ((8 - fx) * (8 - fy) * a + fx * (8 - fy) * a1 + (8 - fx) * fy * b + fx * fy * b1 + 32) >> 6;
        w |= (uint32_t)v << (8 * x);
        a = a1; b = b1;
    }
    c.m->predC[pl][row * 2 + half] = w;
}

/* ---- inter luma residual ------------------------------------------------------------------------------ */
/* lanes 0..15 = luma 4x4 blocks (z-order).  Returns cbp (luma 8x8 bits) in all lanes; *nzmask = per-block non-zero.
 * probe: only decide whether anything survives quantisation + decimation (no levels, no reconstruction). */
__device__ __forceinline__ int inter_luma(const cc_kparams &k, const MbCtx &c, bool probe, unsigned *nzmask) {
    WarpMem *m = c.m;
    const int lane = c.lane, b = lane & 15, bx = c_blk_x[b], by = c_blk_y[b];
    uint32_t cw[4], pw[4];
#pragma unroll
    for (int r = 0; r < 4; r++) { cw[r] = m->curY[(by * 4 + r) * 4 + bx]; pw[r] = m->predY[(by * 4 + r) * 4 + bx]; }
    int d[16], z[16];
    dct4x4_res(cw, pw, d);
    int nnz = quant4x4(d, k.qmf[1], k.qbias[1], false);
    to_scan(d, z, false);
    int score = lane < 16 ? decimate_score(z) : 0;
    int s8 = score + __shfl_xor_sync(FULL, score, 1);
    s8 += __shfl_xor_sync(FULL, s8, 2);
    int tot = s8 + __shfl_xor_sync(FULL, s8, 4);
    tot += __shfl_xor_sync(FULL, tot, 8);
    tot = __shfl_sync(FULL, tot, 0);
    /* [EXT] x264 luma decimation: drop an 8x8 whose score < 4, drop everything if the MB score < 6 */
    const bool kill = tot < 6 || s8 < 4;
    if (kill) nnz = 0;
    unsigned bal = __ballot_sync(FULL, lane < 16 && nnz > 0) & 0xffffu;
    int cbp = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) if ((bal >> (4 * i)) & 15u) cbp |= 1 << i;
    *nzmask = bal;
    if (probe) return cbp;
    if (lane < 16) {
        uint32_t ow[4];
        if (nnz) {
            store_levels(m->coef[b], z);
            dequant4x4(d, c.qp);
            idct4x4_add(d, pw, ow);
        } else {
#pragma unroll
            for (int i = 0; i < 16; i++) z[i] = 0;
            store_levels(m->coef[b], z);
#pragma unroll
            for (int r = 0; r < 4; r++) ow[r] = pw[r];
        }
#pragma unroll
        for (int r = 0; r < 4; r++) m->recY[(by * 4 + r) * 4 + bx] = ow[r];
    }
    __syncwarp();
    return cbp;
}

/* ---- chroma residual (both planes) ---------------------------------------------------------------------- */
/* lanes 0..7 = chroma 4x4 blocks (plane = lane>>2).  Returns chroma cbp (0,1,2) in all lanes. */
__device__ __forceinline__ int encode_chroma(const cc_kparams &k, const MbCtx &c, bool intra, bool probe) {
    WarpMem *m = c.m;
    const int lane = c.lane, pl = (lane >> 2) & 1, b = lane & 3, bx = b & 1, by = b >> 1;
    const uint16_t *mf = k.qmf[intra ? 2 : 3], *bias = k.qbias[intra ? 2 : 3];
    uint32_t cw[4], pw[4];
#pragma unroll
    for (int r = 0; r < 4; r++) { cw[r] = m->curC[pl][(by * 4 + r) * 2 + bx]; pw[r] = m->predC[pl][(by * 4 + r) * 2 + bx]; }
    int d[16], z[16];
    dct4x4_res(cw, pw, d);
    const int dc = d[0];
    int nnz = quant4x4(d, mf, bias, true);
    to_scan(d, z, true);
    int score = lane < 8 ? decimate_score(z) : 0;
    score += __shfl_xor_sync(FULL, score, 1);
    score += __shfl_xor_sync(FULL, score, 2);
    int nac = nnz > 0;
    nac |= __shfl_xor_sync(FULL, nac, 1);
    nac |= __shfl_xor_sync(FULL, nac, 2);
    if (!intra && score < 7 && nac) { nnz = 0; nac = 0; } /* [EXT] x264 chroma AC decimation, inter only */
    const bool any_ac = (__ballot_sync(FULL, lane < 8 && nac) != 0);
    /* DC: 2x2 Hadamard over the plane's four blocks, every lane of the group computes all four */
    int q[4];
    {
        int d0 = __shfl_sync(FULL, dc, 0, 4), d1 = __shfl_sync(FULL, dc, 1, 4), d2 = __shfl_sync(FULL, dc, 2, 4), d3 = __shfl_sync(FULL, dc, 3, 4);
        int a = d0 + d1, bb = d0 - d1, cc = d2 + d3, e = d2 - d3;
        q[0] = quant_one(a + cc, mf[0] >> 1, bias[0] << 1); q[1] = quant_one(bb + e, mf[0] >> 1, bias[0] << 1);
        q[2] = quant_one(a - cc, mf[0] >> 1, bias[0] << 1); q[3] = quant_one(bb - e, mf[0] >> 1, bias[0] << 1);
    }
    const int ndc = (q[0] != 0) + (q[1] != 0) + (q[2] != 0) + (q[3] != 0);
    const bool any_dc = (__ballot_sync(FULL, lane < 8 && ndc > 0) != 0);
    const int cbpc = any_ac ? 2 : (any_dc ? 1 : 0);
    if (probe) return cbpc;
    if (lane < 8) {
        if (!nnz) {
#pragma unroll
            for (int i = 0; i < 16; i++) { z[i] = 0; d[i] = 0; }
        }
        store_levels(m->coef[16 + pl * 4 + b], z);
        if (b == 0) {
            int zz[16];
#pragma unroll
            for (int i = 0; i < 16; i++) zz[i] = i < 4 ? q[i] : 0;
            store_levels(m->coef[25 + pl], zz);
        }
        /* H.264 8.5.11.2: dcC = ((f * LevelScale(qpc%6,0,0)) << (qpc/6)) >> 5 */
        const int dmf = (c_dequant_v[c.qpc % 6][0] * 16) << (c.qpc / 6);
        int a = q[0] + q[1], bb = q[0] - q[1], cc = q[2] + q[3], e = q[2] - q[3];
        int f = b == 0 ? a + cc : (b == 1 ? bb + e : (b == 2 ? a - cc : bb - e));
        dequant4x4(d, c.qpc);
        d[0] = (f * dmf) >> 5;
        uint32_t ow[4];
        idct4x4_add(d, pw, ow);
#pragma unroll
        for (int r = 0; r < 4; r++) m->recC[pl][(by * 4 + r) * 2 + bx] = ow[r];
    }
    __syncwarp();
    return cbpc;
}

/* ---- intra 16x16 ----------------------------------------------------------------------------------------- */
struct I16Params { int dc, a, b, c; };
__device__ __forceinline__ void i16_params(const MbCtx &c, I16Params &p) {
    const WarpMem *m = c.m;
    int st = 0, sl = 0, H = 0, V = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) { st += m->tile[0][4 + i]; sl += m->tile[1 + i][3]; }
    p.dc = (c.T && c.L) ? (st + sl + 16) >> 5 : (c.T ? (st + 8) >> 4 : (c.L ? (sl + 8) >> 4 : 128));
#pragma unroll
    for (int i = 0; i < 8; i++) {
        H += (i + 1) * (m->tile[0][4 + 8 + i] - m->tile[0][4 + 6 - i]);  /* tile[0][3] is the top-left sample (x = -1) */
        V += (i + 1) * (m->tile[1 + 8 + i][3] - (i == 7 ? m->tile[0][3] : m->tile[1 + 6 - i][3]));
    }
    p.a = 16 * (m->tile[16][3] + m->tile[0][4 + 15]); p.b = (5 * H + 32) >> 6; p.c = (5 * V + 32) >> 6;
}
/* predictor words of 4x4 block (bx,by) for intra-16x16 mode (0 V, 1 H, 2 DC, 3 plane) */
__device__ __forceinline__ void i16_pred_block(const MbCtx &c, const I16Params &p, int mode, int bx, int by, uint32_t pw[4]) {
    const WarpMem *m = c.m;
#pragma unroll
    for (int r = 0; r < 4; r++) {
        uint32_t w;
        if (mode == 0) w = *(const uint32_t *)&m->tile[0][4 + bx * 4];
        else if (mode == 1) w = m->tile[1 + by * 4 + r][3] * 0x01010101u;
        else if (mode == 2) w = (uint32_t)p.dc * 0x01010101u;
        else {
            w = 0;
#pragma unroll
            for (int x = 0; x < 4; x++) w |= (uint32_t)cc_clip255((p.a + p.b * (bx * 4 + x - 7) + p.c * (by * 4 + r - 7) + 16) >> 5) << (8 * x);
        }
        pw[r] = w;
    }
}
/* returns best cost, *mode_out; half-warp = mode, lane&15 = block (raster) */
__device__ __forceinline__ int analyse_i16(const MbCtx &c, const I16Params &p, int *mode_out) {
    const WarpMem *m = c.m;
    const int blk = c.lane & 15, half = c.lane >> 4, bx = blk & 3, by = blk >> 2;
    uint32_t cw[4];
#pragma unroll
    for (int r = 0; r < 4; r++) cw[r] = m->curY[(by * 4 + r) * 4 + bx];
    unsigned best = 0xffffffffu;
    for (int pass = 0; pass < 2; pass++) {
        const int mode = pass * 2 + half;
        const bool ok = mode == 2 || (mode == 0 && c.T) || (mode == 1 && c.L) || (mode == 3 && c.T && c.L);
        int satd = 0;
        if (ok) {
            uint32_t pw[4];
            i16_pred_block(c, p, mode, bx, by, pw);
            satd = cc_satd4x4_words(cw, pw);
        }
#pragma unroll
        for (int s = 8; s; s >>= 1) satd += __shfl_xor_sync(FULL, satd, s);
        if (ok) {
            const int bits = mode == 0 ? 1 : (mode == 3 ? 5 : 3);
            unsigned key = ((unsigned)(satd + c.lambda * bits) << 2) | (unsigned)mode;
            best = min(best, key);
        }
    }
    best = __reduce_min_sync(FULL, best);
    *mode_out = (int)(best & 3u);
    return (int)(best >> 2);
}
/* code the MB as intra 16x16 with `mode`: levels -> m->coef, reconstruction -> m->recY; returns luma cbp (0 or 15) */
__device__ __forceinline__ int encode_i16(const cc_kparams &k, const MbCtx &c, const I16Params &p, int mode, unsigned *nzmask) {
    WarpMem *m = c.m;
    const int lane = c.lane, b = lane & 15, bx = c_blk_x[b], by = c_blk_y[b];
    uint32_t cw[4], pw[4];
#pragma unroll
    for (int r = 0; r < 4; r++) cw[r] = m->curY[(by * 4 + r) * 4 + bx];
    i16_pred_block(c, p, mode, bx, by, pw);
    int d[16], z[16];
    dct4x4_res(cw, pw, d);
    if (lane < 16) m->dc[by * 4 + bx] = (int16_t)d[0];
    int nnz = quant4x4(d, k.qmf[0], k.qbias[0], true);
    to_scan(d, z, true);
    unsigned bal = __ballot_sync(FULL, lane < 16 && nnz > 0) & 0xffffu;
    __syncwarp();
    int dcv[16];
#pragma unroll
    for (int i = 0; i < 16; i++) dcv[i] = m->dc[i];
    hadamard4x4(dcv, 1);
#pragma unroll
    for (int i = 0; i < 16; i++) dcv[i] = quant_one(dcv[i], k.qmf[0][0] >> 1, k.qbias[0][0] << 1);
    if (lane == 0) {
        int zz[16];
        to_scan(dcv, zz, false);
        store_levels(m->coef[24], zz);
    }
    hadamard4x4(dcv, 0);
    { /* H.264 8.5.10: scaling of the Intra16x16 DC */
        const int qb = c.qp / 6 - 6, ls = c_dequant_v[c.qp % 6][0] * 16;
#pragma unroll
        for (int i = 0; i < 16; i++) dcv[i] = qb >= 0 ? (dcv[i] * ls) << qb : (dcv[i] * ls + (1 << (-qb - 1))) >> -qb;
    }
    if (lane < 16) {
        store_levels(m->coef[b], z);
        dequant4x4(d, c.qp);
        int mydc = 0;
#pragma unroll
        for (int i = 0; i < 16; i++) if (i == by * 4 + bx) mydc = dcv[i];
        d[0] = mydc;
        uint32_t ow[4];
        idct4x4_add(d, pw, ow);
#pragma unroll
        for (int r = 0; r < 4; r++) m->recY[(by * 4 + r) * 4 + bx] = ow[r];
    }
    __syncwarp();
    *nzmask = bal;
    return bal ? 15 : 0;
}

/* ---- intra 4x4 ---------------------------------------------------------------------------------------------- */
/* analyses and codes the 16 blocks in decoding order; reconstruction goes into m->tile, levels into m->c4,
 * modes into m->m4 / m->m4code / m->n4.  Returns the cost (SATD + lambda*bits) in all lanes. */
__device__ __forceinline__ int analyse_i4(const cc_kparams &k, const MbCtx &c) {
    WarpMem *m = c.m;
    const int lane = c.lane;
    const cc_mbdev *nl = k.mbd + c.mb - 1, *nt = k.mbd + c.mb - k.mb_w;
    int cost = 24 * c.lambda;
    for (int blk = 0; blk < 16; blk++) {
        const int bx = c_blk_x[blk], by = c_blk_y[blk];
        const int has_top = by > 0 || c.T, has_left = bx > 0 || c.L;
        int has_tr;
        if (by == 0) has_tr = bx < 3 ? c.T : c.TR;
        else has_tr = !(bx == 3 || blk == 3 || blk == 11 || blk == 7 || blk == 13 || blk == 15);
        /* edge array */
        if (lane < 13) {
            int v;
            if (lane < 4) v = m->tile[1 + by * 4 + (3 - lane)][3 + bx * 4];
            else if (lane == 4) v = m->tile[by * 4][3 + bx * 4];
            else { int i = lane - 5; v = (i < 4 || has_tr) ? m->tile[by * 4][4 + bx * 4 + i] : m->tile[by * 4][4 + bx * 4 + 3]; }
            m->E[lane] = (uint8_t)v;
        }
        __syncwarp();
        {
            int st = m->E[5] + m->E[6] + m->E[7] + m->E[8], sl = m->E[0] + m->E[1] + m->E[2] + m->E[3];
            int dcv = (has_top && has_left) ? (st + sl + 4) >> 3 : (has_top ? (st + 2) >> 2 : (has_left ? (sl + 2) >> 2 : 128));
            if (lane == 13) m->E[13] = (uint8_t)dcv;
        }
        /* predicted mode, H.264 8.3.1.1 */
        int mA, mB;
        if (bx > 0) mA = m->m4[c_blk_of[by * 4 + bx - 1]];
        else if (c.L) mA = nl->type == CC_MB_I4 ? nl->i4_mode[c_blk_of[by * 4 + 3]] : 2;
        else mA = -1;
        if (by > 0) mB = m->m4[c_blk_of[(by - 1) * 4 + bx]];
        else if (c.T) mB = nt->type == CC_MB_I4 ? nt->i4_mode[c_blk_of[12 + bx]] : 2;
        else mB = -1;
        const int pm = (mA < 0 || mB < 0) ? 2 : min(mA, mB);
        __syncwarp();
        uint32_t cw[4], pw[4] = {0, 0, 0, 0};
#pragma unroll
        for (int r = 0; r < 4; r++) cw[r] = m->curY[(by * 4 + r) * 4 + bx];
        unsigned key = 0xffffffffu;
        {
            const int md = lane < 9 ? lane : 2;
            const bool ok = lane < 9 && (md == 2 || ((md == 0 || md == 3 || md == 7) && has_top) || ((md == 1 || md == 8) && has_left) ||
                                         ((md == 4 || md == 5 || md == 6) && has_top && has_left));
            if (ok) {
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    uint32_t w = 0;
#pragma unroll
                    for (int x = 0; x < 4; x++) {
                        unsigned t = c_i4tab[md][r * 4 + x];
                        int v = (m->E[t & 15] + 2 * m->E[(t >> 4) & 15] + m->E[(t >> 8) & 15] + 2) >> 2;
                        w |= (uint32_t)v << (8 * x);
                    }
                    pw[r] = w;
                }
                int satd = cc_satd4x4_words(cw, pw);
                key = ((unsigned)(satd + c.lambda * (md == pm ? 1 : 4)) << 4) | (unsigned)md;
            }
        }
        const unsigned bk = __reduce_min_sync(FULL, key);
        const int bm = (int)(bk & 15u);
        cost += (int)(bk >> 4);
#pragma unroll
        for (int r = 0; r < 4; r++) pw[r] = __shfl_sync(FULL, pw[r], bm);
        /* every lane runs the block's transform chain on identical data; lane 0 publishes */
        int d[16], z[16];
        dct4x4_res(cw, pw, d);
        const int nnz = quant4x4(d, k.qmf[0], k.qbias[0], false);
        to_scan(d, z, false);
        uint32_t ow[4];
        if (nnz) { dequant4x4(d, c.qp); idct4x4_add(d, pw, ow); }
        else {
#pragma unroll
            for (int r = 0; r < 4; r++) ow[r] = pw[r];
        }
        if (lane == 0) {
            store_levels(m->c4[blk], z);
#pragma unroll
            for (int r = 0; r < 4; r++) *(uint32_t *)&m->tile[1 + by * 4 + r][4 + bx * 4] = ow[r];
            m->m4[blk] = (uint8_t)bm;
            m->m4code[blk] = (uint8_t)(bm == pm ? 8 : (bm < pm ? bm : bm - 1));
            m->n4[blk] = (uint8_t)nnz;
        }
        __syncwarp();
    }
    return cost;
}

/* ---- MV prediction, H.264 8.4.1.3 (one 16x16 partition per MB, one reference picture) ------------------------- */
__device__ __forceinline__ void nb_mv(const cc_mbdev *n, int avail, int &ref, int mv[2]) {
    mv[0] = mv[1] = 0;
    if (!avail) { ref = -2; return; }
    if (n->type >= CC_MB_I16) { ref = -1; return; }
    ref = 0; mv[0] = n->mv[0]; mv[1] = n->mv[1];
}
__device__ __forceinline__ void predict_mv(const cc_kparams &k, const MbCtx &c, int mvp[2], int skip[2]) {
    const cc_mbdev *me = k.mbd + c.mb;
    int rA, rB, rC, a[2], b[2], cc[2];
    nb_mv(me - 1, c.L, rA, a);
    nb_mv(me - k.mb_w, c.T, rB, b);
    nb_mv(me - k.mb_w + 1, c.TR, rC, cc);
    if (rC == -2) nb_mv(me - k.mb_w - 1, c.T && c.L, rC, cc);
    const int rA0 = rA, rB0 = rB, a0x = a[0], a0y = a[1], b0x = b[0], b0y = b[1];
    if (rB == -2 && rC == -2 && rA != -2) { rB = rC = rA; b[0] = cc[0] = a[0]; b[1] = cc[1] = a[1]; }
    const int n = (rA == 0) + (rB == 0) + (rC == 0);
    if (n == 1) {
        if (rA == 0) { mvp[0] = a[0]; mvp[1] = a[1]; }
        else if (rB == 0) { mvp[0] = b[0]; mvp[1] = b[1]; }
        else { mvp[0] = cc[0]; mvp[1] = cc[1]; }
    } else { mvp[0] = cc_median3(a[0], b[0], cc[0]); mvp[1] = cc_median3(a[1], b[1], cc[1]); }
    /* P_Skip vector, 8.4.1.1 */
    if (!c.L || !c.T || (rA0 == 0 && a0x == 0 && a0y == 0) || (rB0 == 0 && b0x == 0 && b0y == 0)) skip[0] = skip[1] = 0;
    else { skip[0] = mvp[0]; skip[1] = mvp[1]; }
}

/* ---- intra chroma ------------------------------------------------------------------------------------------------ */
/* chroma predictor value for mode (0 DC, 1 H, 2 V, 3 plane) at (x,y) of plane pl from m->edgeC */
struct C8Params { int dcv[4]; int a, b, c; };
__device__ __forceinline__ void c8_params(const MbCtx &c, int pl, C8Params &p) {
    const uint8_t *e = c.m->edgeC[pl];
    int st0 = 0, st1 = 0, sl0 = 0, sl1 = 0, H = 0, V = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) { st0 += e[1 + i]; st1 += e[5 + i]; sl0 += e[9 + i]; sl1 += e[13 + i]; }
    const int T = c.T, L = c.L;
    p.dcv[0] = (T && L) ? (st0 + sl0 + 4) >> 3 : (T ? (st0 + 2) >> 2 : (L ? (sl0 + 2) >> 2 : 128));
    p.dcv[1] = T ? (st1 + 2) >> 2 : (L ? (sl0 + 2) >> 2 : 128);
    p.dcv[2] = L ? (sl1 + 2) >> 2 : (T ? (st0 + 2) >> 2 : 128);
    p.dcv[3] = (T && L) ? (st1 + sl1 + 4) >> 3 : (T ? (st1 + 2) >> 2 : (L ? (sl1 + 2) >> 2 : 128));
#pragma unroll
    for (int i = 0; i < 4; i++) {
        H += (i + 1) * (e[1 + 4 + i] - (i == 3 ? e[0] : e[1 + 2 - i]));
        V += (i + 1) * (e[9 + 4 + i] - (i == 3 ? e[0] : e[9 + 2 - i]));
    }
    p.a = 16 * (e[16] + e[8]); p.b = (34 * H + 32) >> 6; p.c = (34 * V + 32) >> 6;
}
__device__ __forceinline__ void c8_pred_block(const MbCtx &c, const C8Params &p, int pl, int mode, int bx, int by, uint32_t pw[4]) {
    const uint8_t *e = c.m->edgeC[pl];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        uint32_t w = 0;
        if (mode == 0) w = (uint32_t)p.dcv[by * 2 + bx] * 0x01010101u;
        else if (mode == 1) w = e[9 + by * 4 + r] * 0x01010101u;
        else if (mode == 2) w = (uint32_t)e[1 + bx * 4] | ((uint32_t)e[2 + bx * 4] << 8) | ((uint32_t)e[3 + bx * 4] << 16) | ((uint32_t)e[4 + bx * 4] << 24);
        else {
#pragma unroll
            for (int x = 0; x < 4; x++) w |= (uint32_t)cc_clip255((p.a + p.b * (bx * 4 + x - 3) + p.c * (by * 4 + r - 3) + 16) >> 5) << (8 * x);
        }
        pw[r] = w;
    }
}
/* picks the chroma mode by SATD over both planes and leaves the predictor in m->predC; returns the mode */
__device__ __forceinline__ int intra_chroma_pred(const MbCtx &c) {
    WarpMem *m = c.m;
    const int lane = c.lane, pl = (lane >> 2) & 1, b = lane & 3, bx = b & 1, by = b >> 1;
    C8Params p;
    c8_params(c, pl, p);
    uint32_t cw[4];
#pragma unroll
    for (int r = 0; r < 4; r++) cw[r] = m->curC[pl][(by * 4 + r) * 2 + bx];
    unsigned best = 0xffffffffu;
    for (int mode = 0; mode < 4; mode++) {
        const bool ok = mode == 0 || (mode == 1 && c.L) || (mode == 2 && c.T) || (mode == 3 && c.T && c.L);
        int satd = 0;
        if (ok && lane < 8) {
            uint32_t pw[4];
            c8_pred_block(c, p, pl, mode, bx, by, pw);
            satd = cc_satd4x4_words(cw, pw);
        }
        satd += __shfl_xor_sync(FULL, satd, 1); satd += __shfl_xor_sync(FULL, satd, 2); satd += __shfl_xor_sync(FULL, satd, 4);
        if (ok) {
            const int bits = mode == 0 ? 1 : (mode == 3 ? 5 : 3);
            best = min(best, ((unsigned)(satd + c.lambda * bits) << 2) | (unsigned)mode);
        }
    }
    best = __shfl_sync(FULL, best, 0);
    const int bm = (int)(best & 3u);
    if (lane < 8) {
        uint32_t pw[4];
        c8_pred_block(c, p, pl, bm, bx, by, pw);
#pragma unroll
        for (int r = 0; r < 4; r++) m->predC[pl][(by * 4 + r) * 2 + bx] = pw[r];
    }
    __syncwarp();
    return bm;
}

/* ---- symbol stream -------------------------------------------------------------------------------------------------- */
/* appends block `blk` of m->coef (n scan positions) to the row payload as {u16 mask, levels...}; returns new offset (in int16 units) */
__device__ __forceinline__ unsigned emit_block(const MbCtx &c, int16_t *out, unsigned off, int blk, int n) {
    const int v = c.lane < n ? c.m->coef[blk][c.lane] : 0;
    const unsigned mask = __ballot_sync(FULL, v != 0);
    if (c.lane == 0) out[off] = (int16_t)mask;
    if (v != 0) out[off + 1 + __popc(mask & ((1u << c.lane) - 1u))] = (int16_t)v;
    return off + 1 + (unsigned)__popc(mask);
}

/* ---- the kernel --------------------------------------------------------------------------------------------------------- */
__device__ __forceinline__ void wait_progress(volatile int *flag, int need, int lane) {
    if (lane == 0) while (*flag < need) CC_SPIN_PAUSE();
    __syncwarp();
    __threadfence_block();
}

/* Unfiltered neighbour samples.  Intra prediction uses the reconstruction BEFORE deblocking (H.264 8.3), while the
 * deblocking warps of the same slice modify k.rec a few macroblocks behind the encode warps.  So the encode warps
 * never read k.rec: the bottom sample row of every macroblock row goes to a shared-memory line (`TopLines`), the
 * right column of the previous macroblock stays in the warp's own tile. */
struct TopLines {
    uint8_t *y, *c; int sy, sc; /* y: rps lines of W+32 bytes (origin +16); c: 2*rps lines of W/2+16 bytes (origin +8) */
    __device__ __forceinline__ uint8_t *Y(int w) const { return y + w * sy + 16; }
    __device__ __forceinline__ uint8_t *C(int pl, int w, int rps) const { return c + (pl * rps + w) * sc + 8; }
};

__device__ void encode_row(const cc_kparams &k, WarpMem *m, const TopLines &tl, volatile int *progress, int w, int row, int row0, int lane) {
    const int mw = k.mb_w, qp = k.qp;
    MbCtx c;
    c.lane = lane; c.mby = row; c.py = row * 16; c.qp = qp; c.qpc = c_chroma_qp[qp]; c.lambda = c_lambda[qp]; c.m = m;
    c.T = row > row0;
    int16_t *out = (int16_t *)(k.rowbuf + (size_t)row * mw * CC_ROW_PAYLOAD_PER_MB);
    unsigned off = 0, n_skip = 0, n_inter = 0, n_intra = 0;
    for (int mbx = 0; mbx < mw; mbx++) {
        c.mbx = mbx; c.px = mbx * 16; c.mb = row * mw + mbx; c.L = mbx > 0; c.TR = c.T && mbx + 1 < mw;
        if (c.T) wait_progress(progress + w - 1, min(mbx + 2, mw), lane);
        /* ---- stage the current MB and its reconstructed neighbourhood ---- */
        {
            const uint8_t *cy = k.cur[0] + (size_t)c.py * k.sy + c.px;
            m->curY[lane] = *(const uint32_t *)(cy + (size_t)(lane >> 2) * k.sy + 4 * (lane & 3));
            m->curY[lane + 32] = *(const uint32_t *)(cy + (size_t)(8 + (lane >> 2)) * k.sy + 4 * (lane & 3));
            const int pl = lane >> 4, r = (lane >> 1) & 7, h = lane & 1;
            m->curC[pl][r * 2 + h] = *(const uint32_t *)(k.cur[1 + pl] + (size_t)(row * 8 + r) * k.sc + mbx * 8 + 4 * h);
            if (lane < 21) { /* top-left, top, top-right: unfiltered line of the row above */
                int x = lane - 1;
                bool ok = c.T && (x < 0 ? c.L : (x < 16 ? true : c.TR));
                m->tile[0][3 + lane] = ok ? tl.Y(w - 1)[c.px + x] : 0;
            }
            if (lane >= 16 && !c.L) m->tile[1 + lane - 16][3] = 0; /* left column: kept from the previous macroblock */
            if (lane < 18) {
                const int p2 = lane >= 9, i = lane - 9 * p2; /* i = 0: top-left, 1..8: top */
                const uint8_t *tc = tl.C(p2, w - 1, k.rps) + mbx * 8;
                m->edgeC[p2][i] = (i == 0 ? (c.T && c.L) : c.T) ? tc[i - 1] : 0;
            } else if (lane < 26 && !c.L) { m->edgeC[0][9 + lane - 18] = 0; m->edgeC[1][9 + lane - 18] = 0; }
        }
        __syncwarp();

        int type, mode16 = 2, chroma_mode = 0, cbp = 0, mvx = 0, mvy = 0, mvdx = 0, mvdy = 0;
        unsigned nzmask = 0;
        int mvp[2] = {0, 0}, skipmv[2] = {0, 0};
        int cost_inter = 1 << 30, cost_i4 = 1 << 30;
        bool done = false, from_tile = false;
        const cc_me me = k.me[c.mb];

        if (!k.is_idr) {
            predict_mv(k, c, mvp, skipmv);
            /* [EXT] x264 probe_skip: code the MB at the P_Skip vector; nothing left after decimation -> P_Skip */
            mc_luma16(k, c, skipmv[0], skipmv[1]);
            __syncwarp();
            unsigned nz;
            if (inter_luma(k, c, true, &nz) == 0) {
                mc_chroma8(k, c, skipmv[0], skipmv[1]);
                __syncwarp();
                if (encode_chroma(k, c, false, true) == 0) {
                    type = CC_MB_PSKIP; mvx = skipmv[0]; mvy = skipmv[1];
                    m->recY[lane] = m->predY[lane]; m->recY[lane + 32] = m->predY[lane + 32];
                    m->recC[lane >> 4][lane & 15] = m->predC[lane >> 4][lane & 15];
                    __syncwarp();
                    done = true;
                }
            }
            if (!done) cost_inter = me.satd + c.lambda * (cc_se_bits(me.mv[0] - mvp[0]) + cc_se_bits(me.mv[1] - mvp[1]));
        }
        if (!done) {
            I16Params ip;
            i16_params(c, ip);
            int cost_i16 = analyse_i16(c, ip, &mode16);
            if (k.is_idr || cost_i16 * 4 <= cost_inter * 5) cost_i4 = analyse_i4(k, c);
            if (!k.is_idr) { cost_i16 += 8 * c.lambda; if (cost_i4 < (1 << 30)) cost_i4 += 8 * c.lambda; }
            if (!k.is_idr && cost_inter <= cost_i16 && cost_inter <= cost_i4) type = CC_MB_P16;
            else type = cost_i16 <= cost_i4 ? CC_MB_I16 : CC_MB_I4;

            if (type == CC_MB_I4) {
                from_tile = true;
                if (lane < 16) {
                    const uint32_t *s = (const uint32_t *)m->c4[lane];
                    uint32_t *d = (uint32_t *)m->coef[lane];
#pragma unroll
                    for (int i = 0; i < 8; i++) d[i] = s[i];
                }
                unsigned bal = __ballot_sync(FULL, lane < 16 && m->n4[lane & 15] > 0) & 0xffffu;
                nzmask = bal;
#pragma unroll
                for (int i = 0; i < 4; i++) if ((bal >> (4 * i)) & 15u) cbp |= 1 << i;
            } else if (type == CC_MB_I16) {
                cbp = encode_i16(k, c, ip, mode16, &nzmask);
            } else {
                mvx = me.mv[0]; mvy = me.mv[1];
                mc_luma16(k, c, mvx, mvy);
                __syncwarp();
                cbp = inter_luma(k, c, false, &nzmask);
            }
            if (type >= CC_MB_I16) {
                chroma_mode = intra_chroma_pred(c);
                cbp |= encode_chroma(k, c, true, false) << 4;
            } else {
                mc_chroma8(k, c, mvx, mvy);
                __syncwarp();
                cbp |= encode_chroma(k, c, false, false) << 4;
                mvdx = mvx - mvp[0]; mvdy = mvy - mvp[1];
                if (cbp == 0 && mvx == skipmv[0] && mvy == skipmv[1]) type = CC_MB_PSKIP;
            }
        }
        /* ---- publish: reconstruction, records, symbols ---- */
        {
            uint8_t *ry = k.rec[0] + (size_t)c.py * k.sy + c.px;
            const int r = lane >> 1, h = lane & 1;
            uint32_t w0, w1;
            if (from_tile) { w0 = *(const uint32_t *)&m->tile[1 + r][4 + h * 8]; w1 = *(const uint32_t *)&m->tile[1 + r][8 + h * 8]; }
            else { w0 = m->recY[r * 4 + h * 2]; w1 = m->recY[r * 4 + h * 2 + 1]; }
            *(uint32_t *)(ry + (size_t)r * k.sy + h * 8) = w0;
            *(uint32_t *)(ry + (size_t)r * k.sy + h * 8 + 4) = w1;
            const int pl = lane >> 4, cr = (lane >> 1) & 7;
            *(uint32_t *)(k.rec[1 + pl] + (size_t)(row * 8 + cr) * k.sc + mbx * 8 + 4 * h) = m->recC[pl][cr * 2 + h];
            /* unfiltered neighbours of the next macroblock (right column) and of the row below (bottom line) */
            if (lane < 16) m->tile[1 + lane][3] = from_tile ? m->tile[1 + lane][19] : (uint8_t)(m->recY[lane * 4 + 3] >> 24);
            else m->edgeC[(lane - 16) >> 3][9 + (lane & 7)] = (uint8_t)(m->recC[(lane - 16) >> 3][(lane & 7) * 2 + 1] >> 24);
            if (lane < 4) *(uint32_t *)(tl.Y(w) + c.px + 4 * lane) = from_tile ? *(const uint32_t *)&m->tile[16][4 + 4 * lane] : m->recY[60 + lane];
            else if (lane < 8) *(uint32_t *)(tl.C((lane - 4) >> 1, w, k.rps) + mbx * 8 + 4 * (lane & 1)) = m->recC[(lane - 4) >> 1][14 + (lane & 1)];
        }
        if (lane == 0) {
            cc_mbdev d;
            d.type = (uint8_t)type; d.pad0 = 0; d.nz_luma = (uint16_t)(type == CC_MB_PSKIP ? 0 : nzmask);
            d.mv[0] = (int16_t)mvx; d.mv[1] = (int16_t)mvy;
#pragma unroll
            for (int i = 0; i < 16; i++) d.i4_mode[i] = type == CC_MB_I4 ? m->m4[i] : 2;
            k.mbd[c.mb] = d;
            cc_mbhost hh;
            hh.type = (uint8_t)type; hh.modes = (uint8_t)(mode16 | (chroma_mode << 2)); hh.cbp = (uint8_t)cbp; hh.pad = 0;
            hh.mvd[0] = (int16_t)mvdx; hh.mvd[1] = (int16_t)mvdy;
            k.mbh[c.mb] = hh;
        }
        if (type == CC_MB_PSKIP) n_skip++;
        else {
            if (type == CC_MB_P16) n_inter++; else n_intra++;
            if (type == CC_MB_I4) {
                if (lane < 4) {
                    unsigned v = 0;
#pragma unroll
                    for (int i = 0; i < 4; i++) v |= (unsigned)m->m4code[lane * 4 + i] << (4 * i);
                    out[off + lane] = (int16_t)v;
                }
                off += 4;
            }
            const int cl = cbp & 15, cc = cbp >> 4;
            if (type == CC_MB_I16 || cbp) {
                if (type == CC_MB_I16) off = emit_block(c, out, off, 24, 16);
                for (int blk = 0; blk < 16; blk++)
                    if (cl & (1 << (blk >> 2))) off = emit_block(c, out, off, blk, type == CC_MB_I16 ? 15 : 16);
                if (cc) {
                    off = emit_block(c, out, off, 25, 4);
                    off = emit_block(c, out, off, 26, 4);
                    if (cc & 2) for (int blk = 16; blk < 24; blk++) off = emit_block(c, out, off, blk, 15);
                }
            }
        }
        __syncwarp();
        __threadfence_block();
        if (lane == 0) progress[w] = mbx + 1;
    }
    if (lane == 0) {
        cc_rowhdr h;
        h.bytes = off * 2; h.n_skip = n_skip; h.n_inter = n_inter; h.n_intra = n_intra;
        k.rowhdr[row] = h;
    }
}

/* CTA = one slice.  Warps [0, rps) encode one macroblock row each; warps [rps, 2*rps) run the in-loop filter over the
 * same rows, each a few macroblocks behind its encode warp and 2 macroblocks behind the filter warp of the row above. */
__global__ void __launch_bounds__(2 * MB_MAX_RPS * 32) k_mb_encode(cc_kparams k) {
    __shared__ WarpMem wm[MB_MAX_RPS];
    __shared__ DbMem dbm[MB_MAX_RPS];
    __shared__ volatile int progress[MB_MAX_RPS], progress2[MB_MAX_RPS];
    CC_DYN_SMEM(dsm);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const bool filt = wid >= k.rps;
    const int w = filt ? wid - k.rps : wid;
    const int row0 = blockIdx.x * k.rps, row = row0 + w;
    const bool valid = row < k.mb_h;
    TopLines tl;
    tl.sy = k.W + 32; tl.sc = (k.W >> 1) + 16;
    tl.y = dsm; tl.c = dsm + (size_t)k.rps * tl.sy;
    if (lane == 0 && !filt) { progress[w] = 0; progress2[w] = 0; }
    __syncthreads();
    if (!valid) return;
    if (!filt) {
        encode_row(k, &wm[w], tl, progress, w, row, row0, lane);
    } else {
        for (int mbx = 0; mbx < k.mb_w; mbx++) {
            wait_progress(progress + w, mbx + 1, lane);
            if (w > 0) wait_progress(progress2 + w - 1, min(mbx + 2, k.mb_w), lane);
            deblock_mb(k, &dbm[w], lane, mbx, row, mbx > 0, w > 0);
            __threadfence_block();
            if (lane == 0) progress2[w] = mbx + 1;
        }
    }
}

/* ---- pack: gather row headers, MB records and row payloads into the mapped host buffer --------------------------------- */
__global__ void k_pack(cc_kparams k) {
    const int nmb = k.mb_w * k.mb_h;
    const unsigned hdr_bytes = (unsigned)((sizeof(cc_rowhdr) * k.mb_h + 15) / 16 * 16);
    const unsigned mbh_bytes = (unsigned)((sizeof(cc_mbhost) * nmb + 15) / 16 * 16);
    const int tid = threadIdx.x, nt = blockDim.x;
    if ((int)blockIdx.x == k.mb_h) { /* extra CTA: headers + MB records */
        uint32_t *dst = (uint32_t *)k.out_host_dev;
        const uint32_t *src = (const uint32_t *)k.rowhdr;
        for (int i = tid; i < (int)(sizeof(cc_rowhdr) * k.mb_h / 4); i += nt) dst[i] = src[i];
        uint2 *d2 = (uint2 *)(k.out_host_dev + hdr_bytes);
        const uint2 *s2 = (const uint2 *)k.mbh;
        for (int i = tid; i < nmb; i += nt) d2[i] = s2[i];
        return;
    }
    const int row = blockIdx.x;
    unsigned offs = 0;
    for (int r = 0; r < row; r++) offs += (k.rowhdr[r].bytes + 15u) & ~15u;
    const unsigned n16 = (k.rowhdr[row].bytes + 15u) >> 4;
    if (hdr_bytes + mbh_bytes + offs + n16 * 16 > k.out_cap) return;
    uint4 *dst = (uint4 *)(k.out_host_dev + hdr_bytes + mbh_bytes + offs);
    const uint4 *src = (const uint4 *)(k.rowbuf + (size_t)row * k.mb_w * CC_ROW_PAYLOAD_PER_MB);
    for (unsigned i = tid; i < n16; i += nt) dst[i] = src[i];
}

extern "C" int cck_encode(const cc_kparams *k, cudaStream_t st) {
    const int ns = (k->mb_h + k->rps - 1) / k->rps;
    const size_t smem = (size_t)k->rps * (2 * (size_t)k->W + 64);
#ifndef CC_EMU
    static size_t smem_set = 0; /* opt in above the 48 KB default once (wide pictures with many rows per slice) */
    if (smem > smem_set) {
        if (cudaFuncSetAttribute(k_mb_encode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
        smem_set = smem;
    }
#endif
    CC_LAUNCH_SMEM(k_mb_encode, dim3((unsigned)ns), dim3((unsigned)k->rps * 64), smem, st, *k);
    return 1;
}
extern "C" int cck_pack(const cc_kparams *k, cudaStream_t st) {
    CC_LAUNCH(k_pack, dim3((unsigned)k->mb_h + 1), dim3(128), st, *k);
    return 1;
}
