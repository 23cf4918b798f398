/*
 * cc_k_mb.cu — macroblock analysis, mode decision, transform/quant/reconstruction and in-loop deblocking
 * (SURVEY.md §8a rows K4, K5, K6, K7, K8).
 *
 * Execution model.  A slice is a run of `rps` macroblock rows and shares nothing with other slices
 * (intra/MV prediction stop at slice edges; deblocking uses disable_deblocking_filter_idc=2).  One CTA = one BAND of up to 8
 * rows, one warp = one macroblock row; a slice of up to 8 rows is one band, a taller slice (rps a multiple of 8, or the whole
 * picture) is several bands whose wavefronts chain through frame-tagged progress words and the per-macroblock records in global
 * memory (the band below spins on the last row of the band above; CTAs of a launch are dispatched in index order, so the band it
 * waits for is already running).  Rows of a slice advance as a wavefront with a
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
#include <stdlib.h>
#include "cc_xform8.cuh"
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
    uint8_t S[2][48];      /* intra-4x4 source arrays of the two blocks in flight (raw edge, DC, 3-tap and 2-tap filtered edge; tools/gen_i4_table.py) */
    uint8_t nbm[8];        /* Intra4x4PredMode of the left MB's column-3 blocks [0..3] and the top MB's row-3 blocks [4..7]; 255 = unavailable */
    uint8_t m4[16];        /* intra-4x4 candidate modes */
    uint8_t m4code[16];    /* syntax nibble per block: 8 = predicted mode, else rem_intra4x4_pred_mode */
    uint8_t n4[16];
    uint4 pf[4][4];        /* ring of prefetched per-macroblock records (cc_me, cc_pre, cc_me8 = 16 + 16 + 32 bytes), filled by cp.async */
    uint4 pc[4][24];       /* ring of prefetched source pixels (k_mb_encode): 16 luma rows x 16 bytes, then 8 rows x 8 bytes of Cb and of Cr */
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
__device__ __forceinline__ void mc_luma16(const cc_kparams &k, const MbCtx &c, int mvx, int mvy, uint32_t *dstY = nullptr) {
    if (!dstY) dstY = c.m->predY;
    const int e = c_qpel[(mvy & 3) * 4 + (mvx & 3)];
    const int row = c.lane >> 1, half = c.lane & 1;
    const ptrdiff_t base = (ptrdiff_t)(c.py + row + (mvy >> 2)) * k.sy + c.px + half * 8 + (mvx >> 2);
    const int pa = e & 3, pb = (e >> 4) & 3;
    const uint8_t *A = (pa == 0 ? k.ref[0] : (pa == 1 ? k.hb : (pa == 2 ? k.hh : k.hj))) + base + ((e >> 2) & 1) + (ptrdiff_t)((e >> 3) & 1) * k.sy;
    const uint8_t *B = (pb == 0 ? k.ref[0] : (pb == 1 ? k.hb : (pb == 2 ? k.hh : k.hj))) + base + ((e >> 6) & 1) + (ptrdiff_t)((e >> 7) & 1) * k.sy;
    uint32_t a0, a1, b0, b1;
    cc_ld64u(A, a0, a1);
    cc_ld64u(B, b0, b1);
    dstY[row * 4 + half * 2] = __vavgu4(a0, b0);
    dstY[row * 4 + half * 2 + 1] = __vavgu4(a1, b1);
}
/* P_8x8: the lane mappings of mc_luma16 (row, 8-pixel half) and mc_chroma8 (plane, row, 4-pixel half) never straddle an 8x8
 * quadrant (4x4 in chroma), so a P_8x8 macroblock is predicted by the same two functions with each lane passing the vector
 * of the quadrant it lies in. */
__device__ __forceinline__ int luma_lane_quad(int lane) { return ((lane >> 4) << 1) | (lane & 1); }
__device__ __forceinline__ int chroma_lane_quad(int lane) { return (((lane >> 3) & 1) << 1) | (lane & 1); }
__device__ __forceinline__ void mc_mb8(const cc_kparams &k, const MbCtx &c, const cc_me8 &m8);
/* 8x8 chroma predictors (both planes), H.264 8.4.2.2.2; lane = (plane, row, half): four pixels from the 5 x 2 samples around them.
 * The samples come as two aligned words per row (five bytes from any alignment fit in eight), and the bilinear sum
 * ((8-fx)(8-fy) A + fx (8-fy) B + (8-fx) fy C + fx fy D + 32) >> 6 runs on two pixels at a time in the 16-bit halves of a register
 * (the weights add up to 64, so a half never exceeds 64 * 255 + 32): the same integers as the per-pixel formula. */
__device__ __forceinline__ void mc_chroma8(const cc_kparams &k, const MbCtx &c, int mvx, int mvy, uint32_t (*dstC)[16] = nullptr) {
    if (!dstC) dstC = c.m->predC;
    const int pl = c.lane >> 4, row = (c.lane >> 1) & 7, half = c.lane & 1, fx = mvx & 7, fy = mvy & 7;
    const uint8_t *p = k.ref[1 + pl] + (ptrdiff_t)(c.mby * 8 + row + (mvy >> 3)) * k.sc + c.mbx * 8 + half * 4 + (mvx >> 3);
    const uintptr_t a = (uintptr_t)p;
    const unsigned sh = (unsigned)(a & 3) * 8u;
    const uint32_t *q0 = (const uint32_t *)(a & ~(uintptr_t)3), *q1 = (const uint32_t *)((const uint8_t *)q0 + k.sc); /* sc is a multiple of 4 */
    const uint32_t a0 = q0[0], a1 = q0[1], b0 = q1[0], b1 = q1[1];
    const uint32_t aL = __funnelshift_r(a0, a1, sh), aR = __funnelshift_rc(a0, a1, sh + 8u);
    const uint32_t bL = __funnelshift_r(b0, b1, sh), bR = __funnelshift_rc(b0, b1, sh + 8u);
    const uint32_t w00 = (uint32_t)((8 - fx) * (8 - fy)), w01 = (uint32_t)(fx * (8 - fy)), w10 = (uint32_t)((8 - fx) * fy), w11 = (uint32_t)(fx * fy);
    const uint32_t M = 0x00ff00ffu;
    const uint32_t ve = (w00 * (aL & M) + w01 * (aR & M) + w10 * (bL & M) + w11 * (bR & M) + 0x00200020u) >> 6;
    const uint32_t vo = (w00 * ((aL >> 8) & M) + w01 * ((aR >> 8) & M) + w10 * ((bL >> 8) & M) + w11 * ((bR >> 8) & M) + 0x00200020u) >> 6;
    dstC[pl][row * 2 + half] = (ve & M) | ((vo & M) << 8);
}

__device__ __forceinline__ void mc_mb8(const cc_kparams &k, const MbCtx &c, const cc_me8 &m8) {
    const int ql = luma_lane_quad(c.lane), qc = chroma_lane_quad(c.lane);
    int lx = m8.mv[0][0], ly = m8.mv[0][1], cx = lx, cy = ly;
#pragma unroll
    for (int q = 1; q < 4; q++) {
        if (ql == q) { lx = m8.mv[q][0]; ly = m8.mv[q][1]; }
        if (qc == q) { cx = m8.mv[q][0]; cy = m8.mv[q][1]; }
    }
    mc_luma16(k, c, lx, ly);
    mc_chroma8(k, c, cx, cy);
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
/* the sums of 8.3.3 over the 16 + 16 + 1 neighbour samples, one sample (or two) per lane and two warp reductions: lane l holds the
 * top sample x = l - 1 (l <= 16; x = -1 is the corner) and the left sample y = l (l < 16).  H and V are the standard's sums
 * sum (i + 1) (p[8 + i] - p[6 - i]) rewritten as sum (x - 7) p[x] over x = -1 .. 15. */
__device__ __forceinline__ void i16_params(const MbCtx &c, I16Params &p) {
    const WarpMem *m = c.m;
    const int l = c.lane;
    const int t = l < 17 ? m->tile[0][3 + l] : 0, lf = l < 16 ? m->tile[1 + l][3] : 0;
    const int sums = __reduce_add_sync(FULL, ((l >= 1 && l < 17) ? t : 0) | (lf << 16)); /* top | left << 16, each at most 4080 */
    const int st = sums & 0xffff, sl = sums >> 16;
    /* H + 65536 V: |H|, |V| <= 72 * 255 */
    const int hv = __reduce_add_sync(FULL, (l - 8) * t + (((l < 16 ? (l - 7) * lf : 0) + (l == 0 ? -8 * t : 0)) << 16));
    const int H = (int)(int16_t)(hv & 0xffff), V = (hv - H) >> 16;
    p.dc = (c.T && c.L) ? (st + sl + 16) >> 5 : (c.T ? (st + 8) >> 4 : (c.L ? (sl + 8) >> 4 : 128));
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
    int tc[16];
    cc_satd4x4_src(cw, tc);
    unsigned best = 0xffffffffu;
    for (int pass = 0; pass < 2; pass++) {
        const int mode = pass * 2 + half;
        const bool ok = mode == 2 || (mode == 0 && c.T) || (mode == 1 && c.L) || (mode == 3 && c.T && c.L);
        int satd = 0;
        if (ok) {
            uint32_t pw[4];
            i16_pred_block(c, p, mode, bx, by, pw);
            satd = cc_satd4x4_pre(tc, pw);
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
/* thresh: Intra4x4 can only be chosen with a cost below it, and the running cost never decreases, so the analysis stops as
 * soon as it is reached ([EXT] x264_mb_analyse_intra's i_satd_thresh early termination).  The decision is the same as
 * without the early exit; only the work differs.  Returns 1 << 30 when abandoned. */
/* proj (P frames): also give up when, from the step that completes the fourth block on, the cost so far projected to sixteen
 * blocks exceeds thresh by a quarter (oracle/ref_mb.c analyse_i4 states the rule and what it costs: it halves the steps taken in
 * P frames, where one analysis in a hundred ends in an Intra4x4 macroblock).  Unlike the exact test this one is a decision rule. */
__device__ __forceinline__ int analyse_i4(const cc_kparams &k, const MbCtx &c, int thresh, bool proj) {
    WarpMem *m = c.m;
    const int lane = c.lane, half = lane >> 4, l = lane & 15;
    /* neighbour macroblocks' modes (global memory): one round trip for the whole macroblock, before the block loop */
    if (lane < 8) {
        const int top = lane >> 2, i = lane & 3;
        int v = 255;
        if (top ? c.T : c.L) {
            const cc_mbdev *n = k.mbd + c.mb - (top ? k.mb_w : 1);
            v = __ldcg(&n->type) == CC_MB_I4 ? __ldcg(&n->i4_mode[c_blk_of[top ? 12 + i : i * 4 + 3]]) : 2;
        }
        m->nbm[lane] = (uint8_t)v;
    }
    /* this lane's mode as 16 source-array indices (lanes 9..15 of a half never use theirs) */
    uint32_t sx[4];
#pragma unroll
    for (int r = 0; r < 4; r++) sx[r] = g_i4idx[l][r];
    const int md = l < 9 ? l : 2;
    /* neighbours this lane's mode predicts from (H.264 8.3.1.2): V, DDL, VL need the top row; H, HU the left column; DDR, VR, HD both */
    const bool need_top = md == 0 || md == 3 || md == 7 || md == 4 || md == 5 || md == 6;
    const bool need_left = md == 1 || md == 8 || md == 4 || md == 5 || md == 6;
    uint8_t *S = m->S[half];
    int cost = 24 * c.lambda;
    /* The block's transform chain is split over its half-warp: lane l owns coefficient l (raster: vertical frequency l >> 2,
     * horizontal l & 3) through forward transform, quantisation and level scaling; the sixteen scaled coefficients are then
     * gathered by shuffles for the inverse transform.  Per-lane constants of that coefficient: */
    const int fu = l & 3, fv = l >> 2;
    const uint32_t cfu = fu == 0 ? CC_CF0 : (fu == 1 ? CC_CF1 : (fu == 2 ? CC_CF2 : CC_CF3));
    const uint32_t cfun = fu == 0 ? CC_CF0N : (fu == 1 ? CC_CF1N : (fu == 2 ? CC_CF2N : CC_CF3N));
    const int av0 = fv == 1 ? 2 : 1, av1 = fv == 0 ? 1 : (fv == 1 ? 1 : (fv == 2 ? -1 : -2)); /* column fv of Cf^T: Cf[fv][0..3] */
    const int av2 = fv == 0 ? 1 : (fv == 1 ? -1 : (fv == 2 ? -1 : 2)), av3 = fv == 0 ? 1 : (fv == 1 ? -2 : (fv == 2 ? 1 : -1));
    const int qcls = ((l & 1) && (fv & 1)) ? 1 : (((l & 1) || (fv & 1)) ? 2 : 0);
    const int qmf = k.qmf[0][qcls], qbias = k.qbias[0][qcls];
    const int dqb = c.qp / 6 - 4, dqsh = dqb < 0 ? -dqb : 0;
    const int dqv = (c_dequant_v[c.qp % 6][qcls] * 16) << (dqb > 0 ? dqb : 0);
    const int dqrnd = dqsh ? 1 << (dqsh - 1) : 0;
    const int spos = (int)((0xfea9db83c7426510ull >> (4 * l)) & 15u); /* scan position of raster coefficient l (inverse zigzag) */
    __syncwarp();
    /* The sixteen blocks depend on their left, top, top-left and top-right neighbours only, so blocks with equal bx + 2*by are
     * independent: 10 steps instead of 16, the two half-warps taking one block each (the predictors available to a block still
     * follow decoding order, H.264 6.4.11.4).  Costs add up in a different order; their sum and every decision are the same. */
    for (int t = 0; t < 10; t++) {
        /* half 0: the block of the step with the smaller by; half 1: the one a block row further down, if any */
        const int by0 = t >= 2 ? (t - 2) >> 1 : 0; /* smallest by with 0 <= t - 2*by <= 3 */
        const int by = by0 + half, bx = t - 2 * by;
        const bool act = by < 4 && bx >= 0 && bx < 4;
        const int bxs = act ? bx : 0, bys = act ? by : 0; /* safe coordinates for the idle half */
        const int blk = c_blk_of[bys * 4 + bxs];
        const int has_top = bys > 0 || c.T, has_left = bxs > 0 || c.L;
        int has_tr;
        if (bys == 0) has_tr = bxs < 3 ? c.T : c.TR;
        else has_tr = !(bxs == 3 || blk == 3 || blk == 11 || blk == 7 || blk == 13 || blk == 15);
        /* edge sample of this lane (E index = l: left3..left0, top-left, top0..7), its filtered forms, the DC value */
        int e = 0;
        if (l < 13) {
            if (l < 4) e = m->tile[1 + bys * 4 + (3 - l)][3 + bxs * 4];
            else if (l == 4) e = m->tile[bys * 4][3 + bxs * 4];
            else { const int i = l - 5; e = m->tile[bys * 4][4 + bxs * 4 + ((i < 4 || has_tr) ? i : 3)]; }
        }
        {
            int a = __shfl_up_sync(FULL, e, 1, 16), cdn = __shfl_down_sync(FULL, e, 1, 16);
            if (l == 0) a = e;
            if (l == 12) cdn = e;
            int sums = (l < 4 ? e : 0) | ((l >= 5 && l < 9 ? e : 0) << 16); /* left sum | top sum << 16, over the half-warp */
#pragma unroll
            for (int sft = 1; sft < 16; sft <<= 1) sums += __shfl_xor_sync(FULL, sums, sft);
            const int sl = sums & 0xffff, st = sums >> 16;
            if (l < 13) {
                S[l] = (uint8_t)e;
                S[16 + l] = (uint8_t)((a + 2 * e + cdn + 2) >> 2);
                if (l < 12) S[32 + l] = (uint8_t)((e + cdn + 1) >> 1);
            } else if (l == 13)
                S[13] = (uint8_t)((has_top && has_left) ? (st + sl + 4) >> 3 : (has_top ? (st + 2) >> 2 : (has_left ? (sl + 2) >> 2 : 128)));
        }
        /* predicted mode, H.264 8.3.1.1 */
        int mA, mB;
        if (bxs > 0) mA = m->m4[c_blk_of[bys * 4 + bxs - 1]];
        else { mA = m->nbm[bys]; if (mA == 255) mA = -1; }
        if (bys > 0) mB = m->m4[c_blk_of[(bys - 1) * 4 + bxs]];
        else { mB = m->nbm[4 + bxs]; if (mB == 255) mB = -1; }
        const int pm = (mA < 0 || mB < 0) ? 2 : min(mA, mB);
        __syncwarp();
        uint32_t cw[4], pw[4] = {0, 0, 0, 0};
#pragma unroll
        for (int r = 0; r < 4; r++) cw[r] = m->curY[(bys * 4 + r) * 4 + bxs];
        unsigned key = 0xffffffffu;
        {
            const bool ok = l < 9 && (has_top || !need_top) && (has_left || !need_left);
            if (ok) {
#pragma unroll
                for (int r = 0; r < 4; r++)
                    pw[r] = (uint32_t)S[sx[r] & 255u] | ((uint32_t)S[(sx[r] >> 8) & 255u] << 8) | ((uint32_t)S[(sx[r] >> 16) & 255u] << 16) |
                            ((uint32_t)S[sx[r] >> 24] << 24);
                int satd = cc_satd4x4_words(cw, pw);
                key = ((unsigned)(satd + c.lambda * (md == pm ? 1 : 4)) << 4) | (unsigned)md;
            }
        }
        unsigned bk = key; /* minimum over the half-warp */
#pragma unroll
        for (int sft = 1; sft < 16; sft <<= 1) bk = min(bk, __shfl_xor_sync(FULL, bk, sft));
        const int bm = (int)(bk & 15u);
        {
            const int mine = act ? (int)(bk >> 4) : 0;
            cost += mine + __shfl_xor_sync(FULL, mine, 16);
        }
        if (cost >= thresh) return 1 << 30; /* uniform across the warp */
        if (proj && t >= 2 && t < 9) {
            const int nd = t < 8 ? 2 * t : 15; /* blocks done after step t: 1, 2, 4, 6, 8, 10, 12, 14, 15, 16 */
            if ((long long)(cost - 24 * c.lambda) * 64 > (long long)(thresh - 24 * c.lambda) * nd * 5) return 1 << 30;
        }
#pragma unroll
        for (int r = 0; r < 4; r++) pw[r] = __shfl_sync(FULL, pw[r], bm, 16);
        /* forward transform of this lane's coefficient (horizontal pass as byte dot products, then its column of Cf^T),
         * quantisation, level scaling */
        int th[4];
#pragma unroll
        for (int r = 0; r < 4; r++) th[r] = cc_dp4a_us(cw[r], cfu, cc_dp4a_us(pw[r], cfun, 0));
        const int coef = av0 * th[0] + av1 * th[1] + av2 * th[2] + av3 * th[3];
        const int level = quant_one(coef, qmf, qbias);
        const int nnz = __popc((__ballot_sync(FULL, level != 0) >> (16 * half)) & 0xffffu);
        const int dq = (level * dqv + dqrnd) >> dqsh;
        /* inverse transform, H.264 8.5.12.2, also split: row pass in lane (row fv, column fu) from the four scaled coefficients
         * of its quad, column pass from the four row-pass values of its column (lanes fu, fu + 4, fu + 8, fu + 12); the lane
         * ends up with the reconstructed sample (fu, fv).  All-zero levels give residual 0: no special case. */
        int px;
        {
            const int q0 = l & ~3;
            const int a0 = __shfl_sync(FULL, dq, q0, 16), a1 = __shfl_sync(FULL, dq, q0 + 1, 16);
            const int a2 = __shfl_sync(FULL, dq, q0 + 2, 16), a3 = __shfl_sync(FULL, dq, q0 + 3, 16);
            const int e0 = a0 + a2, e1 = a0 - a2, e2 = (a1 >> 1) - a3, e3 = a1 + (a3 >> 1);
            const int t = fu == 0 ? e0 + e3 : (fu == 1 ? e1 + e2 : (fu == 2 ? e1 - e2 : e0 - e3));
            const int b0 = __shfl_sync(FULL, t, fu, 16), b1 = __shfl_sync(FULL, t, fu + 4, 16);
            const int b2 = __shfl_sync(FULL, t, fu + 8, 16), b3 = __shfl_sync(FULL, t, fu + 12, 16);
            const int g0 = b0 + b2, g1 = b0 - b2, g2 = (b1 >> 1) - b3, g3 = b1 + (b3 >> 1);
            const int o = fv == 0 ? g0 + g3 : (fv == 1 ? g1 + g2 : (fv == 2 ? g1 - g2 : g0 - g3));
            const uint32_t prow = fv == 0 ? pw[0] : (fv == 1 ? pw[1] : (fv == 2 ? pw[2] : pw[3]));
            px = cc_clip255((int)((prow >> (8 * fu)) & 255u) + ((o + 32) >> 6));
        }
        if (act) {
            m->c4[blk][spos] = (int16_t)level;
            m->tile[1 + by * 4 + fv][4 + bx * 4 + fu] = (uint8_t)px;
            if (l == 0) {
                m->m4[blk] = (uint8_t)bm;
                m->m4code[blk] = (uint8_t)(bm == pm ? 8 : (bm < pm ? bm : bm - 1));
                m->n4[blk] = (uint8_t)nnz;
            }
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
/* rX: -2 unavailable, -1 intra, 0 inter (the one reference picture) */
__device__ __forceinline__ void predict_mv_core(int L, int T, int rA, int a[2], int rB, int b[2], int rC, int cc[2], int mvp[2], int skip[2]) {
    const int rA0 = rA, rB0 = rB, a0x = a[0], a0y = a[1], b0x = b[0], b0y = b[1];
    if (rB == -2 && rC == -2 && rA != -2) { rB = rC = rA; b[0] = cc[0] = a[0]; b[1] = cc[1] = a[1]; }
    const int n = (rA == 0) + (rB == 0) + (rC == 0);
    if (n == 1) {
        if (rA == 0) { mvp[0] = a[0]; mvp[1] = a[1]; }
        else if (rB == 0) { mvp[0] = b[0]; mvp[1] = b[1]; }
        else { mvp[0] = cc[0]; mvp[1] = cc[1]; }
    } else { mvp[0] = cc_median3(a[0], b[0], cc[0]); mvp[1] = cc_median3(a[1], b[1], cc[1]); }
    /* P_Skip vector, 8.4.1.1 */
    if (!L || !T || (rA0 == 0 && a0x == 0 && a0y == 0) || (rB0 == 0 && b0x == 0 && b0y == 0)) skip[0] = skip[1] = 0;
    else { skip[0] = mvp[0]; skip[1] = mvp[1]; }
}
/* quadrant q of a neighbour record (NbRec, cc_deblock.cuh) */
__device__ __forceinline__ void nb_mv2(const NbRec *n, int q, int avail, int &ref, int mv[2]) {
    mv[0] = mv[1] = 0;
    if (!avail) { ref = -2; return; }
    if ((n->x & 255u) >= CC_MB_I16) { ref = -1; return; }
    const uint32_t v = n->mv[q];
    ref = 0; mv[0] = (int)(int16_t)(v & 0xffffu); mv[1] = (int)(int16_t)(v >> 16);
}
/* 16x16 partition: A = left MB quadrant 1, B = top MB quadrant 2, C = top-right MB quadrant 2, else D = top-left MB quadrant 3
 * (the quadrants that contain the neighbouring luma locations of 6.4.11.7).  nbl: this row's line, nbu: the line of the row
 * above (valid where c.T) */
__device__ __forceinline__ void predict_mv(const MbCtx &c, const NbRec *nbl, const NbRec *nbu, int mvp[2], int skip[2]) {
    int rA, rB, rC, a[2], b[2], cc[2];
    nb_mv2(nbl + c.mbx - 1, 1, c.L, rA, a);
    nb_mv2(nbu + c.mbx, 2, c.T, rB, b);
    nb_mv2(nbu + c.mbx + 1, 2, c.TR, rC, cc);
    if (rC == -2) nb_mv2(nbu + c.mbx - 1, 3, c.T && c.L, rC, cc);
    predict_mv_core(c.L, c.T, rA, a, rB, b, rC, cc, mvp, skip);
}
/* quadrant p of a P_8x8 macroblock; own = the macroblock's four vectors (quadrants before p are decided).  oracle/ref_mb.c predict_mv8 */
__device__ __forceinline__ void predict_mv8(const MbCtx &c, const NbRec *nbl, const NbRec *nbu, int p, const cc_me8 &own, int mvp[2]) {
    int rA, rB, rC, a[2], b[2], cc[2], dummy[2];
#define OWN(q, r, v) do { r = 0; v[0] = own.mv[q][0]; v[1] = own.mv[q][1]; } while (0)
    if (p == 0) {
        nb_mv2(nbl + c.mbx - 1, 1, c.L, rA, a);
        nb_mv2(nbu + c.mbx, 2, c.T, rB, b);
        nb_mv2(nbu + c.mbx, 3, c.T, rC, cc);
        if (rC == -2) nb_mv2(nbu + c.mbx - 1, 3, c.T && c.L, rC, cc);
    } else if (p == 1) {
        OWN(0, rA, a);
        nb_mv2(nbu + c.mbx, 3, c.T, rB, b);
        nb_mv2(nbu + c.mbx + 1, 2, c.TR, rC, cc);
        if (rC == -2) nb_mv2(nbu + c.mbx, 2, c.T, rC, cc);
    } else if (p == 2) {
        nb_mv2(nbl + c.mbx - 1, 3, c.L, rA, a);
        OWN(0, rB, b);
        OWN(1, rC, cc);
    } else {
        OWN(2, rA, a);
        OWN(1, rB, b);
        OWN(0, rC, cc); /* C (the macroblock to the right) is not decoded yet -> D = quadrant 0 */
    }
#undef OWN
    predict_mv_core(1, 1, rA, a, rB, b, rC, cc, mvp, dummy);
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
/* picks the chroma mode by SATD over both planes and leaves the predictor in m->predC; returns the mode.
 * All 32 (mode, plane, block) pairs at once: lane = mode * 8 + plane * 4 + block. */
__device__ __forceinline__ int intra_chroma_pred(const MbCtx &c) {
    WarpMem *m = c.m;
    const int lane = c.lane, mode = lane >> 3, pl = (lane >> 2) & 1, b = lane & 3, bx = b & 1, by = b >> 1;
    C8Params p;
    c8_params(c, pl, p);
    uint32_t cw[4], pw[4];
#pragma unroll
    for (int r = 0; r < 4; r++) cw[r] = m->curC[pl][(by * 4 + r) * 2 + bx];
    const bool ok = mode == 0 || (mode == 1 && c.L) || (mode == 2 && c.T) || (mode == 3 && c.T && c.L);
    int satd = 0;
    if (ok) {
        c8_pred_block(c, p, pl, mode, bx, by, pw);
        satd = cc_satd4x4_words(cw, pw);
    }
    satd += __shfl_xor_sync(FULL, satd, 1); satd += __shfl_xor_sync(FULL, satd, 2); satd += __shfl_xor_sync(FULL, satd, 4);
    const int bits = mode == 0 ? 1 : (mode == 3 ? 5 : 3);
    unsigned best = ok ? (((unsigned)(satd + c.lambda * bits) << 2) | (unsigned)mode) : 0xffffffffu;
    best = min(best, __shfl_xor_sync(FULL, best, 8));
    best = min(best, __shfl_xor_sync(FULL, best, 16));
    const int bm = (int)(best & 3u);
    if (lane < 8) {
        c8_pred_block(c, p, pl, bm, bx, by, pw);
#pragma unroll
        for (int r = 0; r < 4; r++) m->predC[pl][(by * 4 + r) * 2 + bx] = pw[r];
    }
    __syncwarp();
    return bm;
}

/* ---- symbol stream -------------------------------------------------------------------------------------------------- */
/* A macroblock's residual in its symbol slot (int16 units, after the slot's prefix: four mode words of an Intra4x4 macroblock
 * or the eight mvd of a P_8x8 one): first one 16-bit significance mask (scan order) per coded block, in the order
 * residual() transmits the blocks (7.3.5.3: [Intra16x16 DC], luma 4x4 blocks of the coded 8x8s, Cb DC, Cr DC, chroma AC), then the
 * non-zero levels of those blocks back to back.  Masks at fixed positions let the entropy coder (k_cavlc_*, or the host
 * writer) find every block's levels with one prefix sum over popcounts instead of a chain of dependent reads. */
__device__ __forceinline__ int mb_nblocks(int type, int cbp) {
    const int cl = cbp & 15, cc = cbp >> 4;
    return (type == CC_MB_I16 ? 1 + (cl ? 16 : 0) : 4 * __popc((unsigned)cl)) + (cc ? 2 : 0) + (cc == 2 ? 8 : 0);
}
/* block `blk` of m->coef (n scan positions): mask -> out[moff], levels -> out[loff...]; returns the new level offset */
__device__ __forceinline__ unsigned emit_block(const MbCtx &c, int16_t *out, unsigned moff, unsigned loff, int blk, int n) {
    const int v = c.lane < n ? c.m->coef[blk][c.lane] : 0;
    const unsigned mask = __ballot_sync(FULL, v != 0);
    if (c.lane == 0) out[moff] = (int16_t)mask;
    if (v != 0) out[loff + __popc(mask & ((1u << c.lane) - 1u))] = (int16_t)v;
    return loff + (unsigned)__popc(mask);
}

/* all residual blocks of one macroblock in syntax order; the slot length is padded to a 4-byte multiple */
template <int UNROLL = 16> /* 1: the block loops stay loops (code size: k_p_code is bound by instruction fetch before anything else) */
__device__ __forceinline__ unsigned emit_mb(const MbCtx &c, int16_t *out, unsigned off, int type, int cbp) {
    const int cl = cbp & 15, cc = cbp >> 4;
    if (type == CC_MB_I16 || cbp) {
        unsigned mo = off;
        off += (unsigned)mb_nblocks(type, cbp);
        if (type == CC_MB_I16) off = emit_block(c, out, mo++, off, 24, 16);
        if (UNROLL == 1) {
#pragma unroll 1
            for (int blk = 0; blk < 16; blk++)
                if (cl & (1 << (blk >> 2))) off = emit_block(c, out, mo++, off, blk, type == CC_MB_I16 ? 15 : 16);
        } else {
            for (int blk = 0; blk < 16; blk++)
                if (cl & (1 << (blk >> 2))) off = emit_block(c, out, mo++, off, blk, type == CC_MB_I16 ? 15 : 16);
        }
        if (cc) {
            off = emit_block(c, out, mo++, off, 25, 4);
            off = emit_block(c, out, mo++, off, 26, 4);
            if (cc & 2) {
                if (UNROLL == 1) {
#pragma unroll 1
                    for (int blk = 16; blk < 24; blk++) off = emit_block(c, out, mo++, off, blk, 15);
                } else {
                    for (int blk = 16; blk < 24; blk++) off = emit_block(c, out, mo++, off, blk, 15);
                }
            }
        }
    }
    if (off & 1u) { if (c.lane == 0) out[off] = 0; off++; }
    return off;
}

/* ---- the kernel --------------------------------------------------------------------------------------------------------- */
__device__ __forceinline__ void wait_progress(volatile int *flag, int need, int lane) {
    if (lane == 0) while (*flag < need) CC_SPIN_PAUSE();
    __syncwarp();
    __threadfence_block();
}

/* the same across CTAs: *flag is a frame-tagged progress word in global memory (tag + macroblocks done); returns the count */
__device__ __forceinline__ int wait_progress_global(const uint32_t *flag, uint32_t tag, int need, int lane) {
    uint32_t v = 0;
    if (lane == 0) {
        for (;;) {
            v = *(const volatile uint32_t *)flag;
            if ((int)(v - (tag + (uint32_t)need)) >= 0) break;
            CC_SPIN_PAUSE();
        }
    }
    v = __shfl_sync(FULL, v, 0);
    __threadfence();
    return (int)(v - tag);
}

/* k_deblock following k_mb_encode: the word must carry THIS frame's tag exactly.  The signed distance test above is right for words
 * that are rewritten every frame; this one is only written while the filter runs next to the encoder, so a handle that switches
 * launch orders could meet a word from 2^19 or more frames ago, which the distance test would take for "ahead" (and cc_enc_submit
 * sees to it that the last word of an earlier stretch never carries the tag of the frame that resumes).  Counts stay below 4096
 * (8192 / 16 macroblocks per row at most). */
__device__ __forceinline__ int wait_encoder_progress(const uint32_t *flag, uint32_t tag, int need, int lane) {
    uint32_t v = 0;
    if (lane == 0) {
        for (;;) {
            v = *(const volatile uint32_t *)flag;
            if (((v ^ tag) >> 12) == 0 && (int)(v & 4095u) >= need) break;
            __nanosleep(1000); /* the word moves every fourth macroblock of the encoder (tens of microseconds): poll slowly */
        }
    }
    v = __shfl_sync(FULL, v, 0);
    __threadfence();
    return (int)(v & 4095u);
}

/* does the macroblock coded at (mvx,mvy) leave nothing after quantisation + decimation?  (P_Skip test) */
/* One pass: lanes 0-15 take the luma blocks, lanes 16-23 the chroma blocks (same arithmetic as inter_luma / encode_chroma in
 * probe mode; the two only differ in quantiser tables, the DC handling of chroma and the decimation rule). */
__device__ __forceinline__ bool probe_eval1(const cc_kparams &k, const MbCtx &c, const uint32_t *pY, const uint32_t (*pC)[16]);
__device__ __forceinline__ bool probe_zero(const cc_kparams &k, const MbCtx &c, int mvx, int mvy) {
    mc_luma16(k, c, mvx, mvy);
    mc_chroma8(k, c, mvx, mvy);
    __syncwarp();
    return probe_eval1(k, c, c.m->predY, c.m->predC);
}
/* the test on a prediction that is in shared memory already */
__device__ __forceinline__ bool probe_eval1(const cc_kparams &k, const MbCtx &c, const uint32_t *pY, const uint32_t (*pC)[16]) {
    WarpMem *m = c.m;
    const int lane = c.lane;
    const bool chroma = lane >= 16;
    const int pl = (lane >> 2) & 1, b = chroma ? (lane & 3) : lane;
    const int bx = chroma ? (b & 1) : c_blk_x[b], by = chroma ? (b >> 1) : c_blk_y[b];
    uint32_t cw[4], pw[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        cw[r] = chroma ? m->curC[pl][(by * 4 + r) * 2 + bx] : m->curY[(by * 4 + r) * 4 + bx];
        pw[r] = chroma ? pC[pl][(by * 4 + r) * 2 + bx] : pY[(by * 4 + r) * 4 + bx];
    }
    uint16_t mf[3], bias[3];
#pragma unroll
    for (int i = 0; i < 3; i++) { mf[i] = chroma ? k.qmf[3][i] : k.qmf[1][i]; bias[i] = chroma ? k.qbias[3][i] : k.qbias[1][i]; }
    int d[16];
    dct4x4_res(cw, pw, d);
    const int dc = d[0];
    /* only the positions of the non-zero levels and whether any exceeds 1 matter here: threshold compares, no quantiser */
    uint16_t thr[3][2];
#pragma unroll
    for (int i = 0; i < 3; i++) { thr[i][0] = chroma ? k.qthr[3][i][0] : k.qthr[1][i][0]; thr[i][1] = chroma ? k.qthr[3][i][1] : k.qthr[1][i][1]; }
    unsigned nzm, big;
    quant_masks(d, thr, nzm, big);
    if (chroma) { nzm >>= 1; big >>= 1; } /* chroma AC blocks: the DC goes through the 2x2 transform below; run lengths start at position 1 */
    const int nnz = __popc(nzm);
    const int score = decimate_masks(nzm, big);
    int s8 = score + __shfl_xor_sync(FULL, score, 1);
    s8 += __shfl_xor_sync(FULL, s8, 2);              /* luma: the 8x8 block's score; chroma: the plane's */
    int tot = s8 + __shfl_xor_sync(FULL, s8, 4);
    tot += __shfl_xor_sync(FULL, tot, 8);            /* luma: the macroblock's score (lanes 0-15) */
    int nac = nnz > 0;
    nac |= __shfl_xor_sync(FULL, nac, 1);
    nac |= __shfl_xor_sync(FULL, nac, 2);
    const int d0 = __shfl_sync(FULL, dc, 0, 4), d1 = __shfl_sync(FULL, dc, 1, 4), d2 = __shfl_sync(FULL, dc, 2, 4), d3 = __shfl_sync(FULL, dc, 3, 4);
    bool nz;
    if (!chroma) nz = nnz > 0 && !(tot < 6 || s8 < 4);
    else {
        const int a = d0 + d1, bb = d0 - d1, cc = d2 + d3, e = d2 - d3;
        const int ndc = (quant_one(a + cc, mf[0] >> 1, bias[0] << 1) != 0) | (quant_one(bb + e, mf[0] >> 1, bias[0] << 1) != 0) |
                        (quant_one(a - cc, mf[0] >> 1, bias[0] << 1) != 0) | (quant_one(bb - e, mf[0] >> 1, bias[0] << 1) != 0);
        nz = (nac && s8 >= 7) || ndc;
    }
    const bool any = __ballot_sync(FULL, lane < 24 && nz) != 0;
    __syncwarp();
    return !any;
}

/* Two P_Skip tests in one pass (the pre-pass probes its two likeliest skip vectors): prediction A in predY / predC, prediction B
 * in the Intra4x4 level scratch c4 (unused by the pre-pass; recY / recC keep the macroblock's reconstruction); lanes 0-15 run the luma blocks of A and lanes 16-31 those of B, lanes 0-7 / 8-15 the chroma
 * blocks.  Same arithmetic as inter_luma / encode_chroma in probe mode.  Returns bit 0: nothing survives at A, bit 1: at B. */
__device__ __forceinline__ int probe_eval2(const cc_kparams &k, const MbCtx &c, const uint32_t *aY, const uint32_t (*aC)[16], const uint32_t *bY, const uint32_t (*bC)[16]);
__device__ __forceinline__ int probe_zero2(const cc_kparams &k, const MbCtx &c, int ax, int ay, int bx2, int by2) {
    WarpMem *m = c.m;
    mc_luma16(k, c, ax, ay);
    mc_chroma8(k, c, ax, ay);
    uint32_t *const bY = (uint32_t *)m->c4;
    uint32_t (*const bC)[16] = (uint32_t (*)[16])(bY + 64);
    mc_luma16(k, c, bx2, by2, bY);
    mc_chroma8(k, c, bx2, by2, bC);
    __syncwarp();
    return probe_eval2(k, c, m->predY, m->predC, bY, bC);
}
/* the two tests on predictions that are in shared memory already: A = (aY, aC), B = (bY, bC) */
__device__ __forceinline__ int probe_eval2(const cc_kparams &k, const MbCtx &c, const uint32_t *aY, const uint32_t (*aC)[16], const uint32_t *bY, const uint32_t (*bC)[16]) {
    WarpMem *m = c.m;
    const int lane = c.lane;
    unsigned nzl;
    { /* luma */
        const int b = lane & 15, bx = c_blk_x[b], by = c_blk_y[b];
        const uint32_t *pred = lane < 16 ? aY : bY;
        uint32_t cw[4], pw[4];
#pragma unroll
        for (int r = 0; r < 4; r++) { cw[r] = m->curY[(by * 4 + r) * 4 + bx]; pw[r] = pred[(by * 4 + r) * 4 + bx]; }
        int d[16];
        dct4x4_res(cw, pw, d);
        unsigned nzm, big;
        quant_masks(d, k.qthr[1], nzm, big);
        int nnz = __popc(nzm);
        const int score = decimate_masks(nzm, big);
        int s8 = score + __shfl_xor_sync(FULL, score, 1);
        s8 += __shfl_xor_sync(FULL, s8, 2);
        int tot = s8 + __shfl_xor_sync(FULL, s8, 4);
        tot += __shfl_xor_sync(FULL, tot, 8); /* sum over the 16 lanes of this half */
        if (tot < 6 || s8 < 4) nnz = 0;
        nzl = __ballot_sync(FULL, nnz > 0);
    }
    unsigned nzc;
    { /* chroma: lanes 0-15 = (probe, plane, block) */
        const int sel = (lane >> 3) & 1, pl = (lane >> 2) & 1, b = lane & 3, bx = b & 1, by = b >> 1;
        const uint32_t *pred = sel ? bC[pl] : aC[pl];
        const uint16_t *mf = k.qmf[3], *bias = k.qbias[3];
        uint32_t cw[4], pw[4];
#pragma unroll
        for (int r = 0; r < 4; r++) { cw[r] = m->curC[pl][(by * 4 + r) * 2 + bx]; pw[r] = pred[(by * 4 + r) * 2 + bx]; }
        int d[16];
        dct4x4_res(cw, pw, d);
        const int dc = d[0];
        unsigned nzm, big;
        quant_masks(d, k.qthr[3], nzm, big);
        nzm >>= 1; big >>= 1; /* AC only */
        const int nnz = __popc(nzm);
        int score = decimate_masks(nzm, big);
        score += __shfl_xor_sync(FULL, score, 1);
        score += __shfl_xor_sync(FULL, score, 2);
        int nac = nnz > 0;
        nac |= __shfl_xor_sync(FULL, nac, 1);
        nac |= __shfl_xor_sync(FULL, nac, 2);
        if (score < 7) nac = 0;
        const int d0 = __shfl_sync(FULL, dc, 0, 4), d1 = __shfl_sync(FULL, dc, 1, 4), d2 = __shfl_sync(FULL, dc, 2, 4), d3 = __shfl_sync(FULL, dc, 3, 4);
        const int a = d0 + d1, bb = d0 - d1, cc = d2 + d3, e = d2 - d3;
        const int ndc = (quant_one(a + cc, mf[0] >> 1, bias[0] << 1) != 0) | (quant_one(bb + e, mf[0] >> 1, bias[0] << 1) != 0) |
                        (quant_one(a - cc, mf[0] >> 1, bias[0] << 1) != 0) | (quant_one(bb - e, mf[0] >> 1, bias[0] << 1) != 0);
        nzc = __ballot_sync(FULL, lane < 16 && (nac || ndc));
    }
    __syncwarp();
    return (((nzl & 0xffffu) | (nzc & 0xffu)) == 0 ? 1 : 0) | (((nzl >> 16) | (nzc & 0xff00u)) == 0 ? 2 : 0);
}

/* Unfiltered neighbour samples.  Intra prediction uses the reconstruction BEFORE deblocking (H.264 8.3), while the
 * filter warps of the same slice modify k.rec a few macroblocks behind the encode warps, and in P frames most
 * macroblocks are reconstructed by the parallel pre-pass, not by this warp.  So whoever reconstructs a macroblock
 * also writes its bottom sample line to k.ubot and its right column to k.uright, and intra analysis reads only those. */
__device__ __forceinline__ void publish_unfiltered(const cc_kparams &k, int lane, int mb, int mbx, int row, const uint32_t *recY /*64 words*/,
                                                   const uint32_t *recC0, const uint32_t *recC1 /*16 words each*/) {
    if (lane < 4) *(uint32_t *)(k.ubot[0] + (size_t)row * k.sy + mbx * 16 + 4 * lane) = recY[60 + lane];
    else if (lane < 6) *(uint32_t *)(k.ubot[1] + (size_t)row * k.sc + mbx * 8 + 4 * (lane - 4)) = recC0[14 + lane - 4];
    else if (lane < 8) *(uint32_t *)(k.ubot[2] + (size_t)row * k.sc + mbx * 8 + 4 * (lane - 6)) = recC1[14 + lane - 6];
    uint8_t *ur = k.uright + (size_t)mb * 32;
    if (lane < 16) ur[lane] = (uint8_t)(recY[lane * 4 + 3] >> 24);
    else if (lane < 24) ur[lane] = (uint8_t)(recC0[(lane - 16) * 2 + 1] >> 24);
    else ur[lane] = (uint8_t)(recC1[(lane - 24) * 2 + 1] >> 24);
}
/* recY/recC of the warp's WarpMem -> k.rec, k.ubot, k.uright, and the macroblock's 8x8 block of the half-resolution reference
 * (2x2 box means of the UNFILTERED reconstruction: the coarse search of the next frame then needs nothing from the in-loop filter
 * and runs next to it; oracle/ref_enc.c writes the same plane) */
__device__ __forceinline__ void publish_pixels(const cc_kparams &k, const MbCtx &c) {
    WarpMem *m = c.m;
    const int lane = c.lane, r = lane >> 1, h = lane & 1;
    uint8_t *ry = k.rec[0] + (size_t)c.py * k.sy + c.px;
    *(uint32_t *)(ry + (size_t)r * k.sy + h * 8) = m->recY[r * 4 + h * 2];
    *(uint32_t *)(ry + (size_t)r * k.sy + h * 8 + 4) = m->recY[r * 4 + h * 2 + 1];
    const int pl = lane >> 4, cr = (lane >> 1) & 7;
    *(uint32_t *)(k.rec[1 + pl] + (size_t)(c.mby * 8 + cr) * k.sc + c.mbx * 8 + 4 * h) = m->recC[pl][cr * 2 + h];
    if (lane < 16) { /* half-resolution row r (lanes 0..15: r = 0..7), 4 pixels of its half h */
        const uint32_t a0 = m->recY[(2 * r) * 4 + 2 * h], a1 = m->recY[(2 * r) * 4 + 2 * h + 1];
        const uint32_t b0 = m->recY[(2 * r + 1) * 4 + 2 * h], b1 = m->recY[(2 * r + 1) * 4 + 2 * h + 1];
        uint32_t v = (uint32_t)((cc_byte(a0, 0) + cc_byte(a0, 1) + cc_byte(b0, 0) + cc_byte(b0, 1) + 2) >> 2);
        v |= (uint32_t)((cc_byte(a0, 2) + cc_byte(a0, 3) + cc_byte(b0, 2) + cc_byte(b0, 3) + 2) >> 2) << 8;
        v |= (uint32_t)((cc_byte(a1, 0) + cc_byte(a1, 1) + cc_byte(b1, 0) + cc_byte(b1, 1) + 2) >> 2) << 16;
        v |= (uint32_t)((cc_byte(a1, 2) + cc_byte(a1, 3) + cc_byte(b1, 2) + cc_byte(b1, 3) + 2) >> 2) << 24;
        *(uint32_t *)(k.l1ref + (size_t)(c.mby * 8 + r) * k.sc + c.mbx * 8 + 4 * h) = v;
    }
    publish_unfiltered(k, lane, c.mb, c.mbx, c.mby, m->recY, m->recC[0], m->recC[1]);
}
__device__ __forceinline__ void stage_cur(const cc_kparams &k, const MbCtx &c) {
    WarpMem *m = c.m;
    const int lane = c.lane;
    const uint8_t *cy = k.cur[0] + (size_t)c.py * k.sy + c.px;
    m->curY[lane] = *(const uint32_t *)(cy + (size_t)(lane >> 2) * k.sy + 4 * (lane & 3));
    m->curY[lane + 32] = *(const uint32_t *)(cy + (size_t)(8 + (lane >> 2)) * k.sy + 4 * (lane & 3));
    const int pl = lane >> 4, r = (lane >> 1) & 7, h = lane & 1;
    m->curC[pl][r * 2 + h] = *(const uint32_t *)(k.cur[1 + pl] + (size_t)(c.mby * 8 + r) * k.sc + c.mbx * 8 + 4 * h);
}
/* neighbour samples for intra prediction into m->tile (row 0, column 3) and m->edgeC */
__device__ __forceinline__ void stage_neighbours(const cc_kparams &k, const MbCtx &c) {
    WarpMem *m = c.m;
    const int lane = c.lane;
    const uint8_t *ur = k.uright + (size_t)(c.mb - 1) * 32;
    if (lane < 21) { /* top-left, top, top-right */
        const int x = lane - 1;
        const bool ok = c.T && (x < 0 ? c.L : (x < 16 ? true : c.TR));
        m->tile[0][3 + lane] = ok ? __ldcg(k.ubot[0] + (size_t)(c.mby - 1) * k.sy + c.px + x) : 0; /* .cg: the row above may be another CTA's */
    }
    if (lane >= 16) m->tile[1 + lane - 16][3] = c.L ? ur[lane - 16] : 0;
    if (lane < 18) {
        const int p2 = lane >= 9, i = lane - 9 * p2; /* i = 0: top-left, 1..8: top */
        const bool ok = i == 0 ? (c.T && c.L) : c.T;
        m->edgeC[p2][i] = ok ? __ldcg(k.ubot[1 + p2] + (size_t)(c.mby - 1) * k.sc + c.mbx * 8 + i - 1) : 0;
    }
    if (lane < 16) m->edgeC[lane >> 3][9 + (lane & 7)] = c.L ? ur[16 + lane] : 0;
}

/* srow0: first row of the slice; last: this is the band's last row and the slice goes on below it (another CTA follows this row) */
__device__ void encode_row(const cc_kparams &k, WarpMem *m, NbRec *nbline, volatile int *progress, int w, int row, int srow0, bool last, int lane) {
    const int mw = k.mb_w, qp = k.qp;
    MbCtx c;
    c.lane = lane; c.mby = row; c.py = row * 16; c.qp = qp; c.qpc = c_chroma_qp[qp]; c.lambda = c_lambda[qp]; c.m = m;
    c.T = row > srow0;
    unsigned total = 0, n_skip = 0, n_inter = 0, n_intra = 0, n_gate = 0, n_probe = 0;
    NbRec *nbl = nbline + (w + 1) * mw; /* line 0 = the row above the band */
    NbRec *nbu = nbl - mw;
    const bool xband = c.T && w == 0;   /* the row above belongs to the band above */
    const uint32_t tag = (uint32_t)k.frame_seq << 12;
    int up_loaded = 0;                  /* records of the row above copied into line 0 so far */
    /* Software pipeline: everything a macroblock needs from global memory that does not depend on its neighbours is requested
     * three macroblocks ahead of its turn by cp.async into shared-memory rings (no register waits on it until it is read), so the
     * latency (HBM when 8 streams share the L2) overlaps earlier macroblocks' work instead of heading the serial chain: the
     * search / pre-pass records and the source pixels (fresh from the ingest kernel, read once: an HBM miss each). */
    auto prefetch = [&](int x) { /* one commit group per call, also when there is nothing left to request */
        if (x < mw) {
            uint32_t *slot = (uint32_t *)m->pc[x & 3];
            const void *src; uint32_t *dst;
            if (lane < 16) { src = k.cur[0] + (size_t)(c.py + lane) * k.sy + x * 16; dst = slot + 4 * lane; }
            else { const int pl = (lane - 16) >> 3, r = (lane - 16) & 7; src = k.cur[1 + pl] + (size_t)(row * 8 + r) * k.sc + x * 8; dst = slot + 64 + 16 * pl + 2 * r; }
#ifndef CC_EMU_BUILD
            if (lane < 16) asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
            else asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
#else
            if (lane < 16) *(uint4 *)dst = *(const uint4 *)src; else *(uint2 *)dst = *(const uint2 *)src;
#endif
            if (!k.is_idr && lane < (k.part8 ? 4 : 2)) {
                const int mb = row * mw + x;
                const void *rs = lane == 0 ? (const void *)(k.me + mb) : (lane == 1 ? (const void *)(k.pre + mb) : (const void *)((const uint4 *)(k.me8 + mb) + (lane - 2)));
#ifndef CC_EMU_BUILD
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(&m->pf[x & 3][lane])), "l"(rs) : "memory");
#else
                m->pf[x & 3][lane] = *(const uint4 *)rs;
#endif
            }
        }
#ifndef CC_EMU_BUILD
        asm volatile("cp.async.commit_group;" ::: "memory");
#endif
    };
    prefetch(0); prefetch(1); prefetch(2);
    long long tprev = k.cyc ? clock64() : 0;
    for (int mbx = 0; mbx < mw; mbx++) {
        c.mbx = mbx; c.px = mbx * 16; c.mb = row * mw + mbx; c.L = mbx > 0; c.TR = c.T && mbx + 1 < mw;
        int16_t *out = (int16_t *)(k.rowbuf + (size_t)c.mb * CC_ROW_PAYLOAD_PER_MB);
        unsigned off = 0;
        cc_me me = {{0, 0}, {0, 0}, 0, 0};
        cc_pre pre = {0, 0, 0, {0, 0}, 0, 0, 0, 0};
        cc_me8 me8 = {{{0, 0}, {0, 0}, {0, 0}, {0, 0}}, {0, 0, 0, 0}}; /* zeros unless me.part (k_me_l0 clears the record of an unsplit macroblock) */
        prefetch(mbx + 3);
#ifndef CC_EMU_BUILD
        asm volatile("cp.async.wait_group 3;" ::: "memory"); /* all but the three newest groups: this macroblock's records and pixels are in */
#endif
        __syncwarp();
        if (!k.is_idr) {
            const uint4 *pf = m->pf[mbx & 3];
            *(uint4 *)&me = pf[0];
            *(uint4 *)&pre = pf[1];
            if (k.part8) { ((uint4 *)&me8)[0] = pf[2]; ((uint4 *)&me8)[1] = pf[3]; }
        }
        const int part = me.part;
        { /* = stage_cur, from the ring */
            const uint32_t *slot = (const uint32_t *)m->pc[mbx & 3];
            m->curY[lane] = slot[lane]; m->curY[lane + 32] = slot[lane + 32]; m->curC[lane >> 4][lane & 15] = slot[64 + lane];
        }
#ifndef CC_EMU_BUILD
        /* The in-loop filter, next in the stream, is launched when every CTA of this kernel has come this far (a no-op unless that
         * launch carries the programmatic attribute).  Not at the top: the filter is more than twice as fast as this kernel, and a
         * filter CTA that was resident from the start would hold its SM's registers for the whole wavefront just to wait; released
         * at k.trail_at macroblocks it runs at its own speed and catches up with the encoder near the end of the row. */
        if (k.trail && w == 0 && mbx == k.trail_at) asm volatile("griddepcontrol.launch_dependents;");
#endif
        const long long t0 = k.cyc ? clock64() : 0;
        if (xband) { /* the band above: its progress word, then the records it published since */
            const int have = min(wait_progress_global(k.gprog + row - 1, tag, min(mbx + 2, mw), lane), mw);
            const uint32_t *src = k.nbrec + (size_t)(row - 1) * mw * 5;
            for (int i = up_loaded * 5 + lane; i < have * 5; i += 32) ((uint32_t *)nbu)[i] = __ldcg(src + i);
            up_loaded = have;
            __syncwarp();
        } else if (c.T) wait_progress(progress + w - 1, min(mbx + 2, mw), lane);
        const long long t1 = k.cyc ? clock64() : 0;
        int path = 0; /* 0 fast, 1 skip probe, 2 intra analysis */

        int type = CC_MB_P16, mode16 = 2, chroma_mode = 0, cbp = 0, mvx = 0, mvy = 0, mvdx = 0, mvdy = 0, t8 = 0;
        unsigned nzmask = 0;
        int mvp[2] = {0, 0}, skipmv[2] = {0, 0};
        int mvp8[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};
        int cost_inter = 1 << 30, cost_i4 = 1 << 30;
        bool done = false;

        if (!k.is_idr) {
            predict_mv(c, nbl, nbu, mvp, skipmv);
            /* [EXT] x264 probe_skip: would the MB coded at the P_Skip vector leave nothing after decimation?  The
             * pre-pass answered that for its search vector and for the two likeliest skip vectors. */
            bool zero, pred_ready = false;
            if (!part && skipmv[0] == me.mv[0] && skipmv[1] == me.mv[1]) zero = pre.cbp == 0;
            else if (skipmv[0] == pre.cand[0] && skipmv[1] == pre.cand[1]) zero = pre.skip_ok & 1;
            else if (skipmv[0] == 0 && skipmv[1] == 0) zero = (pre.skip_ok >> 1) & 1;
            else { n_probe++; path = 1; zero = probe_zero(k, c, skipmv[0], skipmv[1]); pred_ready = true; /* predY / predC hold the prediction at skipmv */ }
            if (zero) {
                type = CC_MB_PSKIP; mvx = skipmv[0]; mvy = skipmv[1];
                /* The reconstruction is the prediction at the skip vector.  k.rec already holds it when the pre-pass published the
                 * prediction at its candidate (pre.skip_ok bit 2) or at the zero vector (bit 3) and that is the vector, or when the vector is the search vector
                 * of an unsplit macroblock that coded to nothing there. */
                const bool spec = (pre.skip_ok & 4) != 0, spec0 = (pre.skip_ok & 8) != 0;
                const bool in_place = spec0 ? (mvx | mvy) == 0 : (spec ? (mvx == pre.cand[0] && mvy == pre.cand[1]) : (!part && mvx == me.mv[0] && mvy == me.mv[1]));
                if (!in_place) {
                    const long long ts = k.cyc ? clock64() : 0;
                    if (!pred_ready) {
                        mc_luma16(k, c, mvx, mvy);
                        mc_chroma8(k, c, mvx, mvy);
                        __syncwarp();
                    }
                    m->recY[lane] = m->predY[lane]; m->recY[lane + 32] = m->predY[lane + 32];
                    m->recC[lane >> 4][lane & 15] = m->predC[lane >> 4][lane & 15];
                    __syncwarp();
                    publish_pixels(k, c);
                    if (k.cyc && lane == 0) {
                        unsigned long long *cy = k.cyc + (size_t)row * CC_CYC_N;
                        cy[17] += (unsigned long long)(clock64() - ts); cy[18] += 1; cy[19] += pred_ready ? 0 : 1;
                        if (!pred_ready) cy[(!part && mvx == me.mv[0] && mvy == me.mv[1]) ? 22 : ((mvx | mvy) == 0 ? 23 : 24)] += 1;
                        if (!pred_ready && (mvx | mvy) == 0 && (pre.skip_ok & 1)) cy[25] += 1;
                    }
                }
                done = true;
            } else if (part) { /* exact cost of the P_8x8 coding: per-quadrant SATD + mvd bits against the real predictors */
                cost_inter = 8 * c.lambda;
#pragma unroll
                for (int p = 0; p < 4; p++) {
                    predict_mv8(c, nbl, nbu, p, me8, mvp8[p]);
                    cost_inter += me8.satd[p] + c.lambda * (cc_se_bits(me8.mv[p][0] - mvp8[p][0]) + cc_se_bits(me8.mv[p][1] - mvp8[p][1]));
                }
            } else cost_inter = me.satd + c.lambda * (cc_se_bits(me.mv[0] - mvp[0]) + cc_se_bits(me.mv[1] - mvp[1]));
        }
        if (!done) {
            /* P frames: real intra analysis only when the pre-pass' source-based estimate is within 12.5 % of inter (oracle/ref_mb.c
             * ref_mb_encode states what the gate costs) */
            const bool gate = k.is_idr || pre.est_intra * 8 <= cost_inter * 9;
            I16Params ip;
            int cost_i16 = 1 << 30;
            if (gate) {
                n_gate++; path = 2;
                const long long ta = k.cyc ? clock64() : 0;
                stage_neighbours(k, c);
                __syncwarp();
                i16_params(c, ip);
                cost_i16 = analyse_i16(c, ip, &mode16);
                const long long tb = k.cyc ? clock64() : 0;
                /* Intra4x4 wins only if cost_i4 < cost_i16 and (P frames) cost_i4 + 8 lambda < cost_inter */
                /* the margin shrinks with the quantiser (oracle/ref_mb.c ref_mb_encode states the rule and what it costs) */
                const bool start_i4 = k.is_idr || (long long)cost_i16 * 8 <= (long long)cost_inter * (8 + (qp < 28 ? 2 : (qp < 34 ? 1 : 0)));
                if (k.cyc && lane == 0 && start_i4) k.cyc[(size_t)row * CC_CYC_N + 20] += 1;
                if (start_i4) cost_i4 = analyse_i4(k, c, k.is_idr ? cost_i16 : min(cost_i16, cost_inter - 8 * c.lambda), !k.is_idr);
                if (k.cyc && lane == 0) {
                    unsigned long long *cy = k.cyc + (size_t)row * CC_CYC_N;
                    cy[11] += (unsigned long long)(tb - ta); cy[12] += (unsigned long long)(clock64() - tb); cy[13] += cost_i4 < (1 << 30);
                }
                if (!k.is_idr) { cost_i16 += 8 * c.lambda; if (cost_i4 < (1 << 30)) cost_i4 += 8 * c.lambda; }
            }
            if (!k.is_idr && cost_inter <= cost_i16 && cost_inter <= cost_i4) type = CC_MB_P16;
            else type = cost_i16 <= cost_i4 ? CC_MB_I16 : CC_MB_I4;

            if (type == CC_MB_P16) { /* already coded by the pre-pass: levels in the slot, pixels in k.rec — or parked beside it */
                if (pre.skip_ok & 12) {
                    const uint32_t *park = (const uint32_t *)(k.park + (size_t)c.mb * 384);
                    m->recY[lane] = park[lane]; m->recY[lane + 32] = park[lane + 32]; m->recC[lane >> 4][lane & 15] = park[64 + lane];
                    __syncwarp();
                    publish_pixels(k, c);
                    if (k.cyc && lane == 0) k.cyc[(size_t)row * CC_CYC_N + 21] += 1;
                }
                if (part) { /* the four mvd go to the 8 int16 the pre-pass left free at the head of the slot */
                    mvx = me8.mv[0][0]; mvy = me8.mv[0][1]; mvdx = mvx - mvp8[0][0]; mvdy = mvy - mvp8[0][1];
                    if (lane < 8) {
                        int v = 0;
#pragma unroll
                        for (int i = 0; i < 8; i++) if (lane == i) v = me8.mv[i >> 1][i & 1] - mvp8[i >> 1][i & 1];
                        out[lane] = (int16_t)v;
                    }
                } else { mvx = me.mv[0]; mvy = me.mv[1]; mvdx = mvx - mvp[0]; mvdy = mvy - mvp[1]; }
                cbp = pre.cbp; nzmask = pre.nz_luma; off = pre.sym_len; t8 = pre.t8;
            } else {
                const long long tc = k.cyc ? clock64() : 0;
                if (type == CC_MB_I4) {
                    if (lane < 16) {
                        const uint32_t *s = (const uint32_t *)m->c4[lane];
                        uint32_t *d = (uint32_t *)m->coef[lane];
#pragma unroll
                        for (int i = 0; i < 8; i++) d[i] = s[i];
                    }
                    const int r = lane >> 1, h = lane & 1;
                    m->recY[r * 4 + h * 2] = *(const uint32_t *)&m->tile[1 + r][4 + h * 8];
                    m->recY[r * 4 + h * 2 + 1] = *(const uint32_t *)&m->tile[1 + r][8 + h * 8];
                    unsigned bal = __ballot_sync(FULL, lane < 16 && m->n4[lane & 15] > 0) & 0xffffu;
                    nzmask = bal;
#pragma unroll
                    for (int i = 0; i < 4; i++) if ((bal >> (4 * i)) & 15u) cbp |= 1 << i;
                } else cbp = encode_i16(k, c, ip, mode16, &nzmask);
                chroma_mode = intra_chroma_pred(c);
                cbp |= encode_chroma(k, c, true, false) << 4;
                __syncwarp();
                publish_pixels(k, c);
                /* symbols */
                if (type == CC_MB_I4) {
                    if (lane < 4) {
                        unsigned v = 0;
#pragma unroll
                        for (int i = 0; i < 4; i++) v |= (unsigned)m->m4code[lane * 4 + i] << (4 * i);
                        out[off + lane] = (int16_t)v;
                    }
                    off += 4;
                }
                const long long td = k.cyc ? clock64() : 0;
                off = emit_mb(c, out, off, type, cbp);
                if (k.cyc && lane == 0) {
                    unsigned long long *cy = k.cyc + (size_t)row * CC_CYC_N;
                    cy[14] += (unsigned long long)(td - tc); cy[15] += (unsigned long long)(clock64() - td);
                    if (!k.is_idr) { cy[26] += (unsigned long long)(td - tc); cy[27] += (unsigned long long)(clock64() - td); cy[28] += 1; }
                }
            }
        }
        { /* publish the macroblock's records: every value is warp-uniform, lane i stores word i of the list below in one go */
            const int p8 = type == CC_MB_P16 && part;
            const uint32_t nz16 = type == CC_MB_PSKIP ? 0u : (nzmask & 0xffffu);
            const uint32_t w_hdr = (uint32_t)type | ((uint32_t)t8 << 8) | (nz16 << 16);
            const uint32_t w_mv = ((uint32_t)mvx & 0xffffu) | ((uint32_t)mvy << 16);
            uint32_t *mbd = (uint32_t *)(k.mbd + c.mb), *nbg = k.nbrec + (size_t)c.mb * 5, *nbs = (uint32_t *)(nbl + mbx);
            if (lane < 2) {        /* cc_mbdev words 0,1: type | t8 | nz_luma, mv */
                mbd[lane] = lane ? w_mv : w_hdr;
            } else if (lane < 6) { /* cc_mbdev words 2..5: Intra4x4PredMode of the 16 blocks (DC for every other type) */
                mbd[lane] = type == CC_MB_I4 ? *(const uint32_t *)&m->m4[4 * (lane - 2)] : 0x02020202u;
            } else if (lane < 16) { /* NbRec {hdr, mv of quadrants 0..3}: global copy for k_deblock (6..10), shared-memory line (11..15) */
                const int i = lane < 11 ? lane - 6 : lane - 11;
                uint32_t v = w_hdr;
                if (i > 0) {
                    v = w_mv;
#pragma unroll
                    for (int q = 0; q < 4; q++) if (p8 && i == q + 1) v = ((uint32_t)me8.mv[q][0] & 0xffffu) | ((uint32_t)me8.mv[q][1] << 16);
                }
                (lane < 11 ? nbg : nbs)[i] = v;
            } else if (lane == 16) { /* cc_mbhost */
                const uint32_t h0 = (uint32_t)type | ((uint32_t)(mode16 | (chroma_mode << 2)) << 8) | ((uint32_t)cbp << 16) | ((uint32_t)(t8 | (p8 << 1)) << 24);
                *(uint2 *)(k.mbh + c.mb) = make_uint2(h0, ((uint32_t)mvdx & 0xffffu) | ((uint32_t)mvdy << 16));
            } else if (lane == 17) k.symlen[c.mb] = (uint16_t)off;
        }
        if (type == CC_MB_PSKIP) n_skip++; else if (type == CC_MB_P16) n_inter++; else n_intra++;
        total += off;
        __syncwarp();
        /* the band below reads what this row published (records, unfiltered lines) from global memory; so does the in-loop filter,
         * which runs next to this kernel and follows the band's bottom row (k_deblock) — for it every fourth macroblock is enough */
        if (last || (k.trail && (w == k.band - 1 || row == k.mb_h - 1) && (((mbx + 1) & 3) == 0 || mbx == mw - 1))) {
            __threadfence();
            if (lane == 0) *(volatile uint32_t *)(k.gprog + row) = tag + (uint32_t)(mbx + 1);
        } else __threadfence_block();
        if (lane == 0) progress[w] = mbx + 1;
        if (k.cyc && lane == 0) {
            const long long t2 = clock64();
            unsigned long long *cy = k.cyc + (size_t)row * CC_CYC_N;
            cy[0] += (unsigned long long)(t1 - t0); cy[1 + path] += (unsigned long long)(t2 - t1); cy[4 + path] += 1;
            cy[16] += (unsigned long long)(t0 - tprev); tprev = t2;
        }
    }
    if (lane == 0) {
        cc_rowhdr h;
        h.bytes = total * 2; h.n_skip = n_skip; h.n_inter = n_inter; h.n_intra = n_intra; h.n_gate = n_gate; h.n_probe = n_probe; h.pad = 0;
        k.rowhdr[row] = h;
    }
}

/* ---- P-frame pre-pass: every macroblock coded as P_L0_16x16 at its search vector, fully parallel ------------------------ */
/* Warps (macroblocks) per CTA.  The kernel is 114 KB of straight-line code and a warp runs through ~50 KB of it per macroblock:
 * with small CTAs the two dozen warps of an SM sit at unrelated places of it and instruction fetch, not data, is what they wait for
 * (ncu, 4 warps per CTA: stall no_instruction 2.7 per issue, more than long_scoreboard) — and the fetches crowd the wavefront
 * kernels' out of the SM's instruction caches as well.  Warps of one CTA start together and stay close: one CTA of 32 warps per SM
 * (64 registers, 214 KB of shared memory) measured +11.6 % / +13.2 % frames/s at 8 / 16 streams over 4-warp CTAs
 * (profiles/cta_size_r02.txt). */
#ifndef PC_WARPS
#define PC_WARPS 32
#endif
#ifndef PC_SYNC
#define PC_SYNC 0
#endif
#if PC_SYNC
#define PC_BAR() __syncthreads()
#else
#define PC_BAR() do { } while (0)
#endif
#ifndef PC_MINB
#define PC_MINB 1
#endif
__global__ void __launch_bounds__(PC_WARPS * 32, PC_MINB) k_p_code(cc_kparams k) {
#if PC_WARPS > 6
    CC_DYN_SMEM(pc_dsm);
    WarpMem *wm = (WarpMem *)pc_dsm;
    X8Mem *x8m = (X8Mem *)(wm + PC_WARPS);
#else
    __shared__ WarpMem wm[PC_WARPS];
    __shared__ X8Mem x8m[PC_WARPS];
#endif
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int mw = k.mb_w;
    /* a warp past the last macroblock codes the last one again (same values to the same places): the CTA's warps meet at barriers */
    const int mb = min((int)(blockIdx.x * PC_WARPS) + wid, mw * k.mb_h - 1);
    X8Mem *x8 = &x8m[wid];
    MbCtx c;
    c.lane = lane; c.mbx = mb % mw; c.mby = mb / mw; c.mb = mb; c.px = c.mbx * 16; c.py = c.mby * 16;
    c.qp = k.qp; c.qpc = c_chroma_qp[k.qp]; c.lambda = c_lambda[k.qp]; c.m = &wm[wid];
    c.L = c.mbx > 0; c.T = c.mby > (c.mby / k.rps) * k.rps; c.TR = c.T && c.mbx + 1 < mw;
    WarpMem *m = c.m;
    /* Everything the macroblock reads from global memory is requested in two rounds at the top, not where it is used (the kernel
     * used to make five dependent round trips per macroblock).  Round 1: the search result, the search results of the three
     * neighbours the P_Skip vector is predicted from (a lane each, 16x16 and 8x8 records both, no dependent second load), the
     * source pixels, the source neighbours of the intra estimate. */
    const cc_me me_r = k.me[mb];
    int nbx = 0, nby = 0; /* lanes 0..2: search vector of the neighbour's quadrant next to this macroblock (A: left, B: top, C: top-right or top-left) */
    {
        const bool useC = c.TR, useD = !c.TR && c.T && c.L;
        const int n = lane == 0 ? mb - 1 : (lane == 1 ? mb - mw : (useC ? mb - mw + 1 : mb - mw - 1));
        const int q = lane == 0 ? 1 : (lane == 1 ? 2 : (useC ? 2 : 3));
        const bool have = lane == 0 ? c.L : (lane == 1 ? c.T : (lane == 2 && (useC || useD)));
        if (have) {
            const cc_me nm = k.me[n];
            const uint32_t v8 = ((const uint32_t *)k.me8[n].mv)[q];
            nbx = nm.part ? (int)(int16_t)(v8 & 0xffffu) : nm.mv[0];
            nby = nm.part ? (int)(int16_t)(v8 >> 16) : nm.mv[1];
        }
    }
    int src_top = 0, src_left = 0; /* source neighbours: lanes 0..16 the row above (from x = -1), lanes 16..31 the column to the left */
    {
        const uint8_t *sy0 = k.cur[0] + (size_t)c.py * k.sy + c.px;
        if (lane < 17 && c.T && (lane > 0 || c.L)) src_top = sy0[-(ptrdiff_t)k.sy + lane - 1];
        if (lane >= 16 && c.L) src_left = sy0[(size_t)(lane - 16) * k.sy - 1];
    }
    x8->zz[lane] = g_zigzag8[lane]; x8->zz[lane + 32] = g_zigzag8[lane + 32];
    stage_cur(k, c);
    const int mvx = me_r.mv[0], mvy = me_r.mv[1], part = me_r.part;
    int satd4 = me_r.satd;
    /* likeliest P_Skip vector: 8.4.1.1 applied to the neighbours' search vectors */
    int cand[2], mvp[2], a[2], b[2], cc[2];
    a[0] = __shfl_sync(FULL, nbx, 0); a[1] = __shfl_sync(FULL, nby, 0);
    b[0] = __shfl_sync(FULL, nbx, 1); b[1] = __shfl_sync(FULL, nby, 1);
    cc[0] = __shfl_sync(FULL, nbx, 2); cc[1] = __shfl_sync(FULL, nby, 2);
    predict_mv_core(c.L, c.T, c.L ? 0 : -2, a, c.T ? 0 : -2, b, (c.TR || (c.T && c.L)) ? 0 : -2, cc, mvp, cand);
    /* a P_8x8 search result says nothing about the 16x16 coding at one vector: both probes run for real */
    const bool c_is_me = !part && cand[0] == mvx && cand[1] == mvy, z_is_me = !part && mvx == 0 && mvy == 0, z_is_c = cand[0] == 0 && cand[1] == 0;
    /* Round 2: the three predictions — at the search vector (predY / predC), at the P_Skip candidate (A, in the record / pixel rings
     * this kernel does not use) and at the zero vector (B, in the Intra4x4 level scratch) */
    uint32_t *const aY = (uint32_t *)m->pc, *const bY = (uint32_t *)m->c4;
    uint32_t (*const aC)[16] = (uint32_t (*)[16])(aY + 64), (*const bC)[16] = (uint32_t (*)[16])(bY + 64);
    if (part) { /* P_8x8: each quadrant predicted with its own vector (chroma: the 4x4 block it covers) */
        const cc_me8 m8 = k.me8[mb];
        mc_mb8(k, c, m8);
        satd4 = m8.satd[0] + m8.satd[1] + m8.satd[2] + m8.satd[3];
    } else {
        mc_luma16(k, c, mvx, mvy);
        mc_chroma8(k, c, mvx, mvy);
    }
    if (!c_is_me) { mc_luma16(k, c, cand[0], cand[1], aY); mc_chroma8(k, c, cand[0], cand[1], aC); }
    if (!z_is_me && !z_is_c) { /* the co-located block: aligned, no interpolation */
        const uint2 v = *(const uint2 *)(k.ref[0] + (size_t)(c.py + (lane >> 1)) * k.sy + c.px + 8 * (lane & 1));
        bY[lane * 2] = v.x; bY[lane * 2 + 1] = v.y;
        bC[lane >> 4][lane & 15] = *(const uint32_t *)(k.ref[1 + (lane >> 4)] + (size_t)(c.mby * 8 + ((lane >> 1) & 7)) * k.sc + c.mbx * 8 + 4 * (lane & 1));
    }
    if (lane < 17) m->tile[0][3 + lane] = (uint8_t)src_top;
    if (lane >= 16) m->tile[1 + lane - 16][3] = (uint8_t)src_left;
    __syncwarp();
    PC_BAR();
    unsigned nzmask;
    /* transform size [EXT: x264_mb_analyse_transform]: 8x8 when the residual's sa8d is below its 4x4 satd */
    int t8 = sa8d_mb(x8, m->curY, m->predY, lane) < satd4;
    int cbp = t8 ? inter_luma8(k, x8, m->curY, m->predY, m->recY, m->coef, c.qp, lane, &nzmask) : inter_luma(k, c, false, &nzmask);
    if (!cbp) t8 = 0; /* the flag is only transmitted with coded luma */
    PC_BAR();
    cbp |= encode_chroma(k, c, false, false) << 4;
    __syncwarp();
    PC_BAR();
    int16_t *out = (int16_t *)(k.rowbuf + (size_t)mb * CC_ROW_PAYLOAD_PER_MB);
    const unsigned off = emit_mb<1>(c, out, part ? 8 : 0, CC_MB_P16, cbp); /* P_8x8: 8 int16 left free for the four mvd */
    PC_BAR();
    /* intra gate: Intra16x16 cost estimated from SOURCE neighbours (no dependency on other macroblocks) */
    int est;
    {
        I16Params ip;
        int dummy;
        i16_params(c, ip);
        est = analyse_i16(c, ip, &dummy);
    }
    PC_BAR();
    /* P_Skip tests at the candidate and at zero: known from the coding above where the vector is the search vector, else evaluated on
     * the predictions staged at the top — both in one pass (lanes split between them) when both are needed */
    int ok = 0;
    {
        const bool needA = !c_is_me, needB = !z_is_me && !z_is_c;
        int r = 0;
        if (needA && needB) r = probe_eval2(k, c, aY, aC, bY, bC);
        else if (needA) r = probe_eval1(k, c, aY, aC) ? 1 : 0;
        else if (needB) r = probe_eval1(k, c, bY, bC) ? 2 : 0;
        ok |= c_is_me ? (cbp == 0) : (r & 1);
        ok |= z_is_me ? ((cbp == 0) << 1) : (z_is_c ? ((ok & 1) << 1) : (r & 2));
    }
    /* Most macroblocks of a moving picture end up P_Skip at the predicted skip vector, which is rarely the search vector to the
     * quarter pel: their reconstruction is the prediction at that vector.  When the probe at `cand` leaves nothing (and cand is not
     * the search vector anyway), the pre-pass therefore publishes THAT prediction and parks the P_L0_16x16 reconstruction in a
     * side buffer.  The wavefront then has nothing to move for the common outcome (P_Skip at cand) and restores the parked pixels
     * in the rare other ones.
     * Which of the two, when both probes pass: the zero vector.  8.4.1.1 forces the skip vector to zero next to a neighbour that
     * does not move, so one macroblock skipped at zero (every first macroblock of a slice row is, if at all) hands the zero vector on
     * to its right and lower neighbours for as long as their zero probe passes — at the coarse quantisers of a rate-starved stream
     * that is most of the picture (measured on the bench workload: 52 of 67 in-wavefront P_Skip reconstructions per macroblock row
     * were at the zero vector, 48 of them with the candidate's prediction published in vain). */
    PC_BAR();
    const bool pub0 = (ok & 2) && !z_is_me && !z_is_c;
    if (pub0 || ((ok & 1) && !c_is_me)) {
        ok |= pub0 ? 8 : 4;
        uint32_t *park = (uint32_t *)(k.park + (size_t)mb * 384);
        park[lane] = m->recY[lane]; park[lane + 32] = m->recY[lane + 32]; park[64 + lane] = m->recC[lane >> 4][lane & 15];
        const uint32_t *sY = pub0 ? bY : aY, *sC = pub0 ? &bC[0][0] : &aC[0][0];
        m->recY[lane] = sY[lane]; m->recY[lane + 32] = sY[lane + 32];
        m->recC[lane >> 4][lane & 15] = sC[lane];
        __syncwarp();
    }
    publish_pixels(k, c);
    if (lane == 0) {
        cc_pre p;
        p.cbp = (uint8_t)cbp; p.skip_ok = (uint8_t)ok; p.nz_luma = (uint16_t)nzmask; p.cand[0] = (int16_t)cand[0]; p.cand[1] = (int16_t)cand[1];
        p.sym_len = (uint16_t)off; p.t8 = (uint8_t)t8; p.part = (uint8_t)part; p.est_intra = est;
        k.pre[mb] = p;
    }
}

/* CTA = one slice, one warp per macroblock row (2-MB-lag wavefront).  The in-loop filter is a second kernel (k_deblock, same
 * shape) launched after k_pack: nothing in the access unit depends on it, so the host writes the bitstream while the filter and
 * the reference-finishing kernels run. */
#ifndef CC_MB_MAXNREG
#define CC_MB_MAXNREG 0
#endif
#if CC_MB_MAXNREG > 0 && !defined(CC_EMU_BUILD) /* experiments: an explicit register cap instead of the launch bound */
#define CC_MB_BOUNDS(rps) __maxnreg__(CC_MB_MAXNREG)
#else
#define CC_MB_BOUNDS(rps) __launch_bounds__((rps) * 32)
#endif
template <int RPS> /* RPS = most rows per slice this instance serves: 4 (256 threads, no 128-register cap) or MB_MAX_RPS */
__global__ void CC_MB_BOUNDS(RPS) k_mb_encode(cc_kparams k) {
    __shared__ WarpMem wm[RPS];
    __shared__ volatile int progress[RPS];
    CC_DYN_SMEM(dsm); /* band + 1 lines of mb_w neighbour records (NbRec): the row above the band, then the band's rows */
    NbRec *nbline = (NbRec *)dsm;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int row0 = blockIdx.x * k.band, row = row0 + w;
    if (lane == 0) progress[w] = 0;
    __syncthreads();
    if (row >= k.mb_h) return;
    const int srow0 = (row / k.rps) * k.rps, send = min(srow0 + k.rps, k.mb_h);
    encode_row(k, &wm[w], nbline, progress, w, row, srow0, w == k.band - 1 && row + 1 < send, lane);
}

/* In-loop filter: CTA = one band of a slice (the same bands as k_mb_encode); per macroblock row one luma warp and one chroma warp,
 * each 2 macroblocks behind the warp of the same kind in the row above (H.264 8.7 is raster-ordered; luma and chroma are
 * independent wavefronts).  A band whose slice started further up chains to the band above through frame-tagged progress words.
 *
 * The kernel runs NEXT TO k_mb_encode (k.trail; cc_enc_submit launches it behind it in the same stream with programmatic stream
 * serialisation: it begins once every CTA of k_mb_encode is resident) and follows it macroblock by macroblock: the filter of a macroblock needs the reconstruction and the records of that
 * macroblock and of those left of and above it, all of which k_mb_encode has published by the time it is done with the macroblock
 * (it keeps its own copies of the unfiltered neighbour samples — ubot / uright — and never reads k.rec back).  What tells the
 * filter how far the encoder is, is the progress word of the band's BOTTOM row (gprog[0]): when that row has finished P >= 1
 * macroblocks, the row j rows above it has finished at least P + j (the wavefront's lag), so no other word is needed — and what
 * those rows stored is visible with it: each row's warp releases its stores to the row below at CTA scope (fence + shared-memory
 * progress word, read + fence), the bottom row releases at GPU scope, and the PTX memory model's causality order is transitive
 * across those scopes (fences are cumulative).  The records come straight from global memory (past the L1, a chunk ahead where the encoder is known to be further). */
#ifndef CC_DB_MINBLOCKS
#define CC_DB_MINBLOCKS 1
#endif
#ifndef CC_DB_MINBLOCKS_TRAIL /* experiment knob: CTAs per SM the instantiation that runs next to the encoder is compiled for */
#define CC_DB_MINBLOCKS_TRAIL CC_DB_MINBLOCKS
#endif
template <bool TRAIL> /* TRAIL: launched next to k_mb_encode (k.trail) */
__global__ void __launch_bounds__(2 * MB_MAX_RPS * 32, TRAIL ? CC_DB_MINBLOCKS_TRAIL : CC_DB_MINBLOCKS) k_deblock(cc_kparams k) {
    __shared__ DbMem dbm[MB_MAX_RPS];
    __shared__ volatile int progress2[2][MB_MAX_RPS];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int pl = wid >= k.band, w = pl ? wid - k.band : wid; /* warps [0, band): luma rows, [band, 2 band): chroma rows */
    const int row0 = blockIdx.x * k.band, row = row0 + w;
    const int rows = min(k.band, k.mb_h - row0), mw = k.mb_w;
    const int srow0 = (row0 / k.rps) * k.rps, send = min(srow0 + k.rps, k.mb_h);
    if (lane == 0) progress2[pl][w] = 0;
    __syncthreads();
    if (row >= k.mb_h) return;
    volatile int *prog = progress2[pl];
    const bool T = row > srow0, xband = T && w == 0, last = w == k.band - 1 && row + 1 < send;
    const uint32_t tag = (uint32_t)k.frame_seq << 12;
    uint32_t *gp = k.gprog + (size_t)(1 + pl) * k.mb_h; /* [0]: k_mb_encode's, [1]: luma filter, [2]: chroma filter */
    const uint32_t *ep = k.gprog + row0 + rows - 1;     /* k_mb_encode's progress word of the band's bottom row */
    const int above = rows - 1 - w;                     /* this row is that many rows above it */
    const uint32_t *rq = k.nbrec + (size_t)row * mw * 5, *rt = rq - (size_t)mw * 5;
    int enc_have = TRAIL ? 0 : mw;                      /* macroblocks of this row k_mb_encode is known to have published */
    /* records by chunks of 4 macroblocks (20 words, one per lane: a single coalesced load for the row and one for the row above), the
     * next chunk on its way while this one is filtered where the encoder is known to be that far; a macroblock's record is then
     * five shuffles.  4 is also the step of the encoder's progress word. */
    uint32_t cq = 0, ct = 0, xq = 0, xt = 0;
    bool have_next = false;
    NbRec nq, nt, nl;
    for (int mbx = 0; mbx < mw; mbx++) {
        const long long t1 = k.cyc ? clock64() : 0;
        const int j = mbx & 3;
        if (j == 0) {
            const int cend = min(mbx + 4, mw);
            if (enc_have < cend) {
                const int p = wait_encoder_progress(ep, tag, max(1, cend - above), lane);
                enc_have = min(p + above, mw);
            }
            const int wi = mbx * 5 + lane; /* the row above is at least as far as this one, in the encoder and hence in its records */
            if (have_next) { cq = xq; ct = xt; }
            else {
                cq = (lane < 20 && wi < mw * 5) ? __ldcg(rq + wi) : 0u;
                ct = (T && lane < 20 && wi < mw * 5) ? __ldcg(rt + wi) : 0u;
            }
            have_next = mbx + 4 < mw && enc_have >= min(mbx + 8, mw);
            if (have_next) {
                xq = (lane < 20 && wi + 20 < mw * 5) ? __ldcg(rq + wi + 20) : 0u;
                xt = (T && lane < 20 && wi + 20 < mw * 5) ? __ldcg(rt + wi + 20) : 0u;
            }
        }
        if (xband) wait_progress_global(gp + row - 1, tag, min(mbx + 2, mw), lane);
        else if (T) wait_progress(prog + w - 1, min(mbx + 2, mw), lane);
        const long long t2 = k.cyc ? clock64() : 0;
        nq.x = __shfl_sync(FULL, cq, j * 5);
        nt.x = __shfl_sync(FULL, ct, j * 5);
#pragma unroll
        for (int i = 0; i < 4; i++) { nq.mv[i] = __shfl_sync(FULL, cq, j * 5 + 1 + i); nt.mv[i] = __shfl_sync(FULL, ct, j * 5 + 1 + i); }
        if (!T) nt = nq;
        if (mbx == 0) nl = nq;
        if (pl) deblock_mb<1>(k, &dbm[w], nq, nl, nt, lane, mbx, row, mbx > 0, T, TRAIL, xband);
        else deblock_mb<0>(k, &dbm[w], nq, nl, nt, lane, mbx, row, mbx > 0, T, TRAIL, xband);
        nl = nq;
        if (last) {
            __threadfence();
            if (lane == 0) *(volatile uint32_t *)(gp + row) = tag + (uint32_t)(mbx + 1);
        } else __threadfence_block();
        if (lane == 0) prog[w] = mbx + 1;
        if (k.cyc && lane == 0 && !pl) {
            unsigned long long *cy = k.cyc + (size_t)row * CC_CYC_N + 8;
            cy[1] += (unsigned long long)(t2 - t1); cy[2] += (unsigned long long)(clock64() - t2);
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
    /* byte offset of this row's payload = sum of the padded sizes of the rows before it: loaded in parallel, reduced per warp
     * (a serial loop would chain up to mb_h dependent global loads) */
    __shared__ unsigned part[4];
    {
        unsigned v = 0;
        for (int r = tid; r < row; r += nt) v += (k.rowhdr[r].bytes + 15u) & ~15u;
#pragma unroll
        for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
        if ((tid & 31) == 0) part[tid >> 5] = v;
    }
    __syncthreads();
    const unsigned offs = part[0] + part[1] + part[2] + part[3];
    if (hdr_bytes + mbh_bytes + offs + ((k.rowhdr[row].bytes + 15u) & ~15u) > k.out_cap) return;
    /* exclusive prefix of the row's per-MB symbol lengths (32-bit words), by warp 0 */
    __shared__ unsigned mboff[513];
    if (tid < 32) {
        unsigned run = 0;
        for (int base = 0; base < k.mb_w; base += 32) {
            const int i = base + tid;
            const unsigned len = i < k.mb_w ? (unsigned)(k.symlen[row * k.mb_w + i] >> 1) : 0u;
            unsigned inc = len;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { unsigned t = __shfl_up_sync(FULL, inc, d); if (tid >= d) inc += t; }
            if (i < k.mb_w) mboff[i] = run + inc - len;
            run += __shfl_sync(FULL, inc, 31);
        }
    }
    __syncthreads();
    uint32_t *dst = (uint32_t *)(k.out_host_dev + hdr_bytes + mbh_bytes + offs);
    const int lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
    for (int i = wid; i < k.mb_w; i += nw) {
        const unsigned n = (unsigned)(k.symlen[row * k.mb_w + i] >> 1);
        const uint32_t *src = (const uint32_t *)(k.rowbuf + (size_t)(row * k.mb_w + i) * CC_ROW_PAYLOAD_PER_MB);
        for (unsigned j = lane; j < n; j += 32) dst[mboff[i] + j] = src[j];
    }
}

extern "C" int cck_encode(const cc_kparams *k, cudaStream_t st) {
    const int ns = (k->mb_h + k->band - 1) / k->band; /* bands */
    size_t dyn = (size_t)(k->band + 1) * k->mb_w * sizeof(NbRec);
#ifndef CC_EMU_BUILD
    { /* experiment knob: pad the request so fewer foreign CTAs share the SM with a wavefront CTA (CC_WAVEFRONT_SMEM_KB) */
        static int pad_kb = -1;
        if (pad_kb < 0) { const char *v = getenv("CC_WAVEFRONT_SMEM_KB"); pad_kb = v ? atoi(v) : 0; }
        if (pad_kb > 0 && dyn < (size_t)pad_kb * 1024) dyn = (size_t)pad_kb * 1024;
        if (dyn > 190 * 1024) dyn = 190 * 1024;
    }
    /* static + dynamic shared memory passes the default 48 KB limit for wide pictures / tall slices: opt in to what this launch needs
     * (per device; cheap).  The request counts against 227 KB together with the kernel's static shared memory. */
    if (dyn > 8 * 1024) {
        cudaError_t er = k->band <= 4 ? cudaFuncSetAttribute(k_mb_encode<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn)
                                     : cudaFuncSetAttribute(k_mb_encode<MB_MAX_RPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
        (void)er; /* a failure shows as a launch error below */
    }
#endif
    if (k->band <= 4) CC_LAUNCH_SMEM(k_mb_encode<4>, dim3((unsigned)ns), dim3((unsigned)k->band * 32), dyn, st, *k);
    else CC_LAUNCH_SMEM(k_mb_encode<MB_MAX_RPS>, dim3((unsigned)ns), dim3((unsigned)k->band * 32), dyn, st, *k);
    return 1;
}
extern "C" int cck_deblock(const cc_kparams *k, cudaStream_t st, int after_trigger) {
    const int ns = (k->mb_h + k->band - 1) / k->band;
#ifndef CC_EMU_BUILD
    if (after_trigger) {
        /* programmatic stream serialisation: this launch may begin when every CTA of the kernel before it in the stream
         * (k_mb_encode) has executed griddepcontrol.launch_dependents — whatever this kernel then spins on (the encoder's progress
         * words) is produced by CTAs that are already running, so the spinning CTAs cannot keep them off the machine.
         * (The same dependency ACROSS streams — cudaLaunchAttributeProgrammaticEvent with triggerAtBlockStart — was measured to
         * release the dependent only when the kernel has ended, on this driver: tools/probes/pdl_probe.cu.) */
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof cfg);
        cfg.gridDim = dim3((unsigned)ns); cfg.blockDim = dim3((unsigned)k->band * 64); cfg.stream = st;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaError_t er = cudaLaunchKernelEx(&cfg, k_deblock<true>, *k);
        (void)er;
        return 1;
    }
#else
    (void)after_trigger;
#endif
    if (k->trail) CC_LAUNCH(k_deblock<true>, dim3((unsigned)ns), dim3((unsigned)k->band * 64), st, *k); /* the emulator's build */
    else CC_LAUNCH(k_deblock<false>, dim3((unsigned)ns), dim3((unsigned)k->band * 64), st, *k);
    return 1;
}
extern "C" int cck_p_code(const cc_kparams *k, cudaStream_t st) {
    const int nmb = k->mb_w * k->mb_h;
#if PC_WARPS > 6
    const size_t dyn = PC_WARPS * (sizeof(WarpMem) + sizeof(X8Mem)) + 16;
#ifndef CC_EMU_BUILD
    /* per device; cheap enough to repeat per launch */
    cudaFuncSetAttribute(k_p_code, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
#endif
    CC_LAUNCH_SMEM(k_p_code, dim3((unsigned)((nmb + PC_WARPS - 1) / PC_WARPS)), dim3(PC_WARPS * 32), dyn, st, *k);
#else
    CC_LAUNCH(k_p_code, dim3((unsigned)((nmb + PC_WARPS - 1) / PC_WARPS)), dim3(PC_WARPS * 32), st, *k);
#endif
    return 1;
}
extern "C" int cck_pack(const cc_kparams *k, cudaStream_t st) {
    CC_LAUNCH(k_pack, dim3((unsigned)k->mb_h + 1), dim3(128), st, *k);
    return 1;
}
