/*
 * cc_deblock.cuh — in-loop deblocking filter of one macroblock by one warp (H.264 8.7).
 *
 * The filter is normatively raster-ordered (each macroblock filters its four vertical then its four horizontal
 * edges on pixels its left / top neighbours already modified), so macroblock rows advance as a 2-MB-lag wavefront
 * exactly like the encode pass.  What this file removes is the per-edge round trip to L2: the warp stages the
 * macroblock plus the 4 pixels left of / above it into shared memory once (172 coalesced word loads), every lane
 * then filters one whole pixel line in registers (lanes 0-15: a luma row / column across its 4 edges, lanes 16-31:
 * a chroma row / column of Cb or Cr across its 2 edges), and the tile is stored back once.
 * Boundary strengths come from the neighbour records k_mb_encode left in global memory (type, nz flags, mv of current, left, top).
 * Oracle counterpart: oracle/ref_mb.c ref_mb_deblock.
 */
#ifndef CC_DEBLOCK_CUH
#define CC_DEBLOCK_CUH
#include "cc_dev.cuh"

struct __align__(16) DbMem {
    uint8_t Y[20][24];    /* Y[r + 4][c + 4] = luma sample (c, r) of the macroblock, r,c in -4..15 */
    uint8_t C[2][12][16]; /* C[pl][r + 4][c + 4], r,c in -4..7 */
};

/* neighbour record of one macroblock as the slice keeps it in shared memory (written by the encode warp, read by MV
 * prediction and by the filter): x = type | transform_size_8x8_flag << 8 | nz_luma << 16, mv[q] = vector of 8x8 quadrant q
 * (x | y << 16; four equal words unless the macroblock is P_8x8) */
struct NbRec { uint32_t x, mv[4]; };

/* boundary strength between 4x4 blocks bp (decoding order, of the record p) and bq (of q), H.264 8.7.2.1, single reference picture */
__device__ __forceinline__ int db_bs(const NbRec &p, int bp, const NbRec &q, int bq, int mb_edge) {
    const uint32_t p0 = p.x, q0 = q.x;
    if ((p0 & 255u) >= CC_MB_I16 || (q0 & 255u) >= CC_MB_I16) return mb_edge ? 4 : 3;
    if (((p0 >> (16 + bp)) | (q0 >> (16 + bq))) & 1u) return 2;
    const int ip = bp >> 2, iq = bq >> 2; /* selects, not indexing: the records live in registers */
    const uint32_t p1 = ip == 0 ? p.mv[0] : (ip == 1 ? p.mv[1] : (ip == 2 ? p.mv[2] : p.mv[3]));
    const uint32_t q1 = iq == 0 ? q.mv[0] : (iq == 1 ? q.mv[1] : (iq == 2 ? q.mv[2] : q.mv[3]));
    const int dx = (int)(int16_t)(p1 & 0xffffu) - (int)(int16_t)(q1 & 0xffffu);
    const int dy = (int)(int16_t)(p1 >> 16) - (int)(int16_t)(q1 >> 16);
    return (cc_iabs(dx) >= 4 || cc_iabs(dy) >= 4) ? 1 : 0;
}

/* raster 4x4 block index (by*4+bx) -> decoding-order index (Figure 6-10), computed (a constant-memory table with a
 * lane-dependent index would serialise) */
__device__ __forceinline__ int db_blk(int i) { return ((i >> 3) << 3) | (((i >> 1) & 1) << 2) | (((i >> 2) & 1) << 1) | (i & 1); }

/* luma edge, 8.7.2.3 (bS < 4) and 8.7.2.4 (bS == 4) */
__device__ __forceinline__ void db_luma(int p3, int &p2, int &p1, int &p0, int &q0, int &q1, int &q2, int q3, int bs, int alpha, int beta, int tc0v) {
    if (cc_iabs(p0 - q0) >= alpha || cc_iabs(p1 - p0) >= beta || cc_iabs(q1 - q0) >= beta) return;
    const int ap = cc_iabs(p2 - p0) < beta, aq = cc_iabs(q2 - q0) < beta;
    if (bs < 4) {
        const int tc = tc0v + ap + aq;
        const int d = cc_clip3(-tc, tc, (((q0 - p0) << 2) + (p1 - q1) + 4) >> 3);
        const int avg = (p0 + q0 + 1) >> 1;
        if (ap) p1 = p1 + cc_clip3(-tc0v, tc0v, (p2 + avg - (p1 << 1)) >> 1);
        if (aq) q1 = q1 + cc_clip3(-tc0v, tc0v, (q2 + avg - (q1 << 1)) >> 1);
        p0 = cc_clip255(p0 + d); q0 = cc_clip255(q0 - d);
    } else {
        const int small = cc_iabs(p0 - q0) < ((alpha >> 2) + 2);
        const int P0 = p0, P1 = p1, P2 = p2, Q0 = q0, Q1 = q1, Q2 = q2;
        if (ap && small) {
            p0 = (P2 + 2 * P1 + 2 * P0 + 2 * Q0 + Q1 + 4) >> 3;
            p1 = (P2 + P1 + P0 + Q0 + 2) >> 2;
            p2 = (2 * p3 + 3 * P2 + P1 + P0 + Q0 + 4) >> 3;
        } else p0 = (2 * P1 + P0 + Q1 + 2) >> 2;
        if (aq && small) {
            q0 = (P1 + 2 * P0 + 2 * Q0 + 2 * Q1 + Q2 + 4) >> 3;
            q1 = (P0 + Q0 + Q1 + Q2 + 2) >> 2;
            q2 = (2 * q3 + 3 * Q2 + Q1 + Q0 + P0 + 4) >> 3;
        } else q0 = (2 * Q1 + Q0 + P1 + 2) >> 2;
    }
}
/* chroma edge */
__device__ __forceinline__ void db_chroma(int p1, int &p0, int &q0, int q1, int bs, int alpha, int beta, int tc0v) {
    if (cc_iabs(p0 - q0) >= alpha || cc_iabs(p1 - p0) >= beta || cc_iabs(q1 - q0) >= beta) return;
    if (bs < 4) {
        const int tc = tc0v + 1;
        const int d = cc_clip3(-tc, tc, (((q0 - p0) << 2) + (p1 - q1) + 4) >> 3);
        p0 = cc_clip255(p0 + d); q0 = cc_clip255(q0 - d);
    } else {
        const int P0 = p0, Q0 = q0;
        p0 = (2 * p1 + P0 + q1 + 2) >> 2; q0 = (2 * q1 + Q0 + p1 + 2) >> 2;
    }
}

/* filters the staged line v[0..19] (luma: positions -4..15) across its 4 edges at positions 0,4,8,12 */
__device__ __forceinline__ void db_line_luma(int v[20], const int bs[4], int first, int alpha, int beta, const int tc0[3]) {
#pragma unroll
    for (int ed = 0; ed < 4; ed++) {
        if (ed < first || !bs[ed]) continue;
        const int b = bs[ed];
        db_luma(v[4 * ed], v[4 * ed + 1], v[4 * ed + 2], v[4 * ed + 3], v[4 * ed + 4], v[4 * ed + 5], v[4 * ed + 6], v[4 * ed + 7], b, alpha, beta,
                tc0[b < 4 ? b - 1 : 0]);
    }
}
/* chroma line v[0..11] (positions -4..7): edges at 0 and 4, strengths of the luma edges 0 and 2 */
__device__ __forceinline__ void db_line_chroma(int v[12], const int bs[4], int first, int alpha, int beta, const int tc0[3]) {
#pragma unroll
    for (int e = 0; e < 2; e++) {
        const int b = bs[2 * e];
        if (2 * e < first || !b) continue;
        db_chroma(v[4 * e + 2], v[4 * e + 3], v[4 * e + 4], v[4 * e + 5], b, alpha, beta, tc0[b < 4 ? b - 1 : 0]);
    }
}

/* One macroblock, one colour component group per warp: PL = 0 luma, PL = 1 both chroma planes (the two run in different warps
 * of the slice's CTA as independent wavefronts — in one warp their different code paths would be serialised).
 * L / T: left / top neighbour inside the slice (disable_deblocking_filter_idc = 2). */
template <int PL>
/* cg_all: all pixel loads go past the L1 (ld.cg) — the kernel is running NEXT TO k_mb_encode, which is still writing the macroblocks
 * to the right of and below this one, and a 32-byte sector fetched for this macroblock's right half also holds 16 samples of the
 * next one: cached in the L1 they would be read back stale when that macroblock's turn comes.
 * xband: the row above was filtered by another CTA (a slice taller than a band): its pixels are read past the L1 in any case. */
__device__ __forceinline__ void deblock_mb(const cc_kparams &k, DbMem *db, const NbRec &nq, const NbRec &nl, const NbRec &nt, int lane, int mbx, int mby, int L, int T, bool cg_all, bool xband) {
    const int qp = k.qp, qpc = c_chroma_qp[qp];
    const int px = mbx * 16, py = mby * 16, cx = mbx * 8, cy = mby * 8;
    if (c_alpha[qp] == 0) return; /* qp < 16: alpha = 0 for luma and chroma (qpc <= qp), the filter is the identity */
    const bool luma = PL == 0, act = lane < 16; /* lanes 16-31 only help with loads and stores */
    const int li = lane & 15;                 /* luma line, or chroma: plane = li >> 3, line = li & 7 */
    const int g = luma ? li >> 2 : (li & 7) >> 1;
    const int cpl = li >> 3, ci = li & 7;
    /* ---- boundary strengths first: most macroblocks of a P frame have nothing to filter ---- */
    /* nq, nl, nt: the records of this macroblock, its left and its top neighbour (= nq where there is none) */
    { /* The common P-frame case first (warp-uniform, ~20 instructions instead of the eight strength computations): the macroblock
       * and its left / top neighbours are inter-coded without luma coefficients and all move with one vector — every boundary
       * strength is 0 (8.7.2.1: no intra, no coefficients, equal vectors), nothing to filter, not even a pixel to load. */
        const uint32_t v = nq.mv[0];
        const bool flat = (nq.x & 0xffff00ffu) < CC_MB_I16 && nq.mv[1] == v && nq.mv[2] == v && nq.mv[3] == v &&
                          (nl.x & 0xffff00ffu) < CC_MB_I16 && nl.mv[0] == v && nl.mv[1] == v && nl.mv[2] == v && nl.mv[3] == v &&
                          (nt.x & 0xffff00ffu) < CC_MB_I16 && nt.mv[0] == v && nt.mv[1] == v && nt.mv[2] == v && nt.mv[3] == v;
        if (flat) return;
    }
    uint32_t ly[4] = {0, 0, 0, 0}, lc[3] = {0, 0, 0};
    if (luma) {
#pragma unroll
        for (int it = 0; it < 4; it++) {
            const int idx = lane + 32 * it, r = idx / 5, cw = idx - r * 5;
            const uint32_t *src = (const uint32_t *)(k.rec[0] + (ptrdiff_t)(py + r - 4) * k.sy + px - 4 + 4 * cw);
            ly[it] = idx < 100 ? ((cg_all || (xband && r < 4)) ? __ldcg(src) : *src) : 0u;
        }
    } else {
#pragma unroll
        for (int it = 0; it < 3; it++) {
            const int idx = lane + 32 * it, pl = idx >= 36, rem = idx - 36 * pl, r = rem / 3, cw = rem - r * 3;
            const uint32_t *src = (const uint32_t *)(k.rec[1 + pl] + (ptrdiff_t)(cy + r - 4) * k.sc + cx - 4 + 4 * cw);
            lc[it] = idx < 72 ? ((cg_all || (xband && r < 4)) ? __ldcg(src) : *src) : 0u;
        }
    }
    const uint32_t q0 = nq.x;
    int bsv[4], bsh[4], any = 0;
    const int blk_row = g * 4; /* raster index of the first 4x4 block of this lane's block row / column */
#pragma unroll
    for (int ed = 0; ed < 4; ed++) {
        const int pe = ed ? ed - 1 : 3;
        bsv[ed] = (ed == 0 && !L) ? 0 : db_bs(ed == 0 ? nl : nq, db_blk(blk_row + pe), nq, db_blk(blk_row + ed), ed == 0);
        bsh[ed] = (ed == 0 && !T) ? 0 : db_bs(ed == 0 ? nt : nq, db_blk(pe * 4 + g), nq, db_blk(ed * 4 + g), ed == 0);
        if ((ed & 1) && ((q0 >> 8) & 1u)) bsv[ed] = bsh[ed] = 0; /* transform_size_8x8_flag: edges 1 and 3 are not transform edges */
        if (luma || !(ed & 1)) any |= bsv[ed] | bsh[ed]; /* chroma filters on the strengths of luma edges 0 and 2 only */
    }
    if (!__any_sync(FULL, act && any)) return;
    /* ---- stage ---- */
    if (luma) {
#pragma unroll
        for (int it = 0; it < 4; it++) {
            const int idx = lane + 32 * it, r = idx / 5, cw = idx - r * 5;
            if (idx < 100) *(uint32_t *)&db->Y[r][4 * cw] = ly[it];
        }
    } else {
#pragma unroll
        for (int it = 0; it < 3; it++) {
            const int idx = lane + 32 * it, pl = idx >= 36, rem = idx - 36 * pl, r = rem / 3, cw = rem - r * 3;
            if (idx < 72) *(uint32_t *)&db->C[pl][r][4 * cw] = lc[it];
        }
    }
    __syncwarp();
    const int alpha = c_alpha[luma ? qp : qpc], beta = c_beta[luma ? qp : qpc];
    int tc0[3];
#pragma unroll
    for (int i = 0; i < 3; i++) tc0[i] = c_tc0[luma ? qp : qpc][i];

    /* ---- vertical edges: one lane = one pixel row ---- */
    if (act) {
        const int *bs = bsv;
        const int first = L ? 0 : 1;
        if (luma) {
            uint32_t w[5];
            int v[20];
#pragma unroll
            for (int i = 0; i < 5; i++) w[i] = *(const uint32_t *)&db->Y[li + 4][4 * i];
#pragma unroll
            for (int i = 0; i < 20; i++) v[i] = cc_byte(w[i >> 2], i & 3);
            db_line_luma(v, bs, first, alpha, beta, tc0);
#pragma unroll
            for (int i = 0; i < 5; i++)
                *(uint32_t *)&db->Y[li + 4][4 * i] = (uint32_t)v[4 * i] | ((uint32_t)v[4 * i + 1] << 8) | ((uint32_t)v[4 * i + 2] << 16) | ((uint32_t)v[4 * i + 3] << 24);
        } else {
            uint32_t w[3];
            int v[12];
#pragma unroll
            for (int i = 0; i < 3; i++) w[i] = *(const uint32_t *)&db->C[cpl][ci + 4][4 * i];
#pragma unroll
            for (int i = 0; i < 12; i++) v[i] = cc_byte(w[i >> 2], i & 3);
            db_line_chroma(v, bs, first, alpha, beta, tc0);
#pragma unroll
            for (int i = 0; i < 3; i++)
                *(uint32_t *)&db->C[cpl][ci + 4][4 * i] = (uint32_t)v[4 * i] | ((uint32_t)v[4 * i + 1] << 8) | ((uint32_t)v[4 * i + 2] << 16) | ((uint32_t)v[4 * i + 3] << 24);
        }
    }
    __syncwarp();
    /* ---- horizontal edges: one lane = one pixel column ---- */
    if (act) {
        const int *bs = bsh;
        const int first = T ? 0 : 1;
        if (luma) {
            int v[20];
#pragma unroll
            for (int i = 0; i < 20; i++) v[i] = db->Y[i][li + 4];
            db_line_luma(v, bs, first, alpha, beta, tc0);
#pragma unroll
            for (int i = 1; i < 20; i++) db->Y[i][li + 4] = (uint8_t)v[i];
        } else {
            int v[12];
#pragma unroll
            for (int i = 0; i < 12; i++) v[i] = db->C[cpl][i][ci + 4];
            db_line_chroma(v, bs, first, alpha, beta, tc0);
#pragma unroll
            for (int i = 2; i < 12; i++) db->C[cpl][i][ci + 4] = (uint8_t)v[i];
        }
    }
    __syncwarp();
    /* ---- store: own macroblock, the 4 columns left of it (if L), the 4 rows above it (if T) ---- */
    if (luma) {
#pragma unroll
        for (int it = 0; it < 4; it++) {
            const int idx = lane + 32 * it, r = idx / 5, cw = idx - r * 5;
            const bool ok = idx < 100 && (r >= 4 ? (cw > 0 || L) : (T && cw > 0));
            if (ok) *(uint32_t *)(k.rec[0] + (ptrdiff_t)(py + r - 4) * k.sy + px - 4 + 4 * cw) = *(const uint32_t *)&db->Y[r][4 * cw];
        }
    } else {
#pragma unroll
        for (int it = 0; it < 3; it++) {
            const int idx = lane + 32 * it, pl = idx >= 36, rem = idx - 36 * pl, r = rem / 3, cw = rem - r * 3;
            const bool ok = idx < 72 && (r >= 4 ? (cw > 0 || L) : (T && cw > 0));
            if (ok) *(uint32_t *)(k.rec[1 + pl] + (ptrdiff_t)(cy + r - 4) * k.sc + cx - 4 + 4 * cw) = *(const uint32_t *)&db->C[pl][r][4 * cw];
        }
    }
    __syncwarp();
}
#endif
