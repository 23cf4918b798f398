/*
 * oracle/ref_mb.c — per-macroblock analysis, residual coding, deblocking and CAVLC
 * macroblock_layer() writer of the CPU reference encoder.  TEST INFRASTRUCTURE ONLY.
 *
 * Follows ITU-T H.264: 8.3 (intra), 8.4.1.3 (mv prediction), 8.5 (transform),
 * 8.7 (deblocking), 7.3.5 (macroblock syntax).  Decision rules (costs, thresholds)
 * follow the published libx264 analyse forms [EXT]; the encoder slot they stand in
 * for is /root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6.
 */
#include "ref_bits.h"
#include "ref_int.h"
#include "ref_tables.h"
#include <string.h>
#include <math.h>

static const uint8_t blk_of[4][4] = {{0, 1, 4, 5}, {2, 3, 6, 7}, {8, 9, 12, 13}, {10, 11, 14, 15}};
static const uint8_t mode_bits4[4] = {1, 3, 3, 5}; /* ue() length of 0..3 */

static inline ref_mb_info *MBI(struct ref_enc *e, int x, int y) { return &e->mb[y * e->mb_w + x]; }
static inline int16_t *COEF(struct ref_enc *e, int x, int y) { return e->coef + (size_t)(y * e->mb_w + x) * COEF_PER_MB; }
static inline int iabs(int x) { return x < 0 ? -x : x; }

/* ---- residual helpers --------------------------------------------------------- */
static int fwd4x4(const uint8_t *cur, int sc, const uint8_t *pred, int sp, const uint16_t *mf,
                  const uint16_t *bias, int16_t *zz, int ac_only, int16_t *dc_out) {
    int16_t d[16];
    int n = 0;
    ref_sub4x4_dct(d, cur, sc, pred, sp);
    if (ac_only) { *dc_out = d[0]; d[0] = 0; }
    ref_quant_4x4(d, mf, bias);
    if (ac_only) { for (int i = 1; i < 16; i++) { zz[i - 1] = d[ref_zigzag4x4[i]]; n += zz[i - 1] != 0; } zz[15] = 0; }
    else for (int i = 0; i < 16; i++) { zz[i] = d[ref_zigzag4x4[i]]; n += zz[i] != 0; }
    return n;
}
/* reconstruct one 4x4: zz levels (ac_only: zz[0..14] are scan pos 1..15 and dc is a dequantised DC) */
static void rec4x4(uint8_t *dst, int sd, const uint8_t *pred, int sp, const int16_t *zz, int qp,
                   int ac_only, int dc) {
    int16_t d[16];
    if (ac_only) { d[0] = 0; for (int i = 1; i < 16; i++) d[ref_zigzag4x4[i]] = zz[i - 1]; }
    else for (int i = 0; i < 16; i++) d[ref_zigzag4x4[i]] = zz[i];
    ref_dequant_4x4(d, qp);
    if (ac_only) d[0] = (int16_t)dc;
    ref_add4x4_idct(dst, sd, pred, sp, d);
}
static void copy_blk(uint8_t *dst, int sd, const uint8_t *src, int ss, int w, int h) {
    for (int y = 0; y < h; y++) memcpy(dst + y * sd, src + y * ss, (size_t)w);
}

/* ---- availability (slices are bands of whole MB rows) ------------------------- */
static inline int av_left(struct ref_enc *e, int x, int y) { (void)e; (void)y; return x > 0; }
static inline int av_top(struct ref_enc *e, int x, int y) { (void)x; return y > ref_slice_first_row(e, y); }
static inline int av_topright(struct ref_enc *e, int x, int y) { return av_top(e, x, y) && x + 1 < e->mb_w; }

/* ---- intra 4x4 analysis + coding (sequential, writes rec luma) ---------------- */
/* proj_thresh > 0 (P frames): the analysis is given up — returns 1 << 30 — when, at one of the product's ten wavefront steps
 * (blocks with equal bx + 2 by are analysed together: 1, 2, 4, 6, ... 16 blocks done), at least four blocks are done and the cost
 * so far, projected to sixteen blocks, exceeds the cost Intra4x4 has to beat by a quarter.  Measured on S2 1080p: 1 % of the
 * analyses started in P frames end in an Intra4x4 macroblock; the projection halves the steps taken and loses about one such
 * macroblock in ten, 0.01 - 0.1 % of the analysed macroblocks' cost (the analysis itself is unchanged: the same block costs). */
static int analyse_i4(struct ref_enc *e, int mbx, int mby, int16_t coef[16][16], uint8_t modes[16], uint8_t nnz[16], int proj_thresh) {
    const int qp = e->qp, lambda = ref_lambda[qp], sy = e->sy;
    const int L = av_left(e, mbx, mby), T = av_top(e, mbx, mby), TR = av_topright(e, mbx, mby);
    uint16_t mf[16], bias[16];
    int cost = 24 * lambda;
    ref_quant_tables(qp, 1, mf, bias);
    for (int blk = 0; blk < 16; blk++) {
        int bx = ref_blk_x[blk], by = ref_blk_y[blk];
        uint8_t *rec = e->rec[0] + (mby * 16 + by * 4) * sy + mbx * 16 + bx * 4;
        const uint8_t *cur = e->cur[0] + (mby * 16 + by * 4) * sy + mbx * 16 + bx * 4;
        int has_top = by > 0 || T, has_left = bx > 0 || L, has_tr;
        uint8_t tb[9] = {0}, left[4] = {0}, *top = tb + 1;
        if (by == 0) has_tr = (bx < 3) ? T : TR;
        else has_tr = !(bx == 3 || blk == 3 || blk == 11 || blk == 7 || blk == 13 || blk == 15);
        if (has_top) {
            for (int i = 0; i < 4; i++) top[i] = rec[-sy + i];
            for (int i = 4; i < 8; i++) top[i] = has_tr ? rec[-sy + i] : top[3];
        }
        if (has_left) for (int i = 0; i < 4; i++) left[i] = rec[i * sy - 1];
        if (has_top && has_left) top[-1] = rec[-sy - 1];
        /* predicted mode, H.264 8.3.1.1 */
        int mA, mB;
        if (bx > 0) mA = modes[blk_of[by][bx - 1]];
        else if (L) { ref_mb_info *n = MBI(e, mbx - 1, mby); mA = n->type == REF_MB_I4 ? n->i4_mode[blk_of[by][3]] : 2; }
        else mA = -1;
        if (by > 0) mB = modes[blk_of[by - 1][bx]];
        else if (T) { ref_mb_info *n = MBI(e, mbx, mby - 1); mB = n->type == REF_MB_I4 ? n->i4_mode[blk_of[3][bx]] : 2; }
        else mB = -1;
        int pm = (mA < 0 || mB < 0) ? 2 : (mA < mB ? mA : mB);
        int best = 1 << 30, bm = 2;
        uint8_t pred[16], bpred[16];
        for (int m = 0; m < 9; m++) {
            int ok = (m == 2) || ((m == 0 || m == 3 || m == 7) && has_top) || ((m == 1 || m == 8) && has_left) ||
                     ((m == 4 || m == 5 || m == 6) && has_top && has_left);
            if (!ok) continue;
            ref_pred4x4(pred, m, top, left, has_top, has_left);
            int c = ref_satd_4x4(cur, sy, pred, 4) + lambda * (m == pm ? 1 : 4);
            if (c < best) { best = c; bm = m; memcpy(bpred, pred, 16); }
        }
        cost += best;
        MBI(e, mbx, mby)->dbg_i4_blk[blk] = (uint16_t)(best > 65535 ? 65535 : best);
        MBI(e, mbx, mby)->dbg_i4_blk_exact[blk] = best;
        modes[blk] = (uint8_t)bm;
        nnz[blk] = (uint8_t)fwd4x4(cur, sy, bpred, 4, mf, bias, coef[blk], 0, 0);
        if (nnz[blk]) rec4x4(rec, sy, bpred, 4, coef[blk], qp, 0, 0);
        else copy_blk(rec, sy, bpred, 4, 4, 4);
    }
    if (proj_thresh > 0) {
        int part = 0, nd = 0;
        for (int t = 0; t < 9; t++) {
            for (int by = 0; by < 4; by++) {
                const int bx = t - 2 * by;
                if (bx >= 0 && bx < 4) { part += MBI(e, mbx, mby)->dbg_i4_blk_exact[blk_of[by][bx]]; nd++; }
            }
            if (24 * lambda + part >= proj_thresh) break; /* the exact test stops the product first; the outcome is "not chosen" either way */
            if (nd >= 4 && (long)part * 64 > (long)(proj_thresh - 24 * lambda) * nd * 5) return 1 << 30;
        }
    }
    return cost;
}

/* nb: the plane the neighbour samples come from — the reconstruction (real analysis) or the source picture
 * (the P-frame intra gate: a neighbour-independent estimate, so it can be computed for all MBs in parallel) */
static int analyse_i16(struct ref_enc *e, int mbx, int mby, int *mode_out, const uint8_t *nb) {
    const int sy = e->sy, lambda = ref_lambda[e->qp];
    const int L = av_left(e, mbx, mby), T = av_top(e, mbx, mby);
    const uint8_t *rec = nb + mby * 16 * sy + mbx * 16, *cur = e->cur[0] + mby * 16 * sy + mbx * 16;
    uint8_t pred[256];
    int best = 1 << 30;
    for (int m = 0; m < 4; m++) {
        int ok = m == 2 || (m == 0 && T) || (m == 1 && L) || (m == 3 && T && L);
        if (!ok) continue;
        ref_pred16x16(pred, 16, m, rec, sy, T, L);
        int c = ref_satd(cur, sy, pred, 16, 16, 16) + lambda * mode_bits4[m];
        if (c < best) { best = c; *mode_out = m; }
    }
    return best;
}

static void encode_i16(struct ref_enc *e, int mbx, int mby, int mode, ref_mb_info *mi, int16_t *coef) {
    const int sy = e->sy, qp = e->qp;
    const int L = av_left(e, mbx, mby), T = av_top(e, mbx, mby);
    uint8_t *rec = e->rec[0] + mby * 16 * sy + mbx * 16;
    const uint8_t *cur = e->cur[0] + mby * 16 * sy + mbx * 16;
    uint8_t pred[256];
    uint16_t mf[16], bias[16];
    int16_t dc[16], dcq[16];
    int any_ac = 0;
    ref_quant_tables(qp, 1, mf, bias);
    ref_pred16x16(pred, 16, mode, rec, sy, T, L);
    for (int blk = 0; blk < 16; blk++) {
        int bx = ref_blk_x[blk], by = ref_blk_y[blk];
        mi->nnz[blk] = (uint8_t)fwd4x4(cur + by * 4 * sy + bx * 4, sy, pred + by * 64 + bx * 4, 16, mf, bias,
                                       coef + blk * 16, 1, &dc[by * 4 + bx]);
        any_ac |= mi->nnz[blk];
    }
    ref_dct4x4dc(dc);
    ref_quant_dc(dc, 16, mf[0] >> 1, bias[0] << 1);
    mi->nnz[24] = 0;
    for (int i = 0; i < 16; i++) { coef[24 * 16 + i] = dc[ref_zigzag4x4[i]]; mi->nnz[24] += dc[ref_zigzag4x4[i]] != 0; }
    memcpy(dcq, dc, sizeof dc);
    ref_idct4x4dc(dcq);
    ref_dequant_4x4_dc(dcq, qp);
    if (!any_ac) for (int blk = 0; blk < 16; blk++) { mi->nnz[blk] = 0; }
    for (int blk = 0; blk < 16; blk++) {
        int bx = ref_blk_x[blk], by = ref_blk_y[blk];
        rec4x4(rec + by * 4 * sy + bx * 4, sy, pred + by * 64 + bx * 4, 16, coef + blk * 16, qp, 1, dcq[by * 4 + bx]);
    }
    mi->cbp = (uint8_t)((mi->cbp & 0x30) | (any_ac ? 15 : 0));
    mi->i16_mode = (uint8_t)mode;
}

/* 8x8-transform variant (High profile, transform_size_8x8_flag = 1).  CAVLC carries an 8x8 block as four interleaved
 * 4x4 blocks: block i of the 8x8 gets scan positions 4k+i (H.264 7.3.5.3.2 / 8.5.6), so coef/nnz keep the 4x4 layout. */
static void encode_inter_luma8(struct ref_enc *e, int mbx, int mby, const uint8_t *pred, ref_mb_info *mi, int16_t *coef) {
    const int sy = e->sy, qp = e->qp;
    uint8_t *rec = e->rec[0] + mby * 16 * sy + mbx * 16;
    const uint8_t *cur = e->cur[0] + mby * 16 * sy + mbx * 16;
    int16_t d[4][64];
    int score8[4], nz8[4], total = 0, cbp = 0;
    for (int i8 = 0; i8 < 4; i8++) {
        const int ox = (i8 & 1) * 8, oy = (i8 >> 1) * 8;
        int16_t zz[64];
        ref_sub8x8_dct8(d[i8], cur + oy * sy + ox, sy, pred + oy * 16 + ox, 16);
        nz8[i8] = ref_quant_8x8(d[i8], qp, 0);
        for (int k = 0; k < 64; k++) zz[k] = d[i8][ref_zigzag8x8[k]];
        score8[i8] = ref_decimate_score64(zz);
        total += score8[i8];
        for (int k = 0; k < 64; k++) coef[(i8 * 4 + (k & 3)) * 16 + (k >> 2)] = zz[k];
    }
    for (int i8 = 0; i8 < 4; i8++) { /* [EXT] x264 luma decimation, same thresholds as the 4x4 path */
        const int ox = (i8 & 1) * 8, oy = (i8 >> 1) * 8;
        if (total < 6 || score8[i8] < 4) nz8[i8] = 0;
        if (nz8[i8]) {
            cbp |= 1 << i8;
            ref_dequant_8x8(d[i8], qp);
            ref_add8x8_idct8(rec + oy * sy + ox, sy, pred + oy * 16 + ox, 16, d[i8]);
        } else {
            memset(coef + i8 * 64, 0, 128);
            copy_blk(rec + oy * sy + ox, sy, pred + oy * 16 + ox, 16, 8, 8);
        }
        for (int i4 = 0; i4 < 4; i4++) {
            int n = 0;
            for (int k = 0; k < 16; k++) n += coef[(i8 * 4 + i4) * 16 + k] != 0;
            mi->nnz[i8 * 4 + i4] = (uint8_t)n;
        }
    }
    mi->cbp = (uint8_t)((mi->cbp & 0x30) | cbp);
}

static void encode_inter_luma(struct ref_enc *e, int mbx, int mby, const uint8_t *pred, ref_mb_info *mi, int16_t *coef) {
    const int sy = e->sy, qp = e->qp;
    uint8_t *rec = e->rec[0] + mby * 16 * sy + mbx * 16;
    const uint8_t *cur = e->cur[0] + mby * 16 * sy + mbx * 16;
    uint16_t mf[16], bias[16];
    int score8[4], total = 0, cbp = 0;
    ref_quant_tables(qp, 0, mf, bias);
    for (int i8 = 0; i8 < 4; i8++) {
        score8[i8] = 0;
        for (int i4 = 0; i4 < 4; i4++) {
            int blk = i8 * 4 + i4, bx = ref_blk_x[blk], by = ref_blk_y[blk];
            mi->nnz[blk] = (uint8_t)fwd4x4(cur + by * 4 * sy + bx * 4, sy, pred + by * 64 + bx * 4, 16, mf, bias, coef + blk * 16, 0, 0);
            score8[i8] += ref_decimate_score(coef + blk * 16, 16);
        }
        total += score8[i8];
    }
    for (int i8 = 0; i8 < 4; i8++) { /* [EXT] x264 luma decimation: 8x8 score < 4, MB score < 6 */
        if (total < 6 || score8[i8] < 4)
            for (int i4 = 0; i4 < 4; i4++) { mi->nnz[i8 * 4 + i4] = 0; memset(coef + (i8 * 4 + i4) * 16, 0, 32); }
        for (int i4 = 0; i4 < 4; i4++) {
            int blk = i8 * 4 + i4, bx = ref_blk_x[blk], by = ref_blk_y[blk];
            if (mi->nnz[blk]) { cbp |= 1 << i8; rec4x4(rec + by * 4 * sy + bx * 4, sy, pred + by * 64 + bx * 4, 16, coef + blk * 16, qp, 0, 0); }
            else copy_blk(rec + by * 4 * sy + bx * 4, sy, pred + by * 64 + bx * 4, 16, 4, 4);
        }
    }
    mi->cbp = (uint8_t)((mi->cbp & 0x30) | cbp);
}

/* chroma residual for both planes; pred[2][64] */
static void encode_chroma(struct ref_enc *e, int mbx, int mby, uint8_t pred[2][64], int intra, ref_mb_info *mi, int16_t *coef) {
    const int sc = e->sc, qpc = ref_chroma_qp[e->qp];
    uint16_t mf[16], bias[16];
    int any_dc = 0, any_ac = 0;
    int16_t dcdq[2][4];
    ref_quant_tables(qpc, intra, mf, bias);
    for (int p = 0; p < 2; p++) {
        const uint8_t *cur = e->cur[1 + p] + mby * 8 * sc + mbx * 8;
        int16_t dc[4];
        int score = 0, nac = 0;
        for (int b = 0; b < 4; b++) {
            int bx = b & 1, by = b >> 1, blk = 16 + p * 4 + b;
            mi->nnz[blk] = (uint8_t)fwd4x4(cur + by * 4 * sc + bx * 4, sc, pred[p] + by * 32 + bx * 4, 8, mf, bias, coef + blk * 16, 1, &dc[b]);
            score += ref_decimate_score(coef + blk * 16, 15);
            nac |= mi->nnz[blk];
        }
        if (!intra && score < 7 && nac) { /* [EXT] x264 chroma AC decimation, inter only */
            for (int b = 0; b < 4; b++) { mi->nnz[16 + p * 4 + b] = 0; memset(coef + (16 + p * 4 + b) * 16, 0, 32); }
            nac = 0;
        }
        any_ac |= nac;
        ref_dct2x2dc(dc);
        ref_quant_dc(dc, 4, mf[0] >> 1, bias[0] << 1);
        mi->nnz[25 + p] = 0;
        for (int i = 0; i < 4; i++) { coef[(25 + p) * 16 + i] = dc[i]; mi->nnz[25 + p] += dc[i] != 0; }
        any_dc |= mi->nnz[25 + p];
        ref_dequant_2x2_dc(dcdq[p], dc, qpc);
    }
    for (int p = 0; p < 2; p++) {
        uint8_t *rec = e->rec[1 + p] + mby * 8 * sc + mbx * 8;
        for (int b = 0; b < 4; b++) {
            int bx = b & 1, by = b >> 1, blk = 16 + p * 4 + b;
            rec4x4(rec + by * 4 * sc + bx * 4, sc, pred[p] + by * 32 + bx * 4, 8, coef + blk * 16, qpc, 1, dcdq[p][b]);
        }
    }
    mi->cbp = (uint8_t)((mi->cbp & 15) | ((any_ac ? 2 : (any_dc ? 1 : 0)) << 4));
}

/* ---- motion vector prediction, H.264 8.4.1.3 (16x16 or four 8x8 partitions per MB, one reference) */
/* neighbour = 8x8 quadrant `part` of MB (x,y): ref -2 unavailable, -1 intra, 0 inter */
static void nb_mv(struct ref_enc *e, int x, int y, int avail, int part, int *ref, int mv[2]) {
    mv[0] = mv[1] = 0;
    if (!avail) { *ref = -2; return; }
    ref_mb_info *n = MBI(e, x, y);
    if (n->type >= REF_MB_I16) { *ref = -1; return; }
    *ref = 0; mv[0] = n->mv4[part][0]; mv[1] = n->mv4[part][1];
}
static void median_pred(int rA, const int *a, int rB, const int *b0, int rC, const int *c0, int mvp[2]) {
    int b[2] = {b0[0], b0[1]}, c[2] = {c0[0], c0[1]};
    if (rB == -2 && rC == -2 && rA != -2) { rB = rC = rA; b[0] = c[0] = a[0]; b[1] = c[1] = a[1]; }
    int n = (rA == 0) + (rB == 0) + (rC == 0);
    if (n == 1) {
        const int *s = rA == 0 ? a : (rB == 0 ? b : c);
        mvp[0] = s[0]; mvp[1] = s[1];
    } else { mvp[0] = ref_median3(a[0], b[0], c[0]); mvp[1] = ref_median3(a[1], b[1], c[1]); }
}
/* 16x16 partition: A = left MB quadrant 1, B = top MB quadrant 2, C = top-right MB quadrant 2 (else D = top-left MB
 * quadrant 3) — the quadrants that contain the neighbouring luma locations of 6.4.11.7 */
static void predict_mv(struct ref_enc *e, int mbx, int mby, int mvp[2], int skip[2]) {
    int rA, rB, rC, a[2], b[2], c[2];
    int L = av_left(e, mbx, mby), T = av_top(e, mbx, mby);
    nb_mv(e, mbx - 1, mby, L, 1, &rA, a);
    nb_mv(e, mbx, mby - 1, T, 2, &rB, b);
    nb_mv(e, mbx + 1, mby - 1, av_topright(e, mbx, mby), 2, &rC, c);
    if (rC == -2) nb_mv(e, mbx - 1, mby - 1, T && L, 3, &rC, c);
    median_pred(rA, a, rB, b, rC, c, mvp);
    /* P_Skip, 8.4.1.1 */
    if (!L || !T || (rA == 0 && a[0] == 0 && a[1] == 0) || (rB == 0 && b[0] == 0 && b[1] == 0)) skip[0] = skip[1] = 0;
    else { skip[0] = mvp[0]; skip[1] = mvp[1]; }
}
/* 8x8 quadrant p of a P_8x8 MB; mv4 = vectors of the quadrants already decided (decoding order 0..3) */
static void predict_mv8(struct ref_enc *e, int mbx, int mby, int p, int16_t (*mv4)[2], int *mvp) {
    int rA, rB, rC, a[2], b[2], c[2];
    int L = av_left(e, mbx, mby), T = av_top(e, mbx, mby);
#define OWN(q, r, v) do { r = 0; v[0] = mv4[q][0]; v[1] = mv4[q][1]; } while (0)
    if (p == 0) {
        nb_mv(e, mbx - 1, mby, L, 1, &rA, a);
        nb_mv(e, mbx, mby - 1, T, 2, &rB, b);
        nb_mv(e, mbx, mby - 1, T, 3, &rC, c);
        if (rC == -2) nb_mv(e, mbx - 1, mby - 1, T && L, 3, &rC, c);
    } else if (p == 1) {
        OWN(0, rA, a);
        nb_mv(e, mbx, mby - 1, T, 3, &rB, b);
        nb_mv(e, mbx + 1, mby - 1, av_topright(e, mbx, mby), 2, &rC, c);
        if (rC == -2) nb_mv(e, mbx, mby - 1, T, 2, &rC, c);
    } else if (p == 2) {
        nb_mv(e, mbx - 1, mby, L, 3, &rA, a);
        OWN(0, rB, b);
        OWN(1, rC, c);
    } else {
        OWN(2, rA, a);
        OWN(1, rB, b);
        OWN(0, rC, c); /* C (the MB to the right) is not yet decoded -> D = quadrant 0 */
    }
#undef OWN
    median_pred(rA, a, rB, b, rC, c, mvp);
}
static void set_mv_all(ref_mb_info *mi, int mx, int my) {
    for (int p = 0; p < 4; p++) { mi->mv4[p][0] = (int16_t)mx; mi->mv4[p][1] = (int16_t)my; }
    mi->mv[0] = (int16_t)mx; mi->mv[1] = (int16_t)my;
}

/* ---- macroblock analysis + encode ---------------------------------------------- */
/* the MB as P_L0_16x16 at its search vector: luma with the transform size chosen by
 * [EXT] x264_mb_analyse_transform (sa8d of the residual < satd of the residual -> 8x8), then chroma */
static void code_inter_at_me(struct ref_enc *e, int mbx, int mby, ref_mb_info *mi, int16_t *coef, int force_t8) {
    const int sy = e->sy, sc = e->sc;
    uint8_t predY[256], predC[2][64];
    const uint8_t *cur = e->cur[0] + mby * 16 * sy + mbx * 16;
    memset(coef, 0, COEF_PER_MB * sizeof(int16_t));
    memset(mi->nnz, 0, sizeof mi->nnz);
    mi->cbp = 0;
    mi->part = mi->me_part;
    int satd4 = mi->me_satd;
    if (mi->part) { /* P_8x8: each quadrant predicted with its own vector (chroma: the 4x4 block it covers) */
        satd4 = 0;
        for (int p = 0; p < 4; p++) {
            const int ox = (p & 1) * 8, oy = (p >> 1) * 8;
            mi->mv4[p][0] = mi->me_mv8[p][0]; mi->mv4[p][1] = mi->me_mv8[p][1];
            ref_mc_luma(predY + oy * 16 + ox, 16, e->ref[0] + (mby * 16 + oy) * sy + mbx * 16 + ox, sy, mi->mv4[p][0], mi->mv4[p][1], 8, 8);
            for (int pl = 0; pl < 2; pl++)
                ref_mc_chroma(predC[pl] + (oy / 2) * 8 + ox / 2, 8, e->ref[1 + pl] + (mby * 8 + oy / 2) * sc + mbx * 8 + ox / 2, sc,
                              mi->mv4[p][0], mi->mv4[p][1], 4, 4);
            satd4 += mi->me_satd8[p];
        }
        mi->mv[0] = mi->mv4[0][0]; mi->mv[1] = mi->mv4[0][1];
    } else {
        set_mv_all(mi, mi->me_mv[0], mi->me_mv[1]);
        ref_mc_luma(predY, 16, e->ref[0] + mby * 16 * sy + mbx * 16, sy, mi->mv[0], mi->mv[1], 16, 16);
        for (int pl = 0; pl < 2; pl++)
            ref_mc_chroma(predC[pl], 8, e->ref[1 + pl] + mby * 8 * sc + mbx * 8, sc, mi->mv[0], mi->mv[1], 8, 8);
    }
    int s8 = 0;
    for (int i8 = 0; i8 < 4; i8++) s8 += ref_sa8d_8x8_raw(cur + (i8 >> 1) * 8 * sy + (i8 & 1) * 8, sy, predY + (i8 >> 1) * 128 + (i8 & 1) * 8, 16);
    mi->t8 = (uint8_t)(((s8 + 2) >> 2) < satd4);
    if (force_t8 >= 0) mi->t8 = (uint8_t)force_t8; /* upper-bound mode only (ref_mb_encode_rd) */
    if (mi->t8) encode_inter_luma8(e, mbx, mby, predY, mi, coef);
    else encode_inter_luma(e, mbx, mby, predY, mi, coef);
    if (!(mi->cbp & 15)) mi->t8 = 0; /* the flag is only transmitted with coded luma */
    encode_chroma(e, mbx, mby, predC, 0, mi, coef);
}

/* chroma of an intra macroblock: mode by SATD over both planes, then the residual */
static void intra_chroma(struct ref_enc *e, int mbx, int mby, ref_mb_info *mi, int16_t *coef) {
    const int sc = e->sc, lambda = ref_lambda[e->qp], L = av_left(e, mbx, mby), T = av_top(e, mbx, mby);
    uint8_t predC[2][64], p[2][64];
    int best = 1 << 30, bm = 0;
    for (int m = 0; m < 4; m++) {
        int ok = m == 0 || (m == 1 && L) || (m == 2 && T) || (m == 3 && T && L), c = lambda * mode_bits4[m];
        if (!ok) continue;
        for (int pl = 0; pl < 2; pl++) {
            ref_pred8x8c(p[pl], 8, m, e->rec[1 + pl] + mby * 8 * sc + mbx * 8, sc, T, L);
            c += ref_satd(e->cur[1 + pl] + mby * 8 * sc + mbx * 8, sc, p[pl], 8, 8, 8);
        }
        if (c < best) { best = c; bm = m; memcpy(predC, p, sizeof p); }
    }
    mi->chroma_mode = (uint8_t)bm;
    encode_chroma(e, mbx, mby, predC, 1, mi, coef);
}

/* decision-rule constants {numerator, denominator}; ref_x_set_rules() overrides them for what-if measurements (tools/rule_eval.py) */
static int g_x_gate[2] = {9, 8};  /* intra gate: source-based estimate within 12.5 % of the inter cost */
static int g_x_i4[2] = {0, 0};    /* Intra4x4 analysis: real Intra16x16 cost within a margin of the inter cost; {0, 0} = the product's rule (below) */
static int g_x_nocbp = 0;
void ref_x_set_nocbp(int on) { g_x_nocbp = on; }
void ref_x_set_rules(int gate_num, int gate_den, int i4_num, int i4_den) { g_x_gate[0] = gate_num; g_x_gate[1] = gate_den; g_x_i4[0] = i4_num; g_x_i4[1] = i4_den; }
void ref_mb_encode(struct ref_enc *e, int mbx, int mby) {
    const int sy = e->sy, sc = e->sc, lambda = ref_lambda[e->qp];
    ref_mb_info *mi = MBI(e, mbx, mby);
    int16_t *coef = COEF(e, mbx, mby);
    uint8_t predY[256], predC[2][64];
    int mvp[2] = {0, 0}, skipmv[2] = {0, 0};
    int cost_inter = 1 << 30, cost_i16, cost_i4 = 1 << 30, mode16 = 2;
    int16_t c4[16][16];
    uint8_t m4[16], n4[16];

    memset(coef, 0, COEF_PER_MB * sizeof(int16_t));
    memset(mi->nnz, 0, sizeof mi->nnz);
    memset(mi->i4_mode, 2, 16);
    mi->cbp = 0; mi->mvd[0] = mi->mvd[1] = 0; mi->i16_mode = 0; mi->chroma_mode = 0; mi->part = 0;
    set_mv_all(mi, 0, 0);
    memset(mi->mvd4, 0, sizeof mi->mvd4);

    mi->t8 = 0;
    mi->dbg_cost_inter = mi->dbg_cost_i16 = mi->dbg_cost_i4 = 1 << 30;
    if (!e->is_idr) {
        predict_mv(e, mbx, mby, mvp, skipmv);
        /* [EXT] x264 probe_skip: would the MB coded at the P_Skip vector leave nothing after decimation?  When the skip
         * vector IS the search vector the answer is the real coding at that vector (with its transform-size choice),
         * otherwise a 4x4-transform probe. */
        if (!mi->me_part && skipmv[0] == mi->me_mv[0] && skipmv[1] == mi->me_mv[1]) {
            code_inter_at_me(e, mbx, mby, mi, coef, -1);
        } else {
            ref_mc_luma(predY, 16, e->ref[0] + mby * 16 * sy + mbx * 16, sy, skipmv[0], skipmv[1], 16, 16);
            encode_inter_luma(e, mbx, mby, predY, mi, coef);
            if ((mi->cbp & 15) == 0) {
                for (int pl = 0; pl < 2; pl++)
                    ref_mc_chroma(predC[pl], 8, e->ref[1 + pl] + mby * 8 * sc + mbx * 8, sc, skipmv[0], skipmv[1], 8, 8);
                encode_chroma(e, mbx, mby, predC, 0, mi, coef);
            } else mi->cbp |= 0x10; /* not a skip; chroma not examined */
        }
        if (mi->cbp == 0) {
            mi->type = REF_MB_PSKIP; set_mv_all(mi, skipmv[0], skipmv[1]); mi->t8 = 0; mi->part = 0;
            return;
        }
        memset(coef, 0, COEF_PER_MB * sizeof(int16_t));
        memset(mi->nnz, 0, sizeof mi->nnz);
        mi->cbp = 0; mi->t8 = 0; mi->part = 0;
        if (mi->me_part) { /* exact cost of the P_8x8 coding: per-quadrant SATD + mvd bits against the real predictors */
            cost_inter = 8 * lambda;
            for (int p = 0; p < 4; p++) {
                int mp[2];
                predict_mv8(e, mbx, mby, p, mi->me_mv8, mp);
                cost_inter += mi->me_satd8[p] + lambda * (ref_mv_bits(mi->me_mv8[p][0] - mp[0]) + ref_mv_bits(mi->me_mv8[p][1] - mp[1]));
            }
        } else
        cost_inter = mi->me_satd + lambda * (ref_mv_bits(mi->me_mv[0] - mvp[0]) + ref_mv_bits(mi->me_mv[1] - mvp[1]));
    }
    /* P frames: intra is analysed for real only when a source-based Intra16x16 estimate comes within 12.5 % of the inter
     * cost [EXT: x264 makes the same kind of early-out in x264_mb_analyse_intra / fast_intra, at 25 %].  Measured on S2 1080p
     * (the oracle over 7 P frames at QP 24 / 30 / 36): against 25 % the tighter gate starts 77 / 50 / 15 % fewer analyses and
     * loses 6 - 8 intra macroblocks per frame, 0.02 - 0.06 % of the coded macroblocks' cost. */
    int gate = 1;
    if (!e->is_idr) {
        int dummy;
        const int est = analyse_i16(e, mbx, mby, &dummy, e->cur[0]);
        mi->dbg_est_intra = est; mi->dbg_cost_inter = cost_inter;
        gate = est * g_x_gate[1] <= cost_inter * g_x_gate[0];
        if (gate && g_x_nocbp) { /* what-if: no intra analysis for a macroblock whose inter coding leaves no coefficient */
            ref_mb_info keep = *mi;
            int16_t keepc[COEF_PER_MB];
            memcpy(keepc, coef, sizeof keepc);
            code_inter_at_me(e, mbx, mby, mi, coef, -1);
            const int cbp0 = mi->cbp;
            *mi = keep; memcpy(coef, keepc, sizeof keepc);
            if (g_x_nocbp == 1 ? cbp0 == 0 : (cbp0 & 15) == 0) gate = 0;
        }
    }
    cost_i16 = 1 << 30;
    if (gate) {
        cost_i16 = analyse_i16(e, mbx, mby, &mode16, e->rec[0]);
        /* P frames: the Intra4x4 analysis starts only when the real Intra16x16 cost is within a margin of the inter cost, and the
         * margin shrinks with the quantiser — 25 % below QP 28, 12.5 % below QP 34, none from there on: at coarse quantisers one
         * analysis in a hundred ends in an Intra4x4 macroblock (QP 32: 0.1 % of the macroblocks, QP 38: 0.03 %), yet 85 - 90 % of the
         * gated macroblocks used to start it.  Measured with tools/rule_eval.py on S2 1080p: +0.10 % BD-rate against a flat 25 %
         * (only QP 34 moves, +0.26 % bits), for a third fewer started analyses at QP 28 - 33 and 55 - 80 % fewer from QP 34. */
        const int i4m = e->qp < 28 ? 2 : (e->qp < 34 ? 1 : 0); /* margin in eighths */
        if (e->is_idr || (g_x_i4[1] ? (long long)cost_i16 * g_x_i4[1] <= (long long)cost_inter * g_x_i4[0] : (long long)cost_i16 * 8 <= (long long)cost_inter * (8 + i4m)))
            cost_i4 = analyse_i4(e, mbx, mby, c4, m4, n4, e->is_idr ? 0 : (cost_i16 < cost_inter - 8 * lambda ? cost_i16 : cost_inter - 8 * lambda));
        if (!e->is_idr) { cost_i16 += 8 * lambda; if (cost_i4 < (1 << 30)) cost_i4 += 8 * lambda; } /* intra mb_type/chroma-mode overhead in P slices */
    }

    mi->dbg_cost_inter = cost_inter; mi->dbg_cost_i16 = cost_i16; mi->dbg_cost_i4 = cost_i4;
    int type;
    if (!e->is_idr && cost_inter <= cost_i16 && cost_inter <= cost_i4) type = REF_MB_P16;
    else type = cost_i16 <= cost_i4 ? REF_MB_I16 : REF_MB_I4;
    mi->type = (uint8_t)type;

    if (type == REF_MB_I4) {
        int cbp = 0;
        memcpy(coef, c4, sizeof c4);
        memcpy(mi->i4_mode, m4, 16);
        memcpy(mi->nnz, n4, 16);
        for (int b = 0; b < 16; b++) if (n4[b]) cbp |= 1 << (b >> 2);
        mi->cbp = (uint8_t)cbp;
    } else if (type == REF_MB_I16) {
        encode_i16(e, mbx, mby, mode16, mi, coef);
    } else {
        code_inter_at_me(e, mbx, mby, mi, coef, -1);
        if (mi->part) {
            for (int p = 0; p < 4; p++) {
                int mp[2];
                predict_mv8(e, mbx, mby, p, mi->mv4, mp);
                mi->mvd4[p][0] = (int16_t)(mi->mv4[p][0] - mp[0]); mi->mvd4[p][1] = (int16_t)(mi->mv4[p][1] - mp[1]);
            }
            mi->mvd[0] = mi->mvd4[0][0]; mi->mvd[1] = mi->mvd4[0][1];
            return;
        }
        mi->mvd[0] = (int16_t)(mi->mv[0] - mvp[0]); mi->mvd[1] = (int16_t)(mi->mv[1] - mvp[1]);
        if (mi->cbp == 0 && mi->mv[0] == skipmv[0] && mi->mv[1] == skipmv[1]) mi->type = REF_MB_PSKIP;
        return;
    }

    if (type >= REF_MB_I16) intra_chroma(e, mbx, mby, mi, coef);
}

/* ---- upper-bound mode decision (ref_enc_set_effort) ------------------------------------------------------------------
 * NOT the parity oracle: a slow encoder over the same toolset (P_Skip, P_L0_16x16 with either transform size, P_8x8,
 * Intra16x16, Intra4x4, CAVLC) that picks, per macroblock, the candidate with the smallest SSD + lambda * bits, the bits counted
 * by writing the macroblock for real.  tools/rd_oracle.py runs it (with an exhaustive full-pel search and one slice per picture)
 * to bound how far the product's fast decisions and its slice structure are from the best this toolset does. */
typedef struct { ref_mb_info mi; int16_t coef[COEF_PER_MB]; uint8_t y[256], c[2][64]; } mb_snap;
static void snap_save(struct ref_enc *e, int mbx, int mby, mb_snap *s) {
    s->mi = *MBI(e, mbx, mby);
    memcpy(s->coef, COEF(e, mbx, mby), sizeof s->coef);
    for (int r = 0; r < 16; r++) memcpy(s->y + 16 * r, e->rec[0] + (mby * 16 + r) * e->sy + mbx * 16, 16);
    for (int p = 0; p < 2; p++) for (int r = 0; r < 8; r++) memcpy(s->c[p] + 8 * r, e->rec[1 + p] + (mby * 8 + r) * e->sc + mbx * 8, 8);
}
static void snap_load(struct ref_enc *e, int mbx, int mby, const mb_snap *s) {
    *MBI(e, mbx, mby) = s->mi;
    memcpy(COEF(e, mbx, mby), s->coef, sizeof s->coef);
    for (int r = 0; r < 16; r++) memcpy(e->rec[0] + (mby * 16 + r) * e->sy + mbx * 16, s->y + 16 * r, 16);
    for (int p = 0; p < 2; p++) for (int r = 0; r < 8; r++) memcpy(e->rec[1 + p] + (mby * 8 + r) * e->sc + mbx * 8, s->c[p] + 8 * r, 8);
}
static double rd_cost(struct ref_enc *e, int mbx, int mby, double lambda2) {
    uint8_t scratch[4096];
    int bp = 0, sr = 0;
    long ssd = 0;
    ref_mb_write(e, mbx, mby, scratch, &bp, &sr);
    for (int r = 0; r < 16; r++) {
        const uint8_t *a = e->cur[0] + (mby * 16 + r) * e->sy + mbx * 16, *b = e->rec[0] + (mby * 16 + r) * e->sy + mbx * 16;
        for (int x = 0; x < 16; x++) ssd += (a[x] - b[x]) * (a[x] - b[x]);
    }
    for (int p = 0; p < 2; p++)
        for (int r = 0; r < 8; r++) {
            const uint8_t *a = e->cur[1 + p] + (mby * 8 + r) * e->sc + mbx * 8, *b = e->rec[1 + p] + (mby * 8 + r) * e->sc + mbx * 8;
            for (int x = 0; x < 8; x++) ssd += (a[x] - b[x]) * (a[x] - b[x]);
        }
    return (double)ssd + lambda2 * bp;
}
void ref_mb_encode_rd(struct ref_enc *e, int mbx, int mby) {
    const int sy = e->sy, sc = e->sc;
    ref_mb_info *mi = MBI(e, mbx, mby);
    int16_t *coef = COEF(e, mbx, mby);
    const double lambda2 = 0.85 * pow(2.0, (e->qp - 12) / 3.0);
    const ref_mb_info keep = *mi; /* search results */
    mb_snap best;
    double bestj = 1e300;
#define RESET() do { *mi = keep; memset(coef, 0, COEF_PER_MB * sizeof(int16_t)); memset(mi->nnz, 0, sizeof mi->nnz); memset(mi->i4_mode, 2, 16); \
                     mi->cbp = 0; mi->mvd[0] = mi->mvd[1] = 0; mi->i16_mode = 0; mi->chroma_mode = 0; mi->part = 0; mi->t8 = 0; set_mv_all(mi, 0, 0); memset(mi->mvd4, 0, sizeof mi->mvd4); } while (0)
#define TRY() do { double j_ = rd_cost(e, mbx, mby, lambda2); if (j_ < bestj) { bestj = j_; snap_save(e, mbx, mby, &best); } } while (0)
    if (!e->is_idr) {
        int mvp[2], skipmv[2];
        predict_mv(e, mbx, mby, mvp, skipmv);
        { /* P_Skip: the prediction at the skip vector, no residual */
            uint8_t py[256], pc[2][64];
            RESET();
            ref_mc_luma(py, 16, e->ref[0] + mby * 16 * sy + mbx * 16, sy, skipmv[0], skipmv[1], 16, 16);
            copy_blk(e->rec[0] + mby * 16 * sy + mbx * 16, sy, py, 16, 16, 16);
            for (int pl = 0; pl < 2; pl++) {
                ref_mc_chroma(pc[pl], 8, e->ref[1 + pl] + mby * 8 * sc + mbx * 8, sc, skipmv[0], skipmv[1], 8, 8);
                copy_blk(e->rec[1 + pl] + mby * 8 * sc + mbx * 8, sc, pc[pl], 8, 8, 8);
            }
            mi->type = REF_MB_PSKIP; set_mv_all(mi, skipmv[0], skipmv[1]);
            TRY();
        }
        for (int part = 0; part <= (e->part8 && keep.me_cost8 > 0); part++)
            for (int t8 = 0; t8 <= 1; t8++) {
                RESET();
                mi->me_part = (uint8_t)part;
                mi->type = REF_MB_P16;
                code_inter_at_me(e, mbx, mby, mi, coef, t8);
                if (mi->part) {
                    for (int p = 0; p < 4; p++) {
                        int mp[2];
                        predict_mv8(e, mbx, mby, p, mi->mv4, mp);
                        mi->mvd4[p][0] = (int16_t)(mi->mv4[p][0] - mp[0]); mi->mvd4[p][1] = (int16_t)(mi->mv4[p][1] - mp[1]);
                    }
                    mi->mvd[0] = mi->mvd4[0][0]; mi->mvd[1] = mi->mvd4[0][1];
                } else { mi->mvd[0] = (int16_t)(mi->mv[0] - mvp[0]); mi->mvd[1] = (int16_t)(mi->mv[1] - mvp[1]); }
                TRY();
            }
    }
    { /* Intra16x16 */
        int mode16;
        RESET();
        analyse_i16(e, mbx, mby, &mode16, e->rec[0]);
        mi->type = REF_MB_I16;
        encode_i16(e, mbx, mby, mode16, mi, coef);
        intra_chroma(e, mbx, mby, mi, coef);
        TRY();
    }
    { /* Intra4x4 */
        int16_t c4[16][16];
        uint8_t m4[16], n4[16];
        int cbp = 0;
        RESET();
        analyse_i4(e, mbx, mby, c4, m4, n4, 0);
        memcpy(coef, c4, sizeof c4); memcpy(mi->i4_mode, m4, 16); memcpy(mi->nnz, n4, 16);
        for (int b = 0; b < 16; b++) if (n4[b]) cbp |= 1 << (b >> 2);
        mi->cbp = (uint8_t)cbp; mi->type = REF_MB_I4;
        intra_chroma(e, mbx, mby, mi, coef);
        TRY();
    }
#undef RESET
#undef TRY
    snap_load(e, mbx, mby, &best);
    MBI(e, mbx, mby)->me_part = keep.me_part;
}

/* ---- deblocking, H.264 8.7 (disable_deblocking_filter_idc = 2: slice-local) ---- */
static inline int clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }
static void filter_edge(uint8_t *pix, int xs, int ys, int n, const int *bS, int chroma, int qp) {
    int alpha = ref_alpha[qp], beta = ref_beta[qp];
    for (int i = 0; i < n; i++) {
        int bs = bS[chroma ? i >> 1 : i >> 2];
        uint8_t *p = pix + i * ys;
        if (!bs) continue;
        int p0 = p[-xs], p1 = p[-2 * xs], q0 = p[0], q1 = p[xs];
        if (iabs(p0 - q0) >= alpha || iabs(p1 - p0) >= beta || iabs(q1 - q0) >= beta) continue;
        if (chroma) {
            if (bs < 4) {
                int tc = ref_tc0[qp][bs - 1] + 1;
                int d = clip3(-tc, tc, (((q0 - p0) << 2) + (p1 - q1) + 4) >> 3);
                p[-xs] = (uint8_t)clip3(0, 255, p0 + d); p[0] = (uint8_t)clip3(0, 255, q0 - d);
            } else { p[-xs] = (uint8_t)((2 * p1 + p0 + q1 + 2) >> 2); p[0] = (uint8_t)((2 * q1 + q0 + p1 + 2) >> 2); }
            continue;
        }
        int p2 = p[-3 * xs], q2 = p[2 * xs];
        int ap = iabs(p2 - p0) < beta, aq = iabs(q2 - q0) < beta;
        if (bs < 4) {
            int tc0 = ref_tc0[qp][bs - 1], tc = tc0 + ap + aq;
            int d = clip3(-tc, tc, (((q0 - p0) << 2) + (p1 - q1) + 4) >> 3);
            p[-xs] = (uint8_t)clip3(0, 255, p0 + d); p[0] = (uint8_t)clip3(0, 255, q0 - d);
            if (ap) p[-2 * xs] = (uint8_t)(p1 + clip3(-tc0, tc0, (p2 + ((p0 + q0 + 1) >> 1) - (p1 << 1)) >> 1));
            if (aq) p[xs] = (uint8_t)(q1 + clip3(-tc0, tc0, (q2 + ((p0 + q0 + 1) >> 1) - (q1 << 1)) >> 1));
        } else {
            int small = iabs(p0 - q0) < ((alpha >> 2) + 2);
            if (ap && small) {
                int p3 = p[-4 * xs];
                p[-xs] = (uint8_t)((p2 + 2 * p1 + 2 * p0 + 2 * q0 + q1 + 4) >> 3);
                p[-2 * xs] = (uint8_t)((p2 + p1 + p0 + q0 + 2) >> 2);
                p[-3 * xs] = (uint8_t)((2 * p3 + 3 * p2 + p1 + p0 + q0 + 4) >> 3);
            } else p[-xs] = (uint8_t)((2 * p1 + p0 + q1 + 2) >> 2);
            if (aq && small) {
                int q3 = p[3 * xs];
                p[0] = (uint8_t)((p1 + 2 * p0 + 2 * q0 + 2 * q1 + q2 + 4) >> 3);
                p[xs] = (uint8_t)((p0 + q0 + q1 + q2 + 2) >> 2);
                p[2 * xs] = (uint8_t)((2 * q3 + 3 * q2 + q1 + q0 + p0 + 4) >> 3);
            } else p[0] = (uint8_t)((2 * q1 + q0 + p1 + 2) >> 2);
        }
    }
}
static int edge_bs(const ref_mb_info *mp, int bp, const ref_mb_info *mq, int bq, int mb_edge) {
    int ip = mp->type >= REF_MB_I16, iq = mq->type >= REF_MB_I16;
    if (ip || iq) return mb_edge ? 4 : 3;
    /* 8x8-transform MBs: 'the 8x8 block containing the sample has non-zero coefficients' (8.7.2.1) */
    if ((mp->t8 ? (mp->cbp >> (bp >> 2)) & 1 : mp->nnz[bp]) || (mq->t8 ? (mq->cbp >> (bq >> 2)) & 1 : mq->nnz[bq])) return 2;
    if (iabs(mp->mv4[bp >> 2][0] - mq->mv4[bq >> 2][0]) >= 4 || iabs(mp->mv4[bp >> 2][1] - mq->mv4[bq >> 2][1]) >= 4) return 1;
    return 0;
}
void ref_mb_deblock(struct ref_enc *e, int mbx, int mby) {
    const int sy = e->sy, sc = e->sc, qp = e->qp, qpc = ref_chroma_qp[qp];
    ref_mb_info *mq = MBI(e, mbx, mby);
    uint8_t *Y = e->rec[0] + mby * 16 * sy + mbx * 16;
    int bS[4];
    for (int dir = 0; dir < 2; dir++) { /* 0: vertical edges (filter across x), 1: horizontal edges */
        for (int ed = 0; ed < 4; ed++) {
            const ref_mb_info *mp = mq;
            if (ed == 0) {
                if (dir == 0) { if (!av_left(e, mbx, mby)) continue; mp = MBI(e, mbx - 1, mby); }
                else { if (!av_top(e, mbx, mby)) continue; mp = MBI(e, mbx, mby - 1); }
            }
            int any = 0;
            for (int k = 0; k < 4; k++) {
                int bq = dir == 0 ? blk_of[k][ed] : blk_of[ed][k];
                int bp = dir == 0 ? blk_of[k][ed ? ed - 1 : 3] : blk_of[ed ? ed - 1 : 3][k];
                bS[k] = edge_bs(mp, bp, mq, bq, ed == 0);
                any |= bS[k];
            }
            if (!any) continue;
            if (!(mq->t8 && (ed & 1))) { /* transform_size_8x8_flag: luma edges 1 and 3 are not transform edges */
                if (dir == 0) filter_edge(Y + ed * 4, 1, sy, 16, bS, 0, qp);
                else filter_edge(Y + ed * 4 * sy, sy, 1, 16, bS, 0, qp);
            }
            if (!(ed & 1))
                for (int pl = 1; pl < 3; pl++) {
                    uint8_t *C = e->rec[pl] + mby * 8 * sc + mbx * 8;
                    if (dir == 0) filter_edge(C + ed * 2, 1, sc, 8, bS, 1, qpc);
                    else filter_edge(C + ed * 2 * sc, sc, 1, 8, bS, 1, qpc);
                }
        }
    }
}

/* ---- macroblock_layer() CAVLC writer, H.264 7.3.5 ------------------------------- */
static int nnz_nb(struct ref_enc *e, int mbx, int mby, int base, int w, int bx, int by) {
    /* total_coeff of block (bx,by) of plane group 'base' (0 luma w=4; 16/20 chroma w=2); -1 unavailable */
    if (bx < 0) { if (!av_left(e, mbx, mby)) return -1; mbx--; bx = w - 1; }
    if (by < 0) { if (!av_top(e, mbx, mby)) return -1; mby--; by = w - 1; }
    return MBI(e, mbx, mby)->nnz[base + (w == 4 ? blk_of[by][bx] : by * 2 + bx)];
}
static int calc_nc(struct ref_enc *e, int mbx, int mby, int base, int w, int bx, int by) {
    int nA = nnz_nb(e, mbx, mby, base, w, bx - 1, by), nB = nnz_nb(e, mbx, mby, base, w, bx, by - 1);
    if (nA >= 0 && nB >= 0) return (nA + nB + 1) >> 1;
    return nA >= 0 ? nA : (nB >= 0 ? nB : 0);
}
void ref_mb_write(struct ref_enc *e, int mbx, int mby, uint8_t *out, int *bp, int *skip_run) {
    ref_mb_info *mi = MBI(e, mbx, mby);
    const int16_t *coef = COEF(e, mbx, mby);
    int start = *bp, P = !e->is_idr;
    if (mi->type == REF_MB_PSKIP) { (*skip_run)++; mi->bits = 0; return; }
    if (P) { ref_put_ue(out, bp, (unsigned)*skip_run); *skip_run = 0; }
    int cl = mi->cbp & 15, cc = mi->cbp >> 4;
    if (mi->type == REF_MB_P16 && mi->part) { /* P_8x8: mb_type 3, four sub_mb_type P_L0_8x8 (0), no ref_idx (one reference), four mvd */
        ref_put_ue(out, bp, 3);
        for (int p = 0; p < 4; p++) ref_put_ue(out, bp, 0);
        for (int p = 0; p < 4; p++) { ref_put_se(out, bp, mi->mvd4[p][0]); ref_put_se(out, bp, mi->mvd4[p][1]); }
    } else if (mi->type == REF_MB_P16) {
        ref_put_ue(out, bp, 0);
        ref_put_se(out, bp, mi->mvd[0]); ref_put_se(out, bp, mi->mvd[1]);
    } else if (mi->type == REF_MB_I4) {
        ref_put_ue(out, bp, P ? 5 : 0);
        ref_put_bits(out, bp, 1, 0); /* transform_size_8x8_flag (PPS transform_8x8_mode_flag = 1): Intra4x4, not Intra8x8 */
        for (int blk = 0; blk < 16; blk++) {
            int bx = ref_blk_x[blk], by = ref_blk_y[blk], mA, mB;
            if (bx > 0) mA = mi->i4_mode[blk_of[by][bx - 1]];
            else if (av_left(e, mbx, mby)) { ref_mb_info *n = MBI(e, mbx - 1, mby); mA = n->type == REF_MB_I4 ? n->i4_mode[blk_of[by][3]] : 2; }
            else mA = -1;
            if (by > 0) mB = mi->i4_mode[blk_of[by - 1][bx]];
            else if (av_top(e, mbx, mby)) { ref_mb_info *n = MBI(e, mbx, mby - 1); mB = n->type == REF_MB_I4 ? n->i4_mode[blk_of[3][bx]] : 2; }
            else mB = -1;
            int pm = (mA < 0 || mB < 0) ? 2 : (mA < mB ? mA : mB), m = mi->i4_mode[blk];
            if (m == pm) ref_put_bits(out, bp, 1, 1);
            else ref_put_bits(out, bp, 4, (uint32_t)(m < pm ? m : m - 1));
        }
    } else {
        ref_put_ue(out, bp, (unsigned)((P ? 5 : 0) + 1 + mi->i16_mode + 4 * cc + (cl ? 12 : 0)));
    }
    if (mi->type >= REF_MB_I16) ref_put_ue(out, bp, mi->chroma_mode);
    if (mi->type != REF_MB_I16) {
        const uint8_t *tab = mi->type == REF_MB_I4 ? ref_golomb_to_intra_cbp : ref_golomb_to_inter_cbp;
        unsigned code = 0;
        while (tab[code] != mi->cbp) code++;
        ref_put_ue(out, bp, code);
        if (mi->type == REF_MB_P16 && cl) ref_put_bits(out, bp, 1, mi->t8); /* transform_size_8x8_flag */
    }
    if (mi->type == REF_MB_I16 || mi->cbp) {
        ref_put_se(out, bp, 0); /* mb_qp_delta */
        if (mi->type == REF_MB_I16) ref_cavlc_block(out, bp, coef + 24 * 16, 16, calc_nc(e, mbx, mby, 0, 4, 0, 0));
        for (int blk = 0; blk < 16; blk++) {
            if (!(cl & (1 << (blk >> 2)))) continue;
            int nC = calc_nc(e, mbx, mby, 0, 4, ref_blk_x[blk], ref_blk_y[blk]);
            ref_cavlc_block(out, bp, coef + blk * 16, mi->type == REF_MB_I16 ? 15 : 16, nC);
        }
        if (cc) for (int p = 0; p < 2; p++) ref_cavlc_block(out, bp, coef + (25 + p) * 16, 4, -1);
        if (cc & 2)
            for (int p = 0; p < 2; p++)
                for (int b = 0; b < 4; b++)
                    ref_cavlc_block(out, bp, coef + (16 + p * 4 + b) * 16, 15, calc_nc(e, mbx, mby, 16 + p * 4, 2, b & 1, b >> 1));
    }
    mi->bits = *bp - start;
}
