/*
 * cc_cavlc.c — host CAVLC writer: macroblock_layer() and residual_block_cavlc()
 * of ITU-T H.264 7.3.5 / 9.2, fed by the compact per-row symbol stream the encode kernel emits
 * (per macroblock: the 16-bit significance masks of its coded 4x4 blocks in scan order, then their non-zero levels).
 * north_star keeps the bitstream writer on the host in C; what crosses PCIe is ~1-3x the final
 * bitstream, not dense coefficients.
 */
#include "cc_host.h"
#include "cc_vlc_tables.h"
#include <string.h>

static inline void put_vlc(cc_bw *b, uint16_t e) { cc_bw_put(b, e >> 8, e & 0xff); }

/* residual_block_cavlc for one block: mask bit i = scan position i is non-zero, lv = the non-zero levels in
 * ascending scan order.  max_coeff 16, 15 or 4 (chroma DC); nC < 0 selects the chroma-DC tables. */
static void write_block(cc_bw *b, unsigned mask, const int16_t *lv, int max_coeff, int nC) {
    const int total = __builtin_popcount(mask);
    if (!total) {
        put_vlc(b, nC < 0 ? cc_vlc_coeff_token_cdc[0][0] : cc_vlc_coeff_token[nC < 2 ? 0 : (nC < 4 ? 1 : (nC < 8 ? 2 : 3))][0]);
        return;
    }
    int t1 = 0;
    while (t1 < 3 && t1 < total && (lv[total - 1 - t1] == 1 || lv[total - 1 - t1] == -1)) t1++;
    put_vlc(b, nC < 0 ? cc_vlc_coeff_token_cdc[0][4 * total + t1]
                      : cc_vlc_coeff_token[nC < 2 ? 0 : (nC < 4 ? 1 : (nC < 8 ? 2 : 3))][4 * total + t1]);
    for (int i = 0; i < t1; i++) cc_bw_put(b, 1, lv[total - 1 - i] < 0);
    /* levels, highest frequency first (9.2.2.1 read backwards) */
    int sl = (total > 10 && t1 < 3) ? 1 : 0;
    for (int i = t1; i < total; i++) {
        int v = lv[total - 1 - i], a = v < 0 ? -v : v;
        int code = 2 * a - 2 + (v < 0);
        if (i == t1 && t1 < 3) code -= 2;
        if (sl == 0) {
            if (code < 14) cc_bw_put(b, code + 1, 1);
            else if (code < 30) { cc_bw_put(b, 15, 1); cc_bw_put(b, 4, (uint32_t)(code - 14)); }
            else { cc_bw_put(b, 16, 1); cc_bw_put(b, 12, (uint32_t)(code - 30)); }
        } else {
            int pre = code >> sl;
            if (pre < 15) { cc_bw_put(b, pre + 1, 1); cc_bw_put(b, sl, (uint32_t)(code & ((1 << sl) - 1))); }
            else { cc_bw_put(b, 16, 1); cc_bw_put(b, 12, (uint32_t)(code - (15 << sl))); }
        }
        if (sl == 0) sl = 1;
        if (a > (3 << (sl - 1)) && sl < 6) sl++;
    }
    const int last = 31 - __builtin_clz(mask); /* highest non-zero scan position */
    int zeros_left = last + 1 - total;
    if (total < max_coeff) put_vlc(b, max_coeff == 4 ? cc_vlc_total_zeros_cdc[total - 1][zeros_left] : cc_vlc_total_zeros[total - 1][zeros_left]);
    unsigned m = mask;
    for (int i = 0; i < total - 1 && zeros_left > 0; i++) {
        int hi = 31 - __builtin_clz(m);
        m &= ~(1u << hi);
        int run = hi - (31 - __builtin_clz(m)) - 1;
        put_vlc(b, cc_vlc_run_before[(zeros_left > 7 ? 7 : zeros_left) - 1][run]);
        zeros_left -= run;
    }
}

/* z-order block index of (bx,by) in 4-pixel units */
static const uint8_t blk_of[4][4] = {{0, 1, 4, 5}, {2, 3, 6, 7}, {8, 9, 12, 13}, {10, 11, 14, 15}};
static const uint8_t blk_x[16] = {0, 1, 0, 1, 2, 3, 2, 3, 0, 1, 0, 1, 2, 3, 2, 3};
static const uint8_t blk_y[16] = {0, 0, 1, 1, 0, 0, 1, 1, 2, 2, 3, 3, 2, 2, 3, 3};

/* nC of 9.2.1: average of the left and top neighbours' total_coeff where available */
static inline int luma_nc(const cc_cavlc_ctx *c, const uint8_t *nz, int mbx, int has_top, int bx, int by) {
    int nA = -1, nB = -1;
    if (bx > 0) nA = nz[blk_of[by][bx - 1]];
    else if (mbx > 0) nA = nz[-24 + blk_of[by][3]];
    if (by > 0) nB = nz[blk_of[by - 1][bx]];
    else if (has_top) nB = nz[-24 * c->mb_w + blk_of[3][bx]];
    if (nA >= 0 && nB >= 0) return (nA + nB + 1) >> 1;
    return nA >= 0 ? nA : (nB >= 0 ? nB : 0);
}
static inline int chroma_nc(const cc_cavlc_ctx *c, const uint8_t *nz, int mbx, int has_top, int base, int bx, int by) {
    int nA = -1, nB = -1;
    if (bx > 0) nA = nz[base + by * 2]; else if (mbx > 0) nA = nz[-24 + base + by * 2 + 1];
    if (by > 0) nB = nz[base + bx]; else if (has_top) nB = nz[-24 * c->mb_w + base + 2 + bx];
    if (nA >= 0 && nB >= 0) return (nA + nB + 1) >> 1;
    return nA >= 0 ? nA : (nB >= 0 ? nB : 0);
}

int cc_cavlc_write_slice(cc_cavlc_ctx *c, cc_bw *b, int is_idr, int row0, int row1, const cc_mbhost *mbh,
                         const cc_rowhdr *rowhdr, const uint8_t *const *row_payload) {
    const int P = !is_idr;
    unsigned skip_run = 0;
    for (int mby = row0; mby < row1; mby++) {
        const uint8_t *pay = row_payload[mby], *pay_end = pay + rowhdr[mby].bytes;
        const int has_top = mby > row0;
        for (int mbx = 0; mbx < c->mb_w; mbx++) {
            const int mb = mby * c->mb_w + mbx;
            const cc_mbhost *m = &mbh[mb];
            uint8_t *nz = c->nnz + (size_t)mb * 24;
            memset(nz, 0, 24);
            if (m->type == CC_MB_PSKIP) { if (!P) return -1; skip_run++; continue; }
            if (P) { cc_bw_ue(b, skip_run); skip_run = 0; }
            pay = row_payload[mby] + (((size_t)(pay - row_payload[mby]) + 3) & ~(size_t)3); /* every MB's symbols start 4-byte aligned */
            const int cl = m->cbp & 15, cc = m->cbp >> 4, i16 = m->type == CC_MB_I16;
            if (m->type == CC_MB_P16 && (m->t8 & 2)) { /* P_8x8: mb_type 3, four sub_mb_type P_L0_8x8 (ue 0), no ref_idx (one reference), four mvd */
                if (!P || pay + 16 > pay_end) return -1;
                cc_bw_ue(b, 3);
                cc_bw_put(b, 4, 15);
                for (int i = 0; i < 8; i++) cc_bw_se(b, ((const int16_t *)pay)[i]);
                pay += 16;
            } else if (m->type == CC_MB_P16) {
                if (!P) return -1;
                cc_bw_ue(b, 0);
                cc_bw_se(b, m->mvd[0]); cc_bw_se(b, m->mvd[1]);
            } else if (m->type == CC_MB_I4) {
                cc_bw_ue(b, P ? 5 : 0);
                cc_bw_put(b, 1, 0); /* transform_size_8x8_flag: Intra4x4 (the PPS enables transform_8x8_mode) */
                if (pay + 8 > pay_end) return -1;
                for (int i = 0; i < 16; i++) { /* nibble: 8 = prev_intra4x4_pred_mode_flag, else rem_intra4x4_pred_mode */
                    int nib = (pay[i >> 1] >> (4 * (i & 1))) & 15;
                    if (nib & 8) cc_bw_put(b, 1, 1); else cc_bw_put(b, 4, (uint32_t)nib);
                }
                pay += 8;
            } else if (i16) {
                cc_bw_ue(b, (unsigned)((P ? 5 : 0) + 1 + (m->modes & 3) + 4 * cc + (cl ? 12 : 0)));
            } else return -1;
            if (m->type >= CC_MB_I16) cc_bw_ue(b, (unsigned)(m->modes >> 2) & 3);
            if (!i16) cc_bw_ue(b, cc_vlc_cbp_code[m->type == CC_MB_I4 ? 0 : 1][m->cbp]);
            if (m->type == CC_MB_P16 && cl) cc_bw_put(b, 1, m->t8 & 1); /* transform_size_8x8_flag */
            if (!i16 && !m->cbp) continue;
            cc_bw_se(b, 0); /* mb_qp_delta */
            /* the slot's residual part: one significance mask per coded block in syntax order, then all levels (cc_k_mb.cu emit_mb) */
            const int nblk = (i16 ? 1 + (cl ? 16 : 0) : 4 * __builtin_popcount((unsigned)cl)) + (cc ? 2 : 0) + (cc == 2 ? 8 : 0);
            if (pay + 2 * nblk > pay_end) return -1;
            const uint8_t *mk = pay;
            pay += 2 * nblk;
#define NEXT_BLOCK(mask_, lv_)                                              \
    do {                                                                    \
        mask_ = (unsigned)mk[0] | ((unsigned)mk[1] << 8);                   \
        mk += 2;                                                            \
        lv_ = (const int16_t *)pay;                                         \
        pay += 2 * __builtin_popcount(mask_);                               \
        if (pay > pay_end) return -1;                                       \
    } while (0)
            unsigned mask;
            const int16_t *lv;
            if (i16) {
                NEXT_BLOCK(mask, lv);
                write_block(b, mask, lv, 16, luma_nc(c, nz, mbx, has_top, 0, 0));
            }
            for (int blk = 0; blk < 16; blk++) {
                if (!(cl & (1 << (blk >> 2)))) continue;
                NEXT_BLOCK(mask, lv);
                write_block(b, mask, lv, i16 ? 15 : 16, luma_nc(c, nz, mbx, has_top, blk_x[blk], blk_y[blk]));
                nz[blk] = (uint8_t)__builtin_popcount(mask);
            }
            if (cc) {
                for (int p = 0; p < 2; p++) { NEXT_BLOCK(mask, lv); write_block(b, mask, lv, 4, -1); }
                if (cc & 2)
                    for (int p = 0; p < 2; p++)
                        for (int i = 0; i < 4; i++) {
                            NEXT_BLOCK(mask, lv);
                            write_block(b, mask, lv, 15, chroma_nc(c, nz, mbx, has_top, 16 + 4 * p, i & 1, i >> 1));
                            nz[16 + 4 * p + i] = (uint8_t)__builtin_popcount(mask);
                        }
            }
        }
    }
    if (skip_run) cc_bw_ue(b, skip_run);
    return 0;
}
