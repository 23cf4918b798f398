/*
 * oracle/ref_cavlc.c — bit writer, Exp-Golomb (H.264 9.1) and CAVLC residual blocks
 * (H.264 9.2 / 7.3.5.3.2 residual_block_cavlc).  TEST INFRASTRUCTURE ONLY.
 * Reference anchor: x264enc slot, /root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6
 * (entropy coder of the external encoder [EXT]); pinned by libavcodec decode equality.
 */
#include "ref_bits.h"
#include "ref_h264.h"
#include "ref_tables.h"

void ref_put_bits(uint8_t *out, int *bitpos, int n, uint32_t v) {
    for (int i = n - 1; i >= 0; i--) {
        int p = (*bitpos)++;
        if (!(p & 7)) out[p >> 3] = 0;
        if ((v >> i) & 1) out[p >> 3] |= (uint8_t)(0x80 >> (p & 7));
    }
}
void ref_put_ue(uint8_t *out, int *bitpos, unsigned v) {
    int n = 0;
    while ((v + 1) >> (n + 1)) n++;
    ref_put_bits(out, bitpos, n, 0);
    ref_put_bits(out, bitpos, n + 1, v + 1);
}
void ref_put_se(uint8_t *out, int *bitpos, int v) { ref_put_ue(out, bitpos, v > 0 ? 2u * v - 1 : 2u * (unsigned)(-v)); }

int ref_cavlc_block(uint8_t *out, int *bitpos, const int16_t *zz, int max_coeff, int nC) {
    int start = *bitpos;
    int level[16], run[16]; /* in reverse scan order: [0] = highest-frequency nonzero */
    int total = 0, zeros_after = 0, total_zeros = 0;
    int last = max_coeff - 1;
    while (last >= 0 && zz[last] == 0) last--;
    for (int i = last; i >= 0; i--) {
        if (zz[i]) { level[total] = zz[i]; run[total] = 0; total++; }
        else { run[total - 1]++; total_zeros++; }
    }
    (void)zeros_after;
    int t1 = 0;
    while (t1 < 3 && t1 < total && (level[t1] == 1 || level[t1] == -1)) t1++;

    /* coeff_token */
    if (nC < 0)
        ref_put_bits(out, bitpos, ref_chroma_dc_coeff_token_len[4 * total + t1], ref_chroma_dc_coeff_token_bits[4 * total + t1]);
    else {
        int tab = nC < 2 ? 0 : (nC < 4 ? 1 : (nC < 8 ? 2 : 3));
        ref_put_bits(out, bitpos, ref_coeff_token_len[tab][4 * total + t1], ref_coeff_token_bits[tab][4 * total + t1]);
    }
    if (!total) return *bitpos - start;
    for (int i = 0; i < t1; i++) ref_put_bits(out, bitpos, 1, level[i] < 0);

    /* levels (H.264 9.2.2.1 inverted) */
    int sl = (total > 10 && t1 < 3) ? 1 : 0;
    for (int i = t1; i < total; i++) {
        int lv = level[i], a = lv < 0 ? -lv : lv;
        int code = 2 * a - 2 + (lv < 0);
        if (i == t1 && t1 < 3) code -= 2;
        if (sl == 0) {
            if (code < 14) ref_put_bits(out, bitpos, code + 1, 1);
            else if (code < 30) { ref_put_bits(out, bitpos, 15, 1); ref_put_bits(out, bitpos, 4, (uint32_t)(code - 14)); }
            else { ref_put_bits(out, bitpos, 16, 1); ref_put_bits(out, bitpos, 12, (uint32_t)(code - 30)); }
        } else {
            if ((code >> sl) < 15) { ref_put_bits(out, bitpos, (code >> sl) + 1, 1); ref_put_bits(out, bitpos, sl, (uint32_t)(code & ((1 << sl) - 1))); }
            else { ref_put_bits(out, bitpos, 16, 1); ref_put_bits(out, bitpos, 12, (uint32_t)(code - (15 << sl))); }
        }
        if (sl == 0) sl = 1;
        if (a > (3 << (sl - 1)) && sl < 6) sl++;
    }

    /* total_zeros */
    if (total < max_coeff) {
        if (max_coeff == 4)
            ref_put_bits(out, bitpos, ref_chroma_dc_total_zeros_len[total - 1][total_zeros], ref_chroma_dc_total_zeros_bits[total - 1][total_zeros]);
        else
            ref_put_bits(out, bitpos, ref_total_zeros_len[total - 1][total_zeros], ref_total_zeros_bits[total - 1][total_zeros]);
    }
    /* run_before */
    int zl = total_zeros;
    for (int i = 0; i < total - 1 && zl > 0; i++) {
        int t = (zl > 7 ? 7 : zl) - 1;
        ref_put_bits(out, bitpos, ref_run_len[t][run[i]], ref_run_bits[t][run[i]]);
        zl -= run[i];
    }
    return *bitpos - start;
}
