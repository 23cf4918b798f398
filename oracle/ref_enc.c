/*
 * oracle/ref_enc.c — frame level of the CPU reference encoder: padding, half-res pyramid,
 * two-level motion search, slice loop, deblock pass, border extension, SPS/PPS/slice
 * headers (H.264 7.3.2.1, 7.3.2.2, 7.3.3) and Annex-B NAL packaging (7.4.1, B.1).
 * TEST INFRASTRUCTURE ONLY (see ref_h264.h).  Stands in for the external x264enc slot at
 * /root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6 (GOP/key-int 60 and
 * 8-bit 4:2:0 from that line); the bitrate that drives qp comes from
 * /root/reference/src/gst/encoder_control.c:45-57.
 */
#include "ref_bits.h"
#include "ref_int.h"
#include "ref_tables.h"
#include <stdlib.h>
#include <string.h>

static inline int iabs(int x) { return x < 0 ? -x : x; }

/* ---- allocation ------------------------------------------------------------------ */
static uint8_t *plane_alloc(int w, int h, int pad, int *stride, uint8_t **base) {
    *stride = w + 2 * pad;
    *base = (uint8_t *)calloc((size_t)(*stride) * (h + 2 * pad), 1);
    return *base + (size_t)pad * (*stride) + pad;
}
ref_enc *ref_enc_open(int width, int height, int fps_n, int fps_d, int rows_per_slice) {
    if (width < 16 || height < 16 || (width & 1) || (height & 1) || rows_per_slice < 1) return 0;
    ref_enc *e = (ref_enc *)calloc(1, sizeof *e);
    e->w = width; e->h = height; e->mb_w = (width + 15) / 16; e->mb_h = (height + 15) / 16;
    e->rps = rows_per_slice; e->fps_n = fps_n; e->fps_d = fps_d; e->part8 = 1;
    int W = e->mb_w * 16, H = e->mb_h * 16, k = 0;
    for (int s = 0; s < 3; s++) { /* cur, rec, ref */
        uint8_t **dst = s == 0 ? e->cur : (s == 1 ? e->rec : e->ref);
        dst[0] = plane_alloc(W, H, REF_PAD, &e->sy, &e->alloc[k++]);
        dst[1] = plane_alloc(W / 2, H / 2, REF_PAD / 2, &e->sc, &e->alloc[k++]);
        dst[2] = plane_alloc(W / 2, H / 2, REF_PAD / 2, &e->sc, &e->alloc[k++]);
    }
    e->l1cur = plane_alloc(W / 2, H / 2, REF_PAD / 2, &e->s1, &e->l1alloc[0]);
    e->l1ref = plane_alloc(W / 2, H / 2, REF_PAD / 2, &e->s1, &e->l1alloc[1]);
    e->mb = (ref_mb_info *)calloc((size_t)e->mb_w * e->mb_h, sizeof(ref_mb_info));
    e->coef = (int16_t *)calloc((size_t)e->mb_w * e->mb_h * COEF_PER_MB, sizeof(int16_t));
    int ns = (e->mb_h + e->rps - 1) / e->rps;
    e->slice_cap = e->rps * e->mb_w * 2048 + 256; /* generous: a CAVLC MB at qp>=10 stays far below 2 KiB */
    e->slice_buf = (uint8_t *)malloc((size_t)ns * e->slice_cap);
    e->slice_bits = (int *)calloc((size_t)ns, sizeof(int));
    return e;
}
void ref_enc_close(ref_enc *e) {
    if (!e) return;
    for (int i = 0; i < 9; i++) free(e->alloc[i]);
    free(e->l1alloc[0]); free(e->l1alloc[1]); free(e->mb); free(e->coef); free(e->slice_buf); free(e->slice_bits);
    free(e);
}
void ref_enc_set_partitions(ref_enc *e, int on) { e->part8 = on != 0; }
void ref_enc_set_effort(ref_enc *e, int on) { e->effort = on != 0; }
#ifdef _OPENMP
#include <omp.h>
void ref_set_threads(int n) { if (n > 0) omp_set_num_threads(n); } /* launchers such as torchrun export OMP_NUM_THREADS=1 */
#else
void ref_set_threads(int n) { (void)n; }
#endif
int ref_enc_mb_w(const ref_enc *e) { return e->mb_w; }
int ref_enc_mb_h(const ref_enc *e) { return e->mb_h; }
const ref_mb_info *ref_enc_mb_info(const ref_enc *e) { return e->mb; }
const uint8_t *ref_enc_l1_cur(const ref_enc *e, int *stride) { *stride = e->s1; return e->l1cur; }
void ref_enc_get_recon(const ref_enc *e, uint8_t *y, uint8_t *u, uint8_t *v) {
    /* after encode() rec/ref were swapped: the newest reconstruction is e->ref */
    int W = e->mb_w * 16, H = e->mb_h * 16;
    for (int r = 0; r < H; r++) memcpy(y + (size_t)r * W, e->ref[0] + (size_t)r * e->sy, (size_t)W);
    for (int r = 0; r < H / 2; r++) {
        memcpy(u + (size_t)r * (W / 2), e->ref[1] + (size_t)r * e->sc, (size_t)W / 2);
        memcpy(v + (size_t)r * (W / 2), e->ref[2] + (size_t)r * e->sc, (size_t)W / 2);
    }
}

/* ---- plane helpers ---------------------------------------------------------------- */
static void load_plane(uint8_t *dst, int sd, const uint8_t *src, int ss, int w, int h, int W, int H) {
    for (int y = 0; y < H; y++) { /* copy + replicate right/bottom edge up to the coded size */
        const uint8_t *s = src + (size_t)(y < h ? y : h - 1) * ss;
        uint8_t *d = dst + (size_t)y * sd;
        memcpy(d, s, (size_t)w);
        for (int x = w; x < W; x++) d[x] = s[w - 1];
    }
}
static void extend_border(uint8_t *p, int s, int W, int H, int pad) {
    for (int y = 0; y < H; y++) {
        uint8_t *r = p + (size_t)y * s;
        memset(r - pad, r[0], (size_t)pad);
        memset(r + W, r[W - 1], (size_t)pad);
    }
    for (int y = 1; y <= pad; y++) {
        memcpy(p - (size_t)y * s - pad, p - pad, (size_t)W + 2 * pad);
        memcpy(p + (size_t)(H - 1 + y) * s - pad, p + (size_t)(H - 1) * s - pad, (size_t)W + 2 * pad);
    }
}
/* half-res plane: 2x2 box mean with rounding, over the picture AND its border */
static void downsample2(uint8_t *dst, int sd, const uint8_t *src, int ss, int W, int H, int pad) {
    for (int y = -pad / 2; y < H / 2 + pad / 2; y++)
        for (int x = -pad / 2; x < W / 2 + pad / 2; x++) {
            const uint8_t *p = src + (size_t)(2 * y) * ss + 2 * x;
            dst[(size_t)y * sd + x] = (uint8_t)((p[0] + p[1] + p[ss] + p[ss + 1] + 2) >> 2);
        }
}

/* ---- motion estimation ------------------------------------------------------------ */
/* level 1: exhaustive +-16 at half resolution on the MB's 8x8 footprint.
 * cost = 4*SAD + |dx| + |dy|; ties -> lowest raster candidate index */
static void me_l1(ref_enc *e, int mbx, int mby) {
    const uint8_t *cur = e->l1cur + (size_t)(mby * 8) * e->s1 + mbx * 8;
    const uint8_t *ref = e->l1ref + (size_t)(mby * 8) * e->s1 + mbx * 8;
    unsigned best = ~0u;
    for (int dy = -16; dy <= 16; dy++)
        for (int dx = -16; dx <= 16; dx++) {
            unsigned sad = (unsigned)ref_sad(cur, e->s1, ref + dy * e->s1 + dx, e->s1, 8, 8);
            unsigned key = ((4 * sad + (unsigned)(iabs(dx) + iabs(dy))) << 11) | (unsigned)((dy + 16) * 33 + dx + 16);
            if (key < best) best = key;
        }
    int idx = (int)(best & 2047);
    ref_mb_info *mi = &e->mb[mby * e->mb_w + mbx];
    mi->l1_mv[0] = (int16_t)(idx % 33 - 16); mi->l1_mv[1] = (int16_t)(idx / 33 - 16);
}
/* level 0: integer refinement (SAD) + half/quarter-pel (SATD) around 2*l1_mv */
static void me_l0(ref_enc *e, int mbx, int mby) {
    const int sy = e->sy, lambda = ref_lambda[e->qp], mw = e->mb_w;
    ref_mb_info *mi = &e->mb[mby * mw + mbx];
    const uint8_t *cur = e->cur[0] + (size_t)(mby * 16) * sy + mbx * 16;
    const uint8_t *ref = e->ref[0] + (size_t)(mby * 16) * sy + mbx * 16;
    int L = mbx > 0, T = mby > ref_slice_first_row(e, mby), C = T && mbx + 1 < mw, pm[2];
    for (int c = 0; c < 2; c++) { /* approximate predictor: median of neighbours' coarse vectors (full-pel) */
        int a = L ? 2 * mi[-1].l1_mv[c] : 0, b = T ? 2 * mi[-mw].l1_mv[c] : 0;
        int cc = C ? 2 * mi[-mw + 1].l1_mv[c] : ((T && L) ? 2 * mi[-mw - 1].l1_mv[c] : 0);
        int cav = C || (T && L);
        pm[c] = (!T && !cav && L) ? a : ref_median3(a, b, cc);
    }
    unsigned best = ~0u;
    int bmx = 0, bmy = 0;
    if (e->effort) { /* upper-bound mode: every full-pel vector within +-24 of the coarse result (and the predictor / zero as below) */
        for (int dy = -24; dy <= 24; dy++)
            for (int dx = -24; dx <= 24; dx++) {
                int mx = 2 * mi->l1_mv[0] + dx, my = 2 * mi->l1_mv[1] + dy;
                if (mx < -40 || mx > 40 || my < -40 || my > 40) continue; /* stay inside the padded reference */
                unsigned c = (unsigned)ref_sad(cur, sy, ref + my * sy + mx, sy, 16, 16) +
                             (unsigned)(lambda * (ref_mv_bits(4 * (mx - pm[0])) + ref_mv_bits(4 * (my - pm[1]))));
                if (c < (best >> 5)) { best = c << 5; bmx = mx; bmy = my; }
            }
    }
    for (int i = 0; i < 27; i++) {
        int mx, my;
        if (i == 0) { mx = pm[0]; my = pm[1]; }
        else if (i == 1) { mx = 0; my = 0; }
        else { mx = 2 * mi->l1_mv[0] + (i - 2) % 5 - 2; my = 2 * mi->l1_mv[1] + (i - 2) / 5 - 2; }
        unsigned c = (unsigned)ref_sad(cur, sy, ref + my * sy + mx, sy, 16, 16) +
                     (unsigned)(lambda * (ref_mv_bits(4 * (mx - pm[0])) + ref_mv_bits(4 * (my - pm[1]))));
        unsigned key = (c << 5) | (unsigned)i;
        if (key < best) { best = key; bmx = mx; bmy = my; }
    }
    /* sub-pel: 9 half-pel positions then 8 quarter-pel, SATD + lambda*bits; first-best wins ties */
    uint8_t pred[256];
    int bx = 4 * bmx, by = 4 * bmy, bcost = 1 << 30, bsatd = 0;
    for (int step = 2; step >= 1; step--) {
        int cx = bx, cy = by;
        for (int i = (step == 2 ? -1 : 0); i < 9; i++) {
            int dx, dy;
            if (i < 0) { dx = dy = 0; }
            else { if (i == 4) continue; dx = (i % 3 - 1) * step; dy = (i / 3 - 1) * step; }
            int mx = cx + dx, my = cy + dy;
            ref_mc_luma(pred, 16, ref, sy, mx, my, 16, 16);
            int satd = ref_satd(cur, sy, pred, 16, 16, 16);
            int c = satd + lambda * (ref_mv_bits(mx - 4 * pm[0]) + ref_mv_bits(my - 4 * pm[1]));
            if (c < bcost) { bcost = c; bsatd = satd; bx = mx; by = my; }
        }
    }
    mi->me_mv[0] = (int16_t)bx; mi->me_mv[1] = (int16_t)by; mi->me_satd = bsatd;
    mi->me_cost16 = bcost; mi->me_cost8 = 0; mi->me_part = 0;
    if (!e->part8) return;
    /* 8x8 partitions (north_star: "motion estimation over 16x16/8x8 partitions"): every quadrant searches +-2 full-pel
     * around the 16x16 integer vector by SAD, then the same half/quarter-pel SATD refinement.  P_8x8 is chosen when the
     * summed cost plus 8 lambda (mb_type + four sub_mb_type code lengths over P_L0_16x16) beats the 16x16 cost. */
    int cost8 = 8 * lambda, f8[4][2];
    unsigned full8 = (unsigned)(8 * lambda);
    /* The quadrants search the 5 x 5 full-pel window of the macroblock's own search (around the coarse vector): the product gets
     * their SADs from the strips it has just read.  When the predictor or the zero vector won from outside that window the
     * macroblock stays one partition.  (The upper-bound mode centres the window on its own winner.) */
    const int cx8 = e->effort ? bmx : 2 * mi->l1_mv[0], cy8 = e->effort ? bmy : 2 * mi->l1_mv[1];
    memset(mi->me_mv8, 0, sizeof mi->me_mv8); memset(mi->me_satd8, 0, sizeof mi->me_satd8);
    if (iabs(bmx - cx8) > 2 || iabs(bmy - cy8) > 2) return;
    for (int p = 0; p < 4; p++) {
        const uint8_t *c8 = cur + (p >> 1) * 8 * sy + (p & 1) * 8, *r8 = ref + (p >> 1) * 8 * sy + (p & 1) * 8;
        unsigned b8 = ~0u;
        int fx = cx8, fy = cy8;
        for (int i = 0; i < 25; i++) {
            int mx = cx8 + i % 5 - 2, my = cy8 + i / 5 - 2;
            unsigned c = (unsigned)ref_sad(c8, sy, r8 + my * sy + mx, sy, 8, 8) +
                         (unsigned)(lambda * (ref_mv_bits(4 * (mx - pm[0])) + ref_mv_bits(4 * (my - pm[1]))));
            unsigned key = (c << 5) | (unsigned)i;
            if (key < b8) { b8 = key; fx = mx; fy = my; }
        }
        f8[p][0] = fx; f8[p][1] = fy;
        full8 += b8 >> 5;
    }
    /* partitions pay only where parts of the macroblock move differently: unless the four full-pel quadrant costs plus the
     * partition overhead already beat the full-pel 16x16 cost, the sub-pel refinement of the quadrants is not run and the
     * macroblock stays one partition */
    memset(mi->me_mv8, 0, sizeof mi->me_mv8); memset(mi->me_satd8, 0, sizeof mi->me_satd8);
    if (full8 >= (best >> 5) && !e->effort) return;
    for (int p = 0; p < 4; p++) {
        const uint8_t *c8 = cur + (p >> 1) * 8 * sy + (p & 1) * 8, *r8 = ref + (p >> 1) * 8 * sy + (p & 1) * 8;
        const int fx = f8[p][0], fy = f8[p][1];
        int qx = 4 * fx, qy = 4 * fy, qcost = 1 << 30, qsatd = 0;
        uint8_t p8[64];
        for (int step = 2; step >= 1; step--) {
            int cx = qx, cy = qy;
            for (int i = (step == 2 ? -1 : 0); i < 9; i++) {
                int dx, dy;
                if (i < 0) { dx = dy = 0; }
                else { if (i == 4) continue; dx = (i % 3 - 1) * step; dy = (i / 3 - 1) * step; }
                int mx = cx + dx, my = cy + dy;
                ref_mc_luma(p8, 8, r8, sy, mx, my, 8, 8);
                int satd = ref_satd(c8, sy, p8, 8, 8, 8);
                int c = satd + lambda * (ref_mv_bits(mx - 4 * pm[0]) + ref_mv_bits(my - 4 * pm[1]));
                if (c < qcost) { qcost = c; qsatd = satd; qx = mx; qy = my; }
            }
        }
        mi->me_mv8[p][0] = (int16_t)qx; mi->me_mv8[p][1] = (int16_t)qy; mi->me_satd8[p] = qsatd;
        cost8 += qcost;
    }
    mi->me_cost8 = cost8;
    mi->me_part = (uint8_t)(cost8 < bcost);
}

/* ---- headers ------------------------------------------------------------------------ */
static int level_idc(const ref_enc *e) {
    /* Table A-1 {level, MaxMBPS, MaxFS, MaxBR (x1250 bit/s, High profile)}; the level covers 30 Mbit/s, the largest bitrate
     * ceracoder may ask for at run time (/root/reference/src/core/bitrate_control.h:31) */
    static const int tab[][4] = {{30, 40500, 1620, 10000},    {31, 108000, 3600, 14000},   {32, 216000, 5120, 20000},
                                 {40, 245760, 8192, 20000},   {41, 245760, 8192, 50000},   {42, 522240, 8704, 50000},
                                 {50, 589824, 22080, 135000}, {51, 983040, 36864, 240000}, {52, 2073600, 36864, 240000}};
    int fd = e->fps_d ? e->fps_d : 1;
    long fs = (long)e->mb_w * e->mb_h, mbps = (fs * e->fps_n + fd - 1) / fd;
    for (int i = 0; i < 9; i++) if (fs <= tab[i][2] && mbps <= tab[i][1] && 30000000L <= 1250L * tab[i][3]) return tab[i][0];
    return 52;
}
static int write_sps(const ref_enc *e, uint8_t *o) {
    int bp = 0, cr = (e->mb_w * 16 - e->w) / 2, cb = (e->mb_h * 16 - e->h) / 2;
    ref_put_bits(o, &bp, 8, 100); ref_put_bits(o, &bp, 8, 0); ref_put_bits(o, &bp, 8, (uint32_t)level_idc(e)); /* High profile */
    ref_put_ue(o, &bp, 0);
    ref_put_ue(o, &bp, 1); ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, 0); ref_put_bits(o, &bp, 1, 0); ref_put_bits(o, &bp, 1, 0); /* 4:2:0, 8 bit, no bypass, no scaling matrices */ ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, 2); ref_put_ue(o, &bp, 1); ref_put_bits(o, &bp, 1, 0);
    ref_put_ue(o, &bp, (unsigned)e->mb_w - 1); ref_put_ue(o, &bp, (unsigned)e->mb_h - 1);
    ref_put_bits(o, &bp, 1, 1); ref_put_bits(o, &bp, 1, 1);
    ref_put_bits(o, &bp, 1, (cr || cb));
    if (cr || cb) { ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, (unsigned)cr); ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, (unsigned)cb); }
    ref_put_bits(o, &bp, 1, 1); /* vui_parameters_present_flag */
    ref_put_bits(o, &bp, 4, 0); /* aspect, overscan, video_signal, chroma_loc */
    ref_put_bits(o, &bp, 1, 1); /* timing_info_present_flag */
    ref_put_bits(o, &bp, 16, (uint32_t)e->fps_d >> 16); ref_put_bits(o, &bp, 16, (uint32_t)e->fps_d & 0xffff);
    ref_put_bits(o, &bp, 16, (uint32_t)(2 * e->fps_n) >> 16); ref_put_bits(o, &bp, 16, (uint32_t)(2 * e->fps_n) & 0xffff);
    ref_put_bits(o, &bp, 1, 1); /* fixed_frame_rate_flag */
    ref_put_bits(o, &bp, 3, 0); /* nal_hrd, vcl_hrd, pic_struct */
    ref_put_bits(o, &bp, 1, 1); /* bitstream_restriction_flag */
    ref_put_bits(o, &bp, 1, 1); ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, 9); ref_put_ue(o, &bp, 9);
    ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, 1);
    ref_put_bits(o, &bp, 1, 1);
    while (bp & 7) ref_put_bits(o, &bp, 1, 0);
    return bp >> 3;
}
static int write_pps(uint8_t *o) {
    int bp = 0;
    ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, 0); ref_put_bits(o, &bp, 1, 0); ref_put_bits(o, &bp, 1, 0);
    ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, 0); ref_put_ue(o, &bp, 0); ref_put_bits(o, &bp, 1, 0); ref_put_bits(o, &bp, 2, 0);
    ref_put_se(o, &bp, 0); ref_put_se(o, &bp, 0); ref_put_se(o, &bp, 0);
    ref_put_bits(o, &bp, 1, 1); ref_put_bits(o, &bp, 1, 0); ref_put_bits(o, &bp, 1, 0);
    ref_put_bits(o, &bp, 1, 1); ref_put_bits(o, &bp, 1, 0); ref_put_se(o, &bp, 0); /* transform_8x8_mode_flag, no scaling lists, second_chroma_qp_index_offset */
    ref_put_bits(o, &bp, 1, 1);
    while (bp & 7) ref_put_bits(o, &bp, 1, 0);
    return bp >> 3;
}
static void write_slice_header(const ref_enc *e, uint8_t *o, int *bp, int first_mb) {
    ref_put_ue(o, bp, (unsigned)first_mb);
    ref_put_ue(o, bp, e->is_idr ? 7 : 5);
    ref_put_ue(o, bp, 0);
    ref_put_bits(o, bp, 4, (uint32_t)e->frame_num & 15);
    if (e->is_idr) ref_put_ue(o, bp, (unsigned)e->idr_pic_id);
    if (!e->is_idr) { ref_put_bits(o, bp, 1, 0); ref_put_bits(o, bp, 1, 0); } /* override, list modification */
    if (e->is_idr) ref_put_bits(o, bp, 2, 0); else ref_put_bits(o, bp, 1, 0); /* dec_ref_pic_marking */
    ref_put_se(o, bp, e->qp - 26);
    ref_put_ue(o, bp, 2); ref_put_se(o, bp, 0); ref_put_se(o, bp, 0); /* deblocking idc=2, offsets 0 */
}
/* Annex B: start code + header + emulation prevention (7.4.1) */
static int write_nal(uint8_t *out, int hdr, const uint8_t *rbsp, int n) {
    int k = 0, zeros = 0;
    out[k++] = 0; out[k++] = 0; out[k++] = 0; out[k++] = 1; out[k++] = (uint8_t)hdr;
    for (int i = 0; i < n; i++) {
        if (zeros >= 2 && rbsp[i] <= 3) { out[k++] = 3; zeros = 0; }
        out[k++] = rbsp[i];
        zeros = rbsp[i] == 0 ? zeros + 1 : 0;
    }
    return k;
}

/* ---- frame ---------------------------------------------------------------------------- */
int ref_enc_encode(ref_enc *e, const uint8_t *y, int sy_, const uint8_t *u, int su, const uint8_t *v,
                   int sv, int is_idr, int qp, uint8_t *out, int out_cap) {
    const int W = e->mb_w * 16, H = e->mb_h * 16, mw = e->mb_w, mh = e->mb_h;
    if (qp < 10 || qp > 51) return -1;
    if (!e->have_ref) is_idr = 1;
    e->qp = qp; e->is_idr = is_idr;
    if (is_idr) { e->frame_num = 0; }
    load_plane(e->cur[0], e->sy, y, sy_, e->w, e->h, W, H);
    load_plane(e->cur[1], e->sc, u, su, e->w / 2, e->h / 2, W / 2, H / 2);
    load_plane(e->cur[2], e->sc, v, sv, e->w / 2, e->h / 2, W / 2, H / 2);

    if (!is_idr) {
        /* l1ref: the half-resolution picture of the previous frame's UNFILTERED reconstruction, written in that frame's slice
         * loop below; its border is edge replication of the half-resolution samples */
        downsample2(e->l1cur, e->s1, e->cur[0], e->sy, W, H, 0);
        extend_border(e->l1ref, e->s1, W / 2, H / 2, REF_PAD / 2);
#pragma omp parallel for schedule(dynamic, 4)
        for (int i = 0; i < mw * mh; i++) me_l1(e, i % mw, i / mw);
#pragma omp parallel for schedule(dynamic, 4)
        for (int i = 0; i < mw * mh; i++) me_l0(e, i % mw, i / mw);
    }

    int ns = (mh + e->rps - 1) / e->rps;
#pragma omp parallel for schedule(dynamic, 1)
    for (int s = 0; s < ns; s++) {
        int r0 = s * e->rps, r1 = r0 + e->rps < mh ? r0 + e->rps : mh;
        for (int my = r0; my < r1; my++)
            for (int mx = 0; mx < mw; mx++) { if (e->effort) ref_mb_encode_rd(e, mx, my); else ref_mb_encode(e, mx, my); }
        /* bitstream before deblocking (needs nothing from rec) */
        uint8_t *sb = e->slice_buf + (size_t)s * e->slice_cap;
        int bp = 0, skip_run = 0;
        write_slice_header(e, sb, &bp, r0 * mw);
        for (int my = r0; my < r1; my++)
            for (int mx = 0; mx < mw; mx++) ref_mb_write(e, mx, my, sb, &bp, &skip_run);
        if (skip_run) ref_put_ue(sb, &bp, (unsigned)skip_run);
        ref_put_bits(sb, &bp, 1, 1);
        while (bp & 7) ref_put_bits(sb, &bp, 1, 0);
        e->slice_bits[s] = bp;
        /* half-resolution reference of the next frame's coarse search: 2x2 box means of this slice's reconstruction BEFORE the
         * in-loop filter (the product writes them where it publishes the reconstruction, so that the coarse search does not have
         * to wait for the filter); the motion search of this frame, which read the plane, is over */
        for (int y = r0 * 8; y < r1 * 8; y++)
            for (int x = 0; x < W / 2; x++) {
                const uint8_t *p = e->rec[0] + (size_t)(2 * y) * e->sy + 2 * x;
                e->l1ref[(size_t)y * e->s1 + x] = (uint8_t)((p[0] + p[1] + p[e->sy] + p[e->sy + 1] + 2) >> 2);
            }
        for (int my = r0; my < r1; my++)
            for (int mx = 0; mx < mw; mx++) ref_mb_deblock(e, mx, my);
    }
    for (int p = 0; p < 3; p++) extend_border(e->rec[p], p ? e->sc : e->sy, p ? W / 2 : W, p ? H / 2 : H, p ? REF_PAD / 2 : REF_PAD);

    /* access unit */
    int k = 0;
    uint8_t hdr[64];
    if (is_idr) {
        if (out_cap < 256) return -2;
        int n = write_sps(e, hdr); k += write_nal(out + k, 0x67, hdr, n);
        n = write_pps(hdr); k += write_nal(out + k, 0x68, hdr, n);
    }
    for (int s = 0; s < ns; s++) {
        int n = e->slice_bits[s] >> 3;
        if (k + n + n / 2 + 16 > out_cap) return -2;
        k += write_nal(out + k, is_idr ? 0x65 : 0x41, e->slice_buf + (size_t)s * e->slice_cap, n);
    }
    for (int p = 0; p < 3; p++) { uint8_t *t = e->rec[p]; e->rec[p] = e->ref[p]; e->ref[p] = t; }
    e->have_ref = 1;
    e->frame_num = (e->frame_num + 1) & 15;
    if (is_idr) e->idr_pic_id = (e->idr_pic_id + 1) & 0xffff;
    return k;
}
