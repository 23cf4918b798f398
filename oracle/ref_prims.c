/*
 * oracle/ref_prims.c — CPU restatement of the H.264 encoder primitives.
 * TEST INFRASTRUCTURE ONLY (see ref_h264.h).
 *
 * Reference anchor: the hot path of CERALIVE/ceracoder is the external x264enc
 * element (/root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6);
 * there is no encoder source under /root/reference.  Non-normative primitives
 * follow the published libx264 C forms [EXT] (pixel_sad_WxH, pixel_satd_WxH,
 * sub4x4_dct, dct4x4dc, quant_4x4, x264_decimate_score); normative ones follow
 * ITU-T H.264 (clause cited at each function).
 */
#include "ref_h264.h"
#include "ref_tables.h"
#include <stdlib.h>
#include <string.h>

static inline int clip255(int x) { return x < 0 ? 0 : (x > 255 ? 255 : x); }
static inline int iabs(int x) { return x < 0 ? -x : x; }

/* [EXT] x264 pixel_sad_WxH: sum of absolute differences */
int ref_sad(const uint8_t *a, int sa, const uint8_t *b, int sb, int w, int h) {
    int s = 0;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) s += iabs(a[y * sa + x] - b[y * sb + x]);
    return s;
}

/* [EXT] x264 pixel_satd_4x4: 4x4 Hadamard of the difference, sum |coef| >> 1 */
int ref_satd_4x4(const uint8_t *a, int sa, const uint8_t *b, int sb) {
    int d[16], t[16], s = 0;
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++) d[y * 4 + x] = a[y * sa + x] - b[y * sb + x];
    for (int i = 0; i < 4; i++) { /* rows */
        int s01 = d[i * 4 + 0] + d[i * 4 + 1], d01 = d[i * 4 + 0] - d[i * 4 + 1];
        int s23 = d[i * 4 + 2] + d[i * 4 + 3], d23 = d[i * 4 + 2] - d[i * 4 + 3];
        t[i * 4 + 0] = s01 + s23; t[i * 4 + 1] = s01 - s23;
        t[i * 4 + 2] = d01 - d23; t[i * 4 + 3] = d01 + d23;
    }
    for (int i = 0; i < 4; i++) { /* columns */
        int s01 = t[0 + i] + t[4 + i], d01 = t[0 + i] - t[4 + i];
        int s23 = t[8 + i] + t[12 + i], d23 = t[8 + i] - t[12 + i];
        s += iabs(s01 + s23) + iabs(s01 - s23) + iabs(d01 - d23) + iabs(d01 + d23);
    }
    return s >> 1;
}

int ref_satd(const uint8_t *a, int sa, const uint8_t *b, int sb, int w, int h) {
    int s = 0;
    for (int y = 0; y < h; y += 4)
        for (int x = 0; x < w; x += 4) s += ref_satd_4x4(a + y * sa + x, sa, b + y * sb + x, sb);
    return s;
}

/* [EXT] x264 sub4x4_dct: forward 4x4 core transform of (cur - pred); H.264 8.5.12 inverse's
 * forward counterpart Cf = [1 1 1 1; 2 1 -1 -2; 1 -1 -1 1; 1 -2 2 -1]; output raster [y*4+x] */
void ref_sub4x4_dct(int16_t dct[16], const uint8_t *cur, int sc, const uint8_t *pred, int sp) {
    int d[16], t[16];
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++) d[y * 4 + x] = cur[y * sc + x] - pred[y * sp + x];
    for (int i = 0; i < 4; i++) { /* horizontal: each row */
        int s03 = d[i * 4 + 0] + d[i * 4 + 3], s12 = d[i * 4 + 1] + d[i * 4 + 2];
        int d03 = d[i * 4 + 0] - d[i * 4 + 3], d12 = d[i * 4 + 1] - d[i * 4 + 2];
        t[i * 4 + 0] = s03 + s12; t[i * 4 + 1] = 2 * d03 + d12;
        t[i * 4 + 2] = s03 - s12; t[i * 4 + 3] = d03 - 2 * d12;
    }
    for (int i = 0; i < 4; i++) { /* vertical: each column */
        int s03 = t[0 + i] + t[12 + i], s12 = t[4 + i] + t[8 + i];
        int d03 = t[0 + i] - t[12 + i], d12 = t[4 + i] - t[8 + i];
        dct[0 + i] = (int16_t)(s03 + s12); dct[4 + i] = (int16_t)(2 * d03 + d12);
        dct[8 + i] = (int16_t)(s03 - s12); dct[12 + i] = (int16_t)(d03 - 2 * d12);
    }
}

static void hadamard4x4(int16_t d[16], int round_half) {
    /* conventional raster layout d[vfreq*4+hfreq] (x264 keeps the same values transposed) */
    int t[16];
    for (int i = 0; i < 4; i++) { /* rows */
        int s01 = d[i * 4 + 0] + d[i * 4 + 1], d01 = d[i * 4 + 0] - d[i * 4 + 1];
        int s23 = d[i * 4 + 2] + d[i * 4 + 3], d23 = d[i * 4 + 2] - d[i * 4 + 3];
        t[i * 4 + 0] = s01 + s23; t[i * 4 + 1] = s01 - s23;
        t[i * 4 + 2] = d01 - d23; t[i * 4 + 3] = d01 + d23;
    }
    for (int i = 0; i < 4; i++) { /* columns */
        int s01 = t[0 + i] + t[4 + i], d01 = t[0 + i] - t[4 + i];
        int s23 = t[8 + i] + t[12 + i], d23 = t[8 + i] - t[12 + i];
        int r = round_half ? 1 : 0;
        d[0 + i] = (int16_t)((s01 + s23 + r) >> r); d[4 + i] = (int16_t)((s01 - s23 + r) >> r);
        d[8 + i] = (int16_t)((d01 - d23 + r) >> r); d[12 + i] = (int16_t)((d01 + d23 + r) >> r);
    }
}
/* [EXT] x264 dct4x4dc: 4x4 Hadamard of the 16 luma DCs with (x+1)>>1 */
void ref_dct4x4dc(int16_t d[16]) { hadamard4x4(d, 1); }
/* H.264 8.5.10 inverse DC transform (no scaling); x264 idct4x4dc */
void ref_idct4x4dc(int16_t d[16]) { hadamard4x4(d, 0); }
/* H.264 8.5.11.1 / x264 dct2x2dc: 2x2 Hadamard, raster order */
void ref_dct2x2dc(int16_t d[4]) {
    int a = d[0] + d[1], b = d[0] - d[1], c = d[2] + d[3], e = d[2] - d[3];
    d[0] = (int16_t)(a + c); d[1] = (int16_t)(b + e); d[2] = (int16_t)(a - c); d[3] = (int16_t)(b - e);
}

static inline int pos_class(int i) { /* 0: (even,even) 1: (odd,odd) 2: mixed */
    int x = i & 3, y = i >> 2;
    return (x & 1) && (y & 1) ? 1 : (!(x & 1) && !(y & 1) ? 0 : 2);
}

/* [EXT] x264 cqm init (flat matrices): quant4_mf / quant4_bias for one qp.
 * deadzone: luma intra 32-11 = 21, luma inter 32-21 = 11 (x264 defaults) */
void ref_quant_tables(int qp, int intra, uint16_t mf[16], uint16_t bias[16]) {
    int dz = intra ? 21 : 11, s = qp / 6 - 1;
    for (int i = 0; i < 16; i++) {
        int m = ref_quant_mf[qp % 6][pos_class(i)];
        int j = s <= 0 ? m << -s : (m + (1 << (s - 1))) >> s;
        int b = ((dz << 10) + (j >> 1)) / j, cap = (1 << 15) / j;
        mf[i] = (uint16_t)j;
        bias[i] = (uint16_t)(b < cap ? b : cap);
    }
}

/* [EXT] x264 quant_4x4 (QUANT_ONE); additionally clamps |level| to 2047 so that the
 * CAVLC level escape (level_prefix <= 15, Baseline) is always representable */
static inline int quant_one(int c, int mf, int f) {
    int l = c > 0 ? ((f + c) * mf) >> 16 : -(((f - c) * mf) >> 16);
    return l > 2047 ? 2047 : (l < -2047 ? -2047 : l);
}
int ref_quant_4x4(int16_t dct[16], const uint16_t mf[16], const uint16_t bias[16]) {
    int nz = 0;
    for (int i = 0; i < 16; i++) { dct[i] = (int16_t)quant_one(dct[i], mf[i], bias[i]); nz |= dct[i]; }
    return nz != 0;
}
/* [EXT] x264 quant_4x4_dc / quant_2x2_dc: caller passes mf = quant4_mf[qp][0]>>1, bias<<1 */
int ref_quant_dc(int16_t *dct, int n, int mf, int bias) {
    int nz = 0;
    for (int i = 0; i < n; i++) { dct[i] = (int16_t)quant_one(dct[i], mf, bias); nz |= dct[i]; }
    return nz != 0;
}

/* H.264 8.5.12.1 scaling of residual 4x4 (flat scaling list 16): LevelScale = 16*v */
void ref_dequant_4x4(int16_t dct[16], int qp) {
    int qb = qp / 6 - 4;
    for (int i = 0; i < 16; i++) {
        int ls = ref_dequant_v[qp % 6][pos_class(i)] * 16;
        dct[i] = (int16_t)(qb >= 0 ? (dct[i] * ls) << qb : (dct[i] * ls + (1 << (-qb - 1))) >> -qb);
    }
}
/* H.264 8.5.10 scaling of Intra16x16 DC */
void ref_dequant_4x4_dc(int16_t dct[16], int qp) {
    int qb = qp / 6 - 6, ls = ref_dequant_v[qp % 6][0] * 16;
    for (int i = 0; i < 16; i++)
        dct[i] = (int16_t)(qb >= 0 ? (dct[i] * ls) << qb : (dct[i] * ls + (1 << (-qb - 1))) >> -qb);
}
/* H.264 8.5.11.2 chroma DC: dcC = ((f * LevelScale(qp%6,0,0)) << (qp/6)) >> 5, f = 2x2 Hadamard */
void ref_dequant_2x2_dc(int16_t out[4], const int16_t in[4], int qp) {
    int16_t t[4] = {in[0], in[1], in[2], in[3]};
    int dmf = (ref_dequant_v[qp % 6][0] * 16) << (qp / 6);
    ref_dct2x2dc(t);
    for (int i = 0; i < 4; i++) out[i] = (int16_t)((t[i] * dmf) >> 5);
}

/* H.264 8.5.12.2 inverse 4x4 transform + 8.5.14 reconstruction: dst = clip(pred + ((r+32)>>6)) */
void ref_add4x4_idct(uint8_t *dst, int sd, const uint8_t *pred, int sp, const int16_t dct[16]) {
    int t[16], r[16];
    for (int i = 0; i < 4; i++) { /* rows */
        int e0 = dct[i * 4 + 0] + dct[i * 4 + 2], e1 = dct[i * 4 + 0] - dct[i * 4 + 2];
        int e2 = (dct[i * 4 + 1] >> 1) - dct[i * 4 + 3], e3 = dct[i * 4 + 1] + (dct[i * 4 + 3] >> 1);
        t[i * 4 + 0] = e0 + e3; t[i * 4 + 1] = e1 + e2; t[i * 4 + 2] = e1 - e2; t[i * 4 + 3] = e0 - e3;
    }
    for (int i = 0; i < 4; i++) { /* columns */
        int e0 = t[0 + i] + t[8 + i], e1 = t[0 + i] - t[8 + i];
        int e2 = (t[4 + i] >> 1) - t[12 + i], e3 = t[4 + i] + (t[12 + i] >> 1);
        r[0 + i] = e0 + e3; r[4 + i] = e1 + e2; r[8 + i] = e1 - e2; r[12 + i] = e0 - e3;
    }
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++) dst[y * sd + x] = (uint8_t)clip255(pred[y * sp + x] + ((r[y * 4 + x] + 32) >> 6));
}

/* [EXT] x264_decimate_score15/16: zz = zigzag-ordered levels, n coefficients */
int ref_decimate_score(const int16_t *zz, int n) {
    static const uint8_t ds[16] = {3, 2, 2, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    int idx = n - 1, score = 0;
    while (idx >= 0 && zz[idx] == 0) idx--;
    while (idx >= 0) {
        int run = 0;
        if (iabs(zz[idx--]) > 1) return 9;
        while (idx >= 0 && zz[idx] == 0) { idx--; run++; }
        score += ds[run];
    }
    return score;
}

/* H.264 8.3.1.2 Intra_4x4 prediction. top[-1] = top-left, top[0..7] (caller replicates top[3]
 * into top[4..7] when the top-right block is unavailable), left[0..3] */
void ref_pred4x4(uint8_t dst[16], int mode, const uint8_t *top, const uint8_t *left, int has_top,
                 int has_left) {
#define T(i) ((i) < 0 ? lt : top[i])
#define L(i) ((i) < 0 ? lt : left[i])
    int lt = top[-1];
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++) {
            int v = 128;
            switch (mode) {
            case 0: v = top[x]; break;
            case 1: v = left[y]; break;
            case 2:
                if (has_top && has_left) v = (top[0] + top[1] + top[2] + top[3] + left[0] + left[1] + left[2] + left[3] + 4) >> 3;
                else if (has_top) v = (top[0] + top[1] + top[2] + top[3] + 2) >> 2;
                else if (has_left) v = (left[0] + left[1] + left[2] + left[3] + 2) >> 2;
                else v = 128;
                break;
            case 3:
                v = (x == 3 && y == 3) ? (top[6] + 3 * top[7] + 2) >> 2
                                       : (top[x + y] + 2 * top[x + y + 1] + top[x + y + 2] + 2) >> 2;
                break;
            case 4:
                if (x > y) v = (T(x - y - 2) + 2 * T(x - y - 1) + T(x - y) + 2) >> 2;
                else if (x < y) v = (L(y - x - 2) + 2 * L(y - x - 1) + L(y - x) + 2) >> 2;
                else v = (top[0] + 2 * lt + left[0] + 2) >> 2;
                break;
            case 5: {
                int z = 2 * x - y, i = x - (y >> 1);
                if (z >= 0 && !(z & 1)) v = (T(i - 1) + T(i) + 1) >> 1;
                else if (z >= 0) v = (T(i - 2) + 2 * T(i - 1) + T(i) + 2) >> 2;
                else if (z == -1) v = (left[0] + 2 * lt + top[0] + 2) >> 2;
                else v = (L(y - 1) + 2 * L(y - 2) + L(y - 3) + 2) >> 2;
                break;
            }
            case 6: {
                int z = 2 * y - x, i = y - (x >> 1);
                if (z >= 0 && !(z & 1)) v = (L(i - 1) + L(i) + 1) >> 1;
                else if (z >= 0) v = (L(i - 2) + 2 * L(i - 1) + L(i) + 2) >> 2;
                else if (z == -1) v = (left[0] + 2 * lt + top[0] + 2) >> 2;
                else v = (T(x - 1) + 2 * T(x - 2) + T(x - 3) + 2) >> 2;
                break;
            }
            case 7: {
                int i = x + (y >> 1);
                v = (y & 1) ? (top[i] + 2 * top[i + 1] + top[i + 2] + 2) >> 2 : (top[i] + top[i + 1] + 1) >> 1;
                break;
            }
            case 8: {
                int z = x + 2 * y, i = y + (x >> 1);
                if (z > 5) v = left[3];
                else if (z == 5) v = (left[2] + 3 * left[3] + 2) >> 2;
                else if (z & 1) v = (left[i] + 2 * left[i + 1] + left[i + 2] + 2) >> 2;
                else v = (left[i] + left[i + 1] + 1) >> 1;
                break;
            }
            }
            dst[y * 4 + x] = (uint8_t)v;
        }
#undef T
#undef L
}

/* H.264 8.3.3 Intra_16x16 prediction; rec points at the MB's top-left sample in the
 * reconstruction plane (neighbours at rec[-sr+x], rec[y*sr-1]). modes 0 V 1 H 2 DC 3 Plane */
void ref_pred16x16(uint8_t *dst, int sd, int mode, const uint8_t *rec, int sr, int has_top,
                   int has_left) {
    int dc = 128, a = 0, b = 0, c = 0;
    if (mode == 2) {
        int st = 0, sl = 0;
        for (int i = 0; i < 16; i++) { if (has_top) st += rec[-sr + i]; if (has_left) sl += rec[i * sr - 1]; }
        if (has_top && has_left) dc = (st + sl + 16) >> 5;
        else if (has_top) dc = (st + 8) >> 4;
        else if (has_left) dc = (sl + 8) >> 4;
    } else if (mode == 3) {
        int H = 0, V = 0;
        for (int i = 0; i < 8; i++) {
            H += (i + 1) * (rec[-sr + 8 + i] - rec[-sr + 6 - i]);
            V += (i + 1) * (rec[(8 + i) * sr - 1] - rec[(6 - i) * sr - 1]);
        }
        a = 16 * (rec[15 * sr - 1] + rec[-sr + 15]); b = (5 * H + 32) >> 6; c = (5 * V + 32) >> 6;
    }
    for (int y = 0; y < 16; y++)
        for (int x = 0; x < 16; x++) {
            int v;
            if (mode == 0) v = rec[-sr + x];
            else if (mode == 1) v = rec[y * sr - 1];
            else if (mode == 2) v = dc;
            else v = clip255((a + b * (x - 7) + c * (y - 7) + 16) >> 5);
            dst[y * sd + x] = (uint8_t)v;
        }
}

/* H.264 8.3.4 chroma intra prediction (8x8, 4:2:0). modes 0 DC 1 H 2 V 3 Plane */
void ref_pred8x8c(uint8_t *dst, int sd, int mode, const uint8_t *rec, int sr, int has_top,
                  int has_left) {
    if (mode == 0) {
        int st[2] = {0, 0}, sl[2] = {0, 0};
        for (int i = 0; i < 4; i++) {
            if (has_top) { st[0] += rec[-sr + i]; st[1] += rec[-sr + 4 + i]; }
            if (has_left) { sl[0] += rec[i * sr - 1]; sl[1] += rec[(4 + i) * sr - 1]; }
        }
        for (int by = 0; by < 2; by++)
            for (int bx = 0; bx < 2; bx++) {
                int dc = 128;
                if (bx == by) {
                    if (has_top && has_left) dc = (st[bx] + sl[by] + 4) >> 3;
                    else if (has_top) dc = (st[bx] + 2) >> 2;
                    else if (has_left) dc = (sl[by] + 2) >> 2;
                } else if (bx == 1) { /* top-right: prefer top */
                    if (has_top) dc = (st[1] + 2) >> 2; else if (has_left) dc = (sl[0] + 2) >> 2;
                } else { /* bottom-left: prefer left */
                    if (has_left) dc = (sl[1] + 2) >> 2; else if (has_top) dc = (st[0] + 2) >> 2;
                }
                for (int y = 0; y < 4; y++)
                    for (int x = 0; x < 4; x++) dst[(by * 4 + y) * sd + bx * 4 + x] = (uint8_t)dc;
            }
        return;
    }
    int a = 0, b = 0, c = 0;
    if (mode == 3) {
        int H = 0, V = 0;
        for (int i = 0; i < 4; i++) {
            H += (i + 1) * (rec[-sr + 4 + i] - rec[-sr + 2 - i]);
            V += (i + 1) * (rec[(4 + i) * sr - 1] - rec[(2 - i) * sr - 1]);
        }
        a = 16 * (rec[7 * sr - 1] + rec[-sr + 7]); b = (34 * H + 32) >> 6; c = (34 * V + 32) >> 6;
    }
    for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++) {
            int v;
            if (mode == 1) v = rec[y * sr - 1];
            else if (mode == 2) v = rec[-sr + x];
            else v = clip255((a + b * (x - 3) + c * (y - 3) + 16) >> 5);
            dst[y * sd + x] = (uint8_t)v;
        }
}

/* H.264 8.4.2.2.1 luma sample interpolation. ref points at the block's integer position
 * (0,0) in a border-extended plane; mvx,mvy in quarter-pel units */
static inline int tap6(int a, int b, int c, int d, int e, int f) { return a - 5 * b + 20 * c + 20 * d - 5 * e + f; }
static inline int hpel_h(const uint8_t *p) { return clip255((tap6(p[-2], p[-1], p[0], p[1], p[2], p[3]) + 16) >> 5); }
static inline int hpel_v(const uint8_t *p, int s) {
    return clip255((tap6(p[-2 * s], p[-s], p[0], p[s], p[2 * s], p[3 * s]) + 16) >> 5);
}
static inline int hpel_c(const uint8_t *p, int s) { /* centre 'j' from unrounded horizontal intermediates */
    int t[6];
    for (int i = 0; i < 6; i++) { const uint8_t *q = p + (i - 2) * s; t[i] = tap6(q[-2], q[-1], q[0], q[1], q[2], q[3]); }
    return clip255((tap6(t[0], t[1], t[2], t[3], t[4], t[5]) + 512) >> 10);
}
void ref_mc_luma(uint8_t *dst, int sd, const uint8_t *ref, int sr, int mvx, int mvy, int w, int h) {
    int fx = mvx & 3, fy = mvy & 3;
    const uint8_t *base = ref + (mvy >> 2) * sr + (mvx >> 2);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            const uint8_t *p = base + y * sr + x;
            int G = p[0], v;
            switch (fy * 4 + fx) {
            case 0: v = G; break;
            case 1: v = (G + hpel_h(p) + 1) >> 1; break;                         /* a */
            case 2: v = hpel_h(p); break;                                       /* b */
            case 3: v = (p[1] + hpel_h(p) + 1) >> 1; break;                      /* c */
            case 4: v = (G + hpel_v(p, sr) + 1) >> 1; break;                     /* d */
            case 5: v = (hpel_h(p) + hpel_v(p, sr) + 1) >> 1; break;             /* e */
            case 6: v = (hpel_h(p) + hpel_c(p, sr) + 1) >> 1; break;             /* f */
            case 7: v = (hpel_h(p) + hpel_v(p + 1, sr) + 1) >> 1; break;         /* g */
            case 8: v = hpel_v(p, sr); break;                                   /* h */
            case 9: v = (hpel_v(p, sr) + hpel_c(p, sr) + 1) >> 1; break;         /* i */
            case 10: v = hpel_c(p, sr); break;                                  /* j */
            case 11: v = (hpel_c(p, sr) + hpel_v(p + 1, sr) + 1) >> 1; break;    /* k */
            case 12: v = (p[sr] + hpel_v(p, sr) + 1) >> 1; break;                /* n */
            case 13: v = (hpel_v(p, sr) + hpel_h(p + sr) + 1) >> 1; break;       /* p */
            case 14: v = (hpel_c(p, sr) + hpel_h(p + sr) + 1) >> 1; break;       /* q */
            default: v = (hpel_v(p + 1, sr) + hpel_h(p + sr) + 1) >> 1; break;   /* r */
            }
            dst[y * sd + x] = (uint8_t)v;
        }
}

/* H.264 8.4.2.2.2 chroma sample interpolation; mvx,mvy = luma quarter-pel mv (1/8 chroma px) */
void ref_mc_chroma(uint8_t *dst, int sd, const uint8_t *ref, int sr, int mvx, int mvy, int w, int h) {
    int fx = mvx & 7, fy = mvy & 7;
    const uint8_t *base = ref + (mvy >> 3) * sr + (mvx >> 3);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            const uint8_t *p = base + y * sr + x;
            dst[y * sd + x] = (uint8_t)(((8 - fx) * (8 - fy) * p[0] + fx * (8 - fy) * p[1] +
                                         (8 - fx) * fy * p[sr] + fx * fy * p[sr + 1] + 32) >> 6);
        }
}

/* bit length of se(v) (H.264 9.1) */
int ref_mv_bits(int d) {
    unsigned k = d > 0 ? 2u * (unsigned)d : 2u * (unsigned)(-d) + 1; /* codeNum+1 */
    if (d == 0) k = 1;
    int n = 0;
    while (k >> (n + 1)) n++;
    return 2 * n + 1;
}
int ref_median3(int a, int b, int c) {
    int mn = a < b ? a : b, mx = a < b ? b : a;
    return c < mn ? mn : (c > mx ? mx : c);
}

/* ---- 8x8 transform path (High profile: transform_size_8x8_flag) ----------------------------------- */
/* [EXT] x264 sub8x8_dct8 1-D butterfly (the forward form matching H.264's 8x8 inverse) */
static void dct8_1d(const int *s, int st, int *d, int dt) {
    const int a0 = s[0] + s[7 * st], a1 = s[st] + s[6 * st], a2 = s[2 * st] + s[5 * st], a3 = s[3 * st] + s[4 * st];
    const int a4 = s[0] - s[7 * st], a5 = s[st] - s[6 * st], a6 = s[2 * st] - s[5 * st], a7 = s[3 * st] - s[4 * st];
    const int b0 = a0 + a3, b1 = a1 + a2, b2 = a0 - a3, b3 = a1 - a2;
    const int b4 = a5 + a6 + ((a4 >> 1) + a4), b5 = a4 - a7 - ((a6 >> 1) + a6);
    const int b6 = a4 + a7 - ((a5 >> 1) + a5), b7 = a5 - a6 + ((a7 >> 1) + a7);
    d[0] = b0 + b1; d[dt] = b4 + (b7 >> 2); d[2 * dt] = b2 + (b3 >> 1); d[3 * dt] = b5 + (b6 >> 2);
    d[4 * dt] = b0 - b1; d[5 * dt] = b6 - (b5 >> 2); d[6 * dt] = (b2 >> 1) - b3; d[7 * dt] = (b4 >> 2) - b7;
}
/* residual -> 8x8 forward transform: rows first, then columns; dct raster (y*8+x) */
void ref_sub8x8_dct8(int16_t dct[64], const uint8_t *cur, int sc, const uint8_t *pred, int sp) {
    int d[64], t[64];
    for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++) d[y * 8 + x] = cur[y * sc + x] - pred[y * sp + x];
    for (int y = 0; y < 8; y++) dct8_1d(d + y * 8, 1, t + y * 8, 1);
    for (int x = 0; x < 8; x++) dct8_1d(t + x, 8, d + x, 8);
    for (int i = 0; i < 64; i++) dct[i] = (int16_t)d[i];
}
/* quantise in place (raster); level = sign * ((|c| + bias) * mf >> 16), mf = quant8_mf >> (qp/6) [EXT: x264 quant_8x8] */
int ref_quant_8x8(int16_t dct[64], int qp, int intra) {
    const int dz = intra ? 21 : 11, s = qp / 6;
    int nz = 0;
    for (int i = 0; i < 64; i++) {
        const int m = ref_quant8_mf[qp % 6][ref_cls8[((i >> 3) & 3) * 4 + (i & 3)]];
        const int j = s <= 0 ? m : (m + (1 << (s - 1))) >> s;
        int b = ((dz << 10) + (j >> 1)) / j;
        const int cap = (1 << 15) / j;
        if (b > cap) b = cap;
        dct[i] = (int16_t)quant_one(dct[i], j, b);
        nz |= dct[i];
    }
    return nz != 0;
}
/* H.264 8.5.13 (scaling of 8x8 blocks, flat matrix): d = (c * 16 * v) << (qp/6 - 6), rounded right shift below qp 36 */
void ref_dequant_8x8(int16_t dct[64], int qp) {
    const int qb = qp / 6 - 6;
    for (int i = 0; i < 64; i++) {
        const int ls = ref_dequant8_v[qp % 6][ref_cls8[((i >> 3) & 3) * 4 + (i & 3)]] * 16;
        dct[i] = (int16_t)(qb >= 0 ? (dct[i] * ls) << qb : (dct[i] * ls + (1 << (-qb - 1))) >> -qb);
    }
}
/* H.264 8.5.13 eq 8-338..8-354: one 1-D inverse pass */
static void idct8_1d(const int *s, int st, int *d, int dt) {
    const int a0 = s[0] + s[4 * st], a2 = s[0] - s[4 * st], a4 = (s[2 * st] >> 1) - s[6 * st], a6 = s[2 * st] + (s[6 * st] >> 1);
    const int a1 = -s[3 * st] + s[5 * st] - s[7 * st] - (s[7 * st] >> 1), a3 = s[st] + s[7 * st] - s[3 * st] - (s[3 * st] >> 1);
    const int a5 = -s[st] + s[7 * st] + s[5 * st] + (s[5 * st] >> 1), a7 = s[3 * st] + s[5 * st] + s[st] + (s[st] >> 1);
    const int b0 = a0 + a6, b2 = a2 + a4, b4 = a2 - a4, b6 = a0 - a6;
    const int b1 = a1 + (a7 >> 2), b3 = a3 + (a5 >> 2), b5 = (a3 >> 2) - a5, b7 = a7 - (a1 >> 2);
    d[0] = b0 + b7; d[dt] = b2 + b5; d[2 * dt] = b4 + b3; d[3 * dt] = b6 + b1;
    d[4 * dt] = b6 - b1; d[5 * dt] = b4 - b3; d[6 * dt] = b2 - b5; d[7 * dt] = b0 - b7;
}
void ref_add8x8_idct8(uint8_t *dst, int sd, const uint8_t *pred, int sp, const int16_t dct[64]) {
    int d[64], t[64];
    for (int i = 0; i < 64; i++) d[i] = dct[i];
    for (int y = 0; y < 8; y++) idct8_1d(d + y * 8, 1, t + y * 8, 1); /* horizontal first (8.5.13) */
    for (int x = 0; x < 8; x++) idct8_1d(t + x, 8, d + x, 8);
    for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++) dst[y * sd + x] = (uint8_t)clip255(pred[y * sp + x] + ((d[y * 8 + x] + 32) >> 6));
}
/* [EXT] x264 pixel_sa8d_8x8 without its final rounding: sum |8x8 Hadamard of the difference| */
int ref_sa8d_8x8_raw(const uint8_t *a, int sa, const uint8_t *b, int sb) {
    int d[64], t[64], s = 0;
    for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++) d[y * 8 + x] = a[y * sa + x] - b[y * sb + x];
    for (int pass = 0; pass < 2; pass++) {
        for (int r = 0; r < 8; r++) {
            const int *p = pass ? t + r : d + r * 8;
            const int st = pass ? 8 : 1;
            int *o = pass ? d + r : t + r * 8;
            const int a0 = p[0] + p[st], a1 = p[0] - p[st], a2 = p[2 * st] + p[3 * st], a3 = p[2 * st] - p[3 * st];
            const int a4 = p[4 * st] + p[5 * st], a5 = p[4 * st] - p[5 * st], a6 = p[6 * st] + p[7 * st], a7 = p[6 * st] - p[7 * st];
            const int b0 = a0 + a2, b1 = a1 + a3, b2 = a0 - a2, b3 = a1 - a3, b4 = a4 + a6, b5 = a5 + a7, b6 = a4 - a6, b7 = a5 - a7;
            o[0] = b0 + b4; o[st] = b1 + b5; o[2 * st] = b2 + b6; o[3 * st] = b3 + b7;
            o[4 * st] = b0 - b4; o[5 * st] = b1 - b5; o[6 * st] = b2 - b6; o[7 * st] = b3 - b7;
        }
    }
    for (int i = 0; i < 64; i++) s += iabs(d[i]);
    return s;
}
/* [EXT] x264_decimate_score64 over zigzag-ordered levels of one 8x8 block */
int ref_decimate_score64(const int16_t *zz) {
    static const uint8_t ds8[64] = {3, 3, 3, 3, 2, 2, 2, 2, 2, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1};
    int idx = 63, score = 0;
    while (idx >= 0 && zz[idx] == 0) idx--;
    while (idx >= 0) {
        int run = 0;
        if (iabs(zz[idx--]) > 1) return 9;
        while (idx >= 0 && zz[idx] == 0) { idx--; run++; }
        score += ds8[run];
    }
    return score;
}
