/*
 * cc_host.h — host-side (plain C) pieces of libceracuda: bit writer, H.264 headers / NAL packaging,
 * CAVLC slice writer, rate control.  north_star: "the per-frame rate-control/QP update and the CAVLC
 * bitstream writer" are host C that call CUDA through the C ABI.
 */
#ifndef CC_HOST_H
#define CC_HOST_H
#include "cc_internal.h"

/* ---- MSB-first bit writer with a 64-bit accumulator ------------------------------------------- */
typedef struct {
    uint8_t *start, *p, *end;
    uint64_t acc; /* the low `n` bits are pending */
    int n;
    int overflow;
} cc_bw;

static inline void cc_bw_init(cc_bw *b, uint8_t *buf, size_t cap) { b->start = b->p = buf; b->end = buf + cap; b->acc = 0; b->n = 0; b->overflow = 0; }
static inline void cc_bw_put(cc_bw *b, int n, uint32_t v) { /* 0 <= n <= 32, v < 2^n */
    b->acc = (b->acc << n) | v;
    b->n += n;
    if (b->n >= 32) {
        b->n -= 32;
        uint32_t w = (uint32_t)(b->acc >> b->n);
        if (b->p + 4 <= b->end) { b->p[0] = (uint8_t)(w >> 24); b->p[1] = (uint8_t)(w >> 16); b->p[2] = (uint8_t)(w >> 8); b->p[3] = (uint8_t)w; b->p += 4; }
        else b->overflow = 1;
    }
}
static inline void cc_bw_ue(cc_bw *b, unsigned v) {
    unsigned k = v + 1;
    int n = 31 - __builtin_clz(k);
    cc_bw_put(b, n, 0);
    cc_bw_put(b, n + 1, k);
}
static inline void cc_bw_se(cc_bw *b, int v) { cc_bw_ue(b, v > 0 ? 2u * (unsigned)v - 1u : 2u * (unsigned)(-v)); }
static inline long cc_bw_bits(const cc_bw *b) { return (long)(b->p - b->start) * 8 + b->n; }
/* rbsp_trailing_bits + flush; returns byte count */
static inline size_t cc_bw_finish(cc_bw *b) {
    cc_bw_put(b, 1, 1);
    if (b->n & 7) cc_bw_put(b, 8 - (b->n & 7), 0);
    while (b->n >= 8) {
        b->n -= 8;
        if (b->p < b->end) *b->p++ = (uint8_t)(b->acc >> b->n); else b->overflow = 1;
    }
    return (size_t)(b->p - b->start);
}

/* ---- headers and NAL packaging (cc_bitstream.c) ------------------------------------------------ */
typedef struct {
    int w, h, mb_w, mb_h, fps_n, fps_d;
} cc_seq;
/* level_idc (Table A-1) that covers a picture of fs macroblocks at mbps macroblocks/s and 30 Mbit/s; 0 = beyond level 5.2 */
int cc_level_idc_for(long fs, long mbps);
size_t cc_write_sps(const cc_seq *s, uint8_t *rbsp, size_t cap);
size_t cc_write_pps(uint8_t *rbsp, size_t cap);
void cc_write_slice_header(cc_bw *b, int first_mb, int is_idr, int frame_num, int idr_pic_id, int qp);
/* Annex-B start code + NAL header byte + payload with emulation prevention; returns bytes written (0 = no room) */
size_t cc_write_nal(uint8_t *out, size_t cap, int nal_hdr, const uint8_t *rbsp, size_t n);

/* ---- CAVLC slice writer (cc_cavlc.c) ------------------------------------------------------------ */
typedef struct {
    int mb_w, mb_h, rps;
    uint8_t *nnz;      /* per MB 24 total_coeff values (16 luma z-order, 4 Cb, 4 Cr); mb_w*mb_h*24 */
} cc_cavlc_ctx;
/* writes slice_data() of MB rows [row0,row1) into b.  mbh: frame array; rowhdr/payload: per-row symbol streams.
 * returns 0, or -1 on malformed symbols */
int cc_cavlc_write_slice(cc_cavlc_ctx *c, cc_bw *b, int is_idr, int row0, int row1, const cc_mbhost *mbh,
                         const cc_rowhdr *rowhdr, const uint8_t *const *row_payload);

/* ---- rate control (cc_rc.c) ---------------------------------------------------------------------- */
typedef struct {
    double fps;
    int qp_min, qp_max;
    int bitrate;          /* bps currently applied */
    double coef[2];       /* bits * qscale of the last frame of each type (0 P, 1 I), smoothed */
    double have[2];
    double fullness;      /* accumulated (actual - target) bits */
    double last_qp_p;
    int frames;
    int key_int;          /* IDR period in frames */
    int step;             /* the target moved by more than a ladder step since the last P frame: follow the model at once */
    double vbv_fill;      /* leaky bucket: bits sent above the target rate that have not drained yet */
    double vbv_seconds;   /* bucket size in seconds of the current bitrate (x264enc vbv-buf-capacity: 600 ms) */
    double ema, max_step; /* weight of a new P frame in the complexity model; largest QP change between P frames */
    int fast;             /* P frames left that update the model at full speed (after a balancer step) */
    double vbv_peak;      /* highest fill seen, in seconds of bitrate (reported by tests / tools) */
} cc_rc;
void cc_rc_init(cc_rc *rc, double fps, int bitrate, int qp_min, int qp_max, int key_int);
int cc_rc_frame_qp(cc_rc *rc, int bitrate_now, int is_idr, int nmb);
void cc_rc_update(cc_rc *rc, int is_idr, int qp, long bits);

/* quantiser tables for one qp: mf[3], bias[3] per class (0 even-even, 1 odd-odd, 2 mixed) */
void cc_quant_tables(int qp, int intra, uint16_t mf[3], uint16_t bias[3]);
void cc_quant8_tables(int qp, uint16_t mf[6], uint16_t bias[6]);
#endif
