/*
 * oracle/ref_h264.h — public interface of the CPU oracle (libceraoracle.so).
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library; the product
 * (libceracuda.so) never links, loads or calls it.
 *
 * PARITY PINNING: the reference (CERALIVE/ceracoder) contains no encoder source
 * and no encoder golden vectors; its hot path is the external `x264enc`
 * element named at /root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6
 * (libx264, un-vendored, version un-pinned; absent from this image).  At the
 * reference level parity is therefore UNPINNED.  This oracle restates
 *   - ITU-T H.264 clauses 7.3, 8.3, 8.4, 8.5, 8.7, 9.1, 9.2 for every normative stage,
 *     pinned by libavcodec's h264 decoder (tests/test_oracle_decode.py: decoded
 *     picture == oracle reconstruction, every frame), and
 *   - the published libx264 C forms of the non-normative primitives
 *     (pixel_sad, pixel_satd, sub4x4_dct, dct4x4dc, quant_4x4, decimate_score) [EXT],
 *     pinned by definition-level known answers (tests/test_oracle_prims.py).
 */
#ifndef REF_H264_H
#define REF_H264_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define REF_PAD 48 /* luma border (px) around the coded picture; chroma = REF_PAD/2 */

/* ---- primitives (batch-free, single block) -------------------------------- */
int ref_sad(const uint8_t *a, int sa, const uint8_t *b, int sb, int w, int h);
int ref_satd_4x4(const uint8_t *a, int sa, const uint8_t *b, int sb);
int ref_satd(const uint8_t *a, int sa, const uint8_t *b, int sb, int w, int h);
void ref_sub4x4_dct(int16_t dct[16], const uint8_t *cur, int sc, const uint8_t *pred, int sp);
void ref_dct4x4dc(int16_t d[16]);
void ref_idct4x4dc(int16_t d[16]);
void ref_dct2x2dc(int16_t d[4]);
void ref_quant_tables(int qp, int intra, uint16_t mf[16], uint16_t bias[16]);
int ref_quant_4x4(int16_t dct[16], const uint16_t mf[16], const uint16_t bias[16]);
int ref_quant_dc(int16_t *dct, int n, int mf, int bias);
void ref_dequant_4x4(int16_t dct[16], int qp);
void ref_dequant_4x4_dc(int16_t dct[16], int qp);
void ref_dequant_2x2_dc(int16_t out[4], const int16_t in[4], int qp);
void ref_add4x4_idct(uint8_t *dst, int sd, const uint8_t *pred, int sp, const int16_t dct[16]);
int ref_decimate_score(const int16_t *zz, int n);
/* 8x8 transform path (High profile) */
void ref_sub8x8_dct8(int16_t dct[64], const uint8_t *cur, int sc, const uint8_t *pred, int sp);
int ref_quant_8x8(int16_t dct[64], int qp, int intra);
void ref_dequant_8x8(int16_t dct[64], int qp);
void ref_add8x8_idct8(uint8_t *dst, int sd, const uint8_t *pred, int sp, const int16_t dct[64]);
int ref_sa8d_8x8_raw(const uint8_t *a, int sa, const uint8_t *b, int sb);
int ref_decimate_score64(const int16_t *zz);
void ref_pred4x4(uint8_t dst[16], int mode, const uint8_t *top /*8, [-1]=topleft valid*/,
                 const uint8_t *left /*4*/, int has_top, int has_left);
void ref_pred16x16(uint8_t *dst, int sd, int mode, const uint8_t *rec, int sr, int has_top,
                   int has_left);
void ref_pred8x8c(uint8_t *dst, int sd, int mode, const uint8_t *rec, int sr, int has_top,
                  int has_left);
void ref_mc_luma(uint8_t *dst, int sd, const uint8_t *ref, int sr, int mvx, int mvy, int w,
                 int h);
void ref_mc_chroma(uint8_t *dst, int sd, const uint8_t *ref, int sr, int mvx, int mvy, int w,
                   int h);
int ref_mv_bits(int d);
int ref_median3(int a, int b, int c);

/* ---- CAVLC ----------------------------------------------------------------- */
/* writes residual_block_cavlc() for zigzag-ordered levels; returns bits written.
 * out is an MSB-first bit buffer, *bitpos advanced. nC = -1 for chroma DC. */
int ref_cavlc_block(uint8_t *out, int *bitpos, const int16_t *zz, int max_coeff, int nC);

/* ---- encoder ------------------------------------------------------------------ */
enum { REF_MB_PSKIP = 0, REF_MB_P16 = 1, REF_MB_I16 = 2, REF_MB_I4 = 3 };

typedef struct {
    uint8_t type, i16_mode, chroma_mode, cbp; /* cbp: bits0-3 luma 8x8, bits4-5 chroma */
    uint8_t i4_mode[16];
    int16_t mv[2], mvd[2];
    uint8_t nnz[28]; /* total_coeff: 0-15 luma 4x4, 16-19 Cb AC, 20-23 Cr AC, 24 lumaDC, 25/26 chroma DC */
    int16_t me_mv[2];
    int32_t me_satd;
    int16_t l1_mv[2];
    int32_t bits; /* bits this MB contributed to slice data */
    uint8_t t8;   /* transform_size_8x8_flag (inter MBs only) */
    uint8_t part; /* 0: one 16x16 partition, 1: P_8x8 (four 8x8 sub-macroblocks, sub_mb_type P_L0_8x8) */
    uint8_t me_part; /* partitioning chosen by the motion search */
    uint8_t pad8;
    int16_t mv4[4][2], mvd4[4][2]; /* final vector / mvd per 8x8 quadrant (all equal for 16x16 and P_Skip) */
    int16_t me_mv8[4][2];          /* 8x8 search vectors, quarter-pel */
    int32_t me_satd8[4];
    int32_t me_cost16, me_cost8;   /* search costs (SATD + lambda * approximate mvd bits [+ partition overhead]) */
    uint16_t dbg_i4_blk[16];                           /* per 4x4 block (decoding order): best Intra4x4 cost (SATD + lambda * mode bits), clamped */
    int32_t dbg_i4_blk_exact[16];                      /* the same, not clamped (the projection test of analyse_i4 reads these) */
    int32_t dbg_est_intra;                             /* P frames: the source-based Intra16x16 estimate behind the intra gate */
    int32_t dbg_cost_inter, dbg_cost_i16, dbg_cost_i4; /* mode-decision costs of the last ref_mb_encode (1 << 30 = not evaluated); analysis tools only */
} ref_mb_info;

typedef struct ref_enc ref_enc;

ref_enc *ref_enc_open(int width, int height, int fps_n, int fps_d, int rows_per_slice);
void ref_enc_close(ref_enc *e);
/* enable (1) / disable (0) P_8x8 partitions; ref_enc_open leaves them on.  Mirrors cc_enc_params.partitions of the product. */
void ref_enc_set_partitions(ref_enc *e, int on);
/* 1: upper-bound mode — NOT the parity oracle: exhaustive +-24 full-pel search, 8x8 partitions always refined, and per macroblock the
 * candidate (P_Skip, P_L0_16x16 with either transform size, P_8x8, Intra16x16, Intra4x4) with the smallest SSD + lambda * real CAVLC
 * bits.  tools/rd_oracle.py measures the product's RD curve against it. */
void ref_enc_set_effort(ref_enc *e, int on);
/* OpenMP threads the slice / macroblock loops use (0 = leave as is) */
void ref_set_threads(int n);
/* encode one frame (I420, arbitrary strides). is_idr: 1 = IDR, 0 = P. qp 10..51.
 * returns AU size in bytes (Annex-B, SPS/PPS prepended on IDR), <0 on error */
int ref_enc_encode(ref_enc *e, const uint8_t *y, int sy, const uint8_t *u, int su,
                   const uint8_t *v, int sv, int is_idr, int qp, uint8_t *out, int out_cap);
/* copy out the (deblocked) reconstruction, coded size (mb_w*16 x mb_h*16), tight strides */
void ref_enc_get_recon(const ref_enc *e, uint8_t *y, uint8_t *u, uint8_t *v);
const ref_mb_info *ref_enc_mb_info(const ref_enc *e);
int ref_enc_mb_w(const ref_enc *e);
int ref_enc_mb_h(const ref_enc *e);
/* half-resolution planes used by the coarse search (debug / parity of the pyramid kernel) */
const uint8_t *ref_enc_l1_cur(const ref_enc *e, int *stride);

#ifdef __cplusplus
}
#endif
#endif
