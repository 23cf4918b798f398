/* oracle/ref_int.h — internal state of the CPU reference encoder. TEST INFRASTRUCTURE ONLY. */
#ifndef REF_INT_H
#define REF_INT_H
#include "ref_h264.h"

#define COEF_PER_MB (27 * 16)

struct ref_enc {
    int w, h, mb_w, mb_h, rps, fps_n, fps_d;
    int frame_num, idr_pic_id, have_ref;
    int sy, sc;                 /* strides of bordered luma / chroma planes */
    uint8_t *cur[3];            /* source, padded to MB multiple (bordered layout, border unused) */
    uint8_t *rec[3], *ref[3];   /* reconstruction being built / previous reconstruction; point at (0,0) */
    uint8_t *alloc[9];
    int s1;                     /* stride of half-res planes (border REF_PAD/2) */
    uint8_t *l1cur, *l1ref, *l1alloc[2];
    ref_mb_info *mb;
    int16_t *coef;              /* [mb][27][16], zigzag order; AC-only blocks use [0..14] */
    int qp, is_idr;
    uint8_t *slice_buf;         /* per-slice RBSP scratch, slice_cap bytes each */
    int slice_cap;
    int *slice_bits;
    int part8;                  /* 1: the motion search may choose P_8x8 (four 8x8 partitions) */
    int effort;                 /* 0: the parity oracle.  1: upper-bound mode (ref_enc_set_effort): exhaustive full-pel search, RD mode decision */
};

/* ref_mb.c */
void ref_mb_encode(struct ref_enc *e, int mbx, int mby);
void ref_mb_encode_rd(struct ref_enc *e, int mbx, int mby);
void ref_mb_deblock(struct ref_enc *e, int mbx, int mby);
void ref_mb_write(struct ref_enc *e, int mbx, int mby, uint8_t *out, int *bitpos, int *skip_run);

static inline int ref_slice_first_row(const struct ref_enc *e, int mby) { return (mby / e->rps) * e->rps; }
#endif
