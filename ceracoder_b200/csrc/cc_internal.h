/*
 * cc_internal.h — structures shared by the host C code and the CUDA kernels of libceracuda.
 * Plain C (included from .c and .cu).
 */
#ifndef CC_INTERNAL_H
#define CC_INTERNAL_H
#include <stddef.h>
#include <stdint.h>
#ifdef CC_EMU_BUILD
#include "cuda_emu_rt.h"
#else
#include <cuda_runtime_api.h>
#endif
#ifdef __cplusplus
extern "C" {
#endif

/* Plane geometry in HBM.  Every plane is stored with a replicated border so that motion vectors may
 * point outside the picture (unrestricted MVs) without clamping in the inner loops:
 *   luma-sized planes (cur Y, rec Y, half-pel b/h/j): origin at byte 48 of row 48, pitch = roundup(W+96,128)
 *   chroma planes: origin at byte 32 of row 32, pitch = roundup(W/2+64,64)
 *   half-resolution luma (coarse search): same layout as chroma planes.
 * Origins are 16-byte aligned so that vector loads of whole macroblock rows are aligned. */
#define CC_PAD 48
#define CC_PADC 32
#define CC_HPAD 40 /* half-pel planes are valid on [-40, W+40) x [-40, H+40) */
#define CC_L1R 16  /* coarse search range in half-res pixels */

enum { CC_MB_PSKIP = 0, CC_MB_P16 = 1, CC_MB_I16 = 2, CC_MB_I4 = 3 };

#define CC_ALIGN16 __attribute__((aligned(16))) /* lets the kernels fetch a record with one 128-bit load */

/* motion-search result per macroblock */
typedef struct CC_ALIGN16 {
    int16_t l1_mv[2]; /* coarse vector, half-res pixels */
    int16_t mv[2];    /* refined vector, quarter-pel */
    int32_t satd;     /* SATD of the 16x16 prediction at mv */
    int32_t part;     /* 1: four 8x8 partitions (cc_me8) cost less than the 16x16 partition -> P_8x8 */
} cc_me; /* 16 bytes */

/* 8x8-partition search result per macroblock (quadrants in decoding order: 0 top-left, 1 top-right, 2 bottom-left, 3 bottom-right) */
typedef struct CC_ALIGN16 {
    int16_t mv[4][2]; /* quarter-pel */
    int32_t satd[4];
} cc_me8; /* 32 bytes */

/* device-side macroblock record: neighbour context for prediction and deblocking */
typedef struct {
    uint8_t type, pad0;
    uint16_t nz_luma;   /* bit blk (z-order): luma 4x4 block has a non-zero level */
    int16_t mv[2];
    uint8_t i4_mode[16];
} cc_mbdev; /* 24 bytes */

/* host-side macroblock record (what the CAVLC writer needs besides the coefficient payload) */
typedef struct {
    uint8_t type, modes /* i16_mode | chroma_mode<<2 */, cbp, t8 /* bit 0: transform_size_8x8_flag, bit 1: P_8x8 */;
    int16_t mvd[2]; /* P_8x8: the four quadrants' mvd are the first 8 int16 of the macroblock's symbol slot */
} cc_mbhost; /* 8 bytes */

/* result of the parallel P-frame pre-pass (k_p_code) for one macroblock: the MB coded as P_L0_16x16 at the motion
 * search vector (levels in the MB's symbol slot, reconstruction in k.rec), plus skip probes at the two most likely
 * P_Skip vectors.  The wavefront kernel only decides; it recomputes nothing for macroblocks that stay inter. */
typedef struct CC_ALIGN16 {
    uint8_t cbp;        /* luma 8x8 bits | chroma << 4, coded at me.mv */
    uint8_t skip_ok;    /* bit 0: nothing survives at cand[] ; bit 1: nothing survives at (0,0) ; bit 2: k.rec holds the prediction at cand[]
                         * and the P_L0_16x16 reconstruction is parked in k.park (the pre-pass expects P_Skip at cand) ; bit 3: the same
                         * with the prediction at (0,0) (never together with bit 2) */
    uint16_t nz_luma;
    int16_t cand[2];    /* P_Skip vector predicted from the neighbours' search vectors */
    uint16_t sym_len;   /* int16 units used in the MB's symbol slot */
    uint8_t t8;         /* coded with the 8x8 transform */
    uint8_t part;       /* coded as P_8x8 (levels start 8 int16 into the slot, leaving room for the four mvd) */
    int32_t est_intra;  /* Intra16x16 SATD cost with SOURCE neighbours: gate of the real intra analysis */
} cc_pre; /* 16 bytes */

/* per-MB-row symbol stream header, written by the encode kernel */
typedef struct {
    uint32_t bytes;   /* payload bytes used by this row (multiple of 2) */
    uint32_t n_skip, n_inter, n_intra;
    uint32_t n_gate, n_probe; /* P frames: MBs that ran the real intra analysis / the in-wavefront skip probe */
    uint32_t pad;
} cc_rowhdr;

#define CC_CYC_N 32
#define CC_ROW_PAYLOAD_PER_MB 1024 /* worst case: 8 (i4 modes) + 27 blocks * (2 + 32) = 926 bytes */
/* worst-case CAVLC bits of one macroblock: 16 (coeff_token) + 28 per level (escape code) + 9 (total_zeros) per block, 27 blocks
 * with 400 levels, plus ~300 header bits = ~11.7 kbit; 1536 bytes per macroblock of slice bit buffer can never overflow */
#define CC_BS_BYTES_PER_MB 1536

typedef struct {
    int w, h;          /* display size */
    int W, H;          /* coded size (multiples of 16) */
    int mb_w, mb_h, rps; /* rps: macroblock rows per slice */
    int band;            /* macroblock rows per CTA of the wavefront kernels: min(rps, 8); a taller slice spans several bands */
    int frame_seq;       /* frame counter (from 1): tag of the cross-band progress words */
    uint32_t *gprog;     /* [3][mb_h] progress words (tag << 12 | macroblocks done) of k_mb_encode, the luma filter, the chroma filter */
    int sy, sc;        /* pitches: luma-sized planes, chroma-sized planes (also half-res luma) */
    uint8_t *cur[3];   /* origins (pixel 0,0) */
    uint8_t *rec[3];   /* reconstruction being written */
    uint8_t *ref[3];   /* previous reconstruction (deblocked, borders extended) */
    uint8_t *hb, *hh, *hj; /* half-pel planes of ref: horizontal, vertical, centre */
    uint8_t *l1cur, *l1ref;
    unsigned long long tmap_l1[16]; /* CUtensorMap (128 B, opaque here) over the allocation of l1ref: k_me_l1 loads its search windows with TMA */
    int tmap_l1_ok;
    const void *tmaps;       /* host memory: five CUtensorMap over the luma-sized planes {rec A, rec B, b, h, j} (k_me_l0's sub-pel windows) */
    const uint8_t *tmap_plane[2]; /* origins of rec A / rec B: which map belongs to this frame's reference */
    uint8_t *uvstage;  /* NV12 interleaved chroma staging (pitch sy), or NULL */
    uint8_t *pkstage;  /* packed-input staging (YUY2: 2 bytes / pixel, BGRx: 4), pitch pks, or NULL */
    int pks;
    cc_me *me;
    cc_me8 *me8;
    int part8;         /* 1: the motion search may choose P_8x8 */
    cc_mbdev *mbd;
    uint32_t *nbrec;   /* per MB 5 words {type | t8 << 8 | nz_luma << 16, mv of quadrants 0..3}: what k_deblock needs from k_mb_encode */
    cc_mbhost *mbh;    /* device copy */
    cc_pre *pre;
    int trail;           /* 1: k_deblock runs next to k_mb_encode and follows its progress words (0: it is launched after it) */
    int trail_at;        /* k_mb_encode releases the launch of k_deblock when its CTAs' top rows reach this macroblock column */
    uint8_t *park;     /* per MB 384 bytes (16x16 luma, 8x8 Cb, 8x8 Cr as in WarpMem.recY / recC): see cc_pre.skip_ok bit 2 */
    uint16_t *symlen;  /* per MB: int16 units used in its symbol slot (even; 0 for P_Skip) */
    uint8_t *ubot[3];  /* unfiltered bottom sample line of every MB row: [mb_h][sy] luma, [mb_h][sc] per chroma plane */
    uint8_t *uright;   /* unfiltered right column of every MB: 32 bytes {16 luma, 8 Cb, 8 Cr} */
    uint8_t *rowbuf;   /* one symbol slot of CC_ROW_PAYLOAD_PER_MB bytes per macroblock */
    cc_rowhdr *rowhdr; /* device */
    int qp, is_idr;
    uint16_t qmf[4][3], qbias[4][3]; /* quantiser: 0 luma intra, 1 luma inter, 2 chroma intra, 3 chroma inter; class-indexed */
    uint16_t qthr[4][3][2];          /* smallest |coefficient| that quantises to a level of magnitude >= 1 / >= 2 (P_Skip probes only need that much) */
    uint16_t q8mf[6], q8bias[6];     /* 8x8 transform, luma inter; class-indexed (H.264 8.5.9) */
    /* packed output in mapped pinned host memory: [cc_rowhdr * mb_h][cc_mbhost * nmb][payload rows, 16-byte aligned] */
    uint8_t *out_host_dev; /* device pointer of that mapping */
    uint32_t out_cap;
    /* device-side CAVLC (cc_k_cavlc.cu): */
    int dev_entropy;       /* 1: the slices' RBSPs are written on the device, 0: packed symbols go to the host writer */
    int frame_num, idr_pic_id; /* slice header fields of the frame being coded */
    uint32_t *mbbits;      /* per MB: bit length of its macroblock_layer() (incl. the mb_skip_run before it) */
    struct CvRec *cvrec;   /* per MB 32 bytes: what k_cavlc_size found outside the macroblock (nC per block, mb_skip_run) for k_cavlc_write */
    uint32_t *bs;          /* per slice a zero-initialised bit buffer of bs_slice_words words (MSB-first words) */
    uint32_t bs_slice_words;
    uint32_t *bs_ctl;      /* [0, ns): row CTAs of the slice that finished; [ns, 2 ns): the slice's RBSP bytes */
    int dbg;               /* debug switch (CC_DBG_ME): 1 = k_me_l0 stages its sub-pel windows with plain loads instead of TMA */
    unsigned long long *cyc; /* debug (CC_DEBUG_CYCLES=1): per MB row CC_CYC_N cycle / event counters of the wavefront kernels, or NULL */
} cc_kparams;

/* kernel launch wrappers (cc_kernels.cu); each returns the number of kernels launched */
int cck_ingest(const cc_kparams *k, int fmt, cudaStream_t st);
int cck_finish_ref(const cc_kparams *k, cudaStream_t st); /* rec -> borders, half-pel planes */
int cck_border(const cc_kparams *k, cudaStream_t st);
int cck_hpel(const cc_kparams *k, cudaStream_t st);
int cck_hpel_border(const cc_kparams *k, cudaStream_t st); /* both in one launch (k_hpel reads the picture only) */
int cck_l1border(const cc_kparams *k, cudaStream_t st); /* edge replication of the half-resolution reference (before the coarse search) */
int cck_me_l1_prepare(cc_kparams *k); /* builds k->tmap_l1 (after the planes are allocated); CC_OK or CC_ERR_CUDA */
int cck_me_l1(const cc_kparams *k, cudaStream_t st);
int cck_me_l0(const cc_kparams *k, cudaStream_t st);
int cck_me_l0_prepare(cc_kparams *k, uint8_t *const planes[5], void *maps); /* tensor maps of the sub-pel windows; 0 or CC_ERR_CUDA */
int cck_motion(const cc_kparams *k, cudaStream_t st);
int cck_p_code(const cc_kparams *k, cudaStream_t st);
int cck_encode(const cc_kparams *k, cudaStream_t st);
int cck_deblock(const cc_kparams *k, cudaStream_t st, int after_trigger); /* after_trigger: programmatic stream serialisation behind k_mb_encode, the kernel before it in st (0: a plain launch) */
int cck_pack(const cc_kparams *k, cudaStream_t st);
int cck_cavlc(const cc_kparams *k, cudaStream_t st); /* k_cavlc_size + k_cavlc_write: slice RBSPs into the mapped host buffer */

#ifdef __cplusplus
}
#endif
#endif
