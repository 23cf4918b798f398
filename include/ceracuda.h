/*
 * include/ceracuda.h — C ABI of libceracuda.so, the B200-native H.264 encoder that stands in the
 * `x264enc` slot of a ceracoder pipeline.
 *
 * What this replaces in the reference (CERALIVE/ceracoder, paths under /root/reference):
 *   - the encoder element named in pipeline text, `x264enc speed-preset=3 key-int-max=60 name=venc_kbps`
 *     (pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6 and siblings; instantiated by
 *     gst_parse_launch at src/io/pipeline_loader.c:59).  GStreamer's x264enc binds libx264's
 *     x264_encoder_open / x264_encoder_encode / x264_encoder_reconfig / x264_encoder_close [EXT];
 *     cc_enc_open / cc_enc_encode / cc_enc_set_bitrate / cc_enc_close are those four entry points.
 *   - the one place ceracoder touches the encoder at run time: g_object_set(element, "bps", ...) at
 *     src/gst/encoder_control.c:53 (20 ms timer, src/ceracoder.c:263) -> cc_enc_set_bitrate().
 * The GStreamer element `cudah264enc` (ceracoder_b200/gst/gstcudah264enc.c) and the ctypes mirror
 * (ceracoder_b200/encoder.py) sit on top of exactly these symbols.
 *
 * Conventions follow the reference's own C modules (src/net/srt_client.c:40-109,
 * src/gst/encoder_control.c:39-57): opaque handle, int return, 0 = success, small negative codes,
 * diagnostics on stderr.  No torch / C++ types cross this boundary.
 *
 * There is NO CPU fallback: every entry point that computes needs a CUDA device and fails with
 * CC_ERR_CUDA otherwise.
 */
#ifndef CERACUDA_H
#define CERACUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define CC_OK 0
#define CC_ERR_PARAM (-1)  /* bad argument (size, format, null pointer) */
#define CC_ERR_CUDA (-2)   /* CUDA runtime error or no device */
#define CC_ERR_NOMEM (-3)  /* host or device allocation failed */
#define CC_ERR_STATE (-4)  /* call order violated (collect without submit, ...) */
#define CC_ERR_OVERFLOW (-5) /* access unit did not fit the internal output buffer */

/* CC_FMT_YUY2 / CC_FMT_BGRX: packed input (plane[0], stride[0] only), converted to 4:2:0 inside the ingest kernel — the
 * `videoconvert` in front of x264enc in /root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:3-5.
 * YUY2: luma as is, chroma = rounded mean of the two rows.  BGRx: 8-bit limited-range matrix, BT.709 above 576 lines else
 * BT.601 (GStreamer's default colorimetry for YUV video of that size), chroma from the rounded 2x2 mean of R, G, B.
 * The exact integer arithmetic is stated in oracle/ref_convert.c. */
enum { CC_FMT_I420 = 0, CC_FMT_NV12 = 1, CC_FMT_YUY2 = 2, CC_FMT_BGRX = 3 };
enum { CC_FORCE_IDR = 1u, CC_FRAME_DEVICE = 2u /* plane pointers are device memory */ };

typedef struct cc_enc cc_enc;

typedef struct {
    int device;          /* CUDA ordinal */
    int width, height;   /* even, >= 16 */
    int fps_n, fps_d;
    int format;          /* CC_FMT_* of the frames handed to cc_enc_encode */
    int bitrate_bps;     /* initial target bitrate (ceracoder writes max_bitrate before PLAYING, src/ceracoder.c:515-518) */
    int key_int_max;     /* IDR period in frames; 60 in every reference pipeline file; 0 -> 60 */
    int rows_per_slice;  /* macroblock rows per slice: 1..8, a multiple of 8, or anything >= the picture's macroblock rows (one slice per
                          * picture); 0 -> 8 (nine slices per 1080p picture).  A slice is an independent unit of prediction: P_Skip cannot
                          * carry motion in a slice's first row (8.4.1.1), so fewer, taller slices cost fewer bits — measured on S2 1080p:
                          * 8 rows -8.7 %, 2 rows +13 %, 1 row +34 % against 4 rows at equal Y-PSNR, one slice per picture another -15 %
                          * (profiles/rd_slices_r02.txt) — and more time: the macroblock wavefront of a slice is mb_w + 2 (rows - 1) steps long.
                          * Up to 8 rows a slice is one CTA of the wavefront kernels; taller slices chain 8-row bands across CTAs */
    int qp_fixed;        /* > 0: constant-QP mode (rate control off); used by parity tests */
    int qp_min, qp_max;  /* 0 -> 10 / 51 */
    int partitions;      /* 1: the motion search may split a macroblock into four 8x8 partitions (P_8x8); cc_enc_params_default sets 1 */
    int threads;         /* host threads that write the slices of one access unit (x264enc's `threads` property: its sliced-threads
                          * under tune=zerolatency).  0 / 1: the calling thread alone — right when one thread per stream already keeps
                          * the host busy; 2..16: a single large stream (2160p) whose bitstream writing would otherwise take longer
                          * than its device work.  The access unit is the same bytes either way */
    int host_entropy;    /* 0 (default): CAVLC runs on the device: finished slice RBSPs cross PCIe and the host only adds
                          * start codes, NAL headers and emulation prevention.  1: the device hands over packed symbols and the host C
                          * writer (cc_cavlc.c) codes them — the round-1 path, kept as a cross-check.  Same bytes either way */
} cc_enc_params;

typedef struct {
    const uint8_t *plane[3]; /* I420: Y,U,V.  NV12: Y, interleaved UV, unused */
    int stride[3];           /* bytes per row; GStreamer strides are not tight (GstVideoInfo) */
    int64_t pts;
} cc_frame;

typedef struct {
    const uint8_t *data; /* Annex-B access unit, owned by the handle, valid until the next encode/collect/close */
    int size;
    int is_idr;
    int qp;
    int64_t pts;
    int n_skip, n_inter, n_intra; /* macroblock statistics of the frame */
    int sym_bytes;       /* bytes the device wrote to host memory for this frame (slice RBSPs, or packed symbols with host_entropy) */
    float gpu_ms;        /* device time of the frame's kernels (CUDA events) */
    float host_ms;       /* host entropy coding + packaging time */
    int n_intra_analysed, n_skip_probed; /* P frames: MBs that took the slow paths of the wavefront kernel */
    float submit_ms;     /* host time spent inside cc_enc_submit issuing the frame's copies and launches */
    float wait_ms;       /* host time cc_enc_collect spent waiting for the device */
} cc_packet;

const char *cc_version(void);
const char *cc_strerror(int code);

void cc_enc_params_default(cc_enc_params *p);
int cc_enc_open(cc_enc **h, const cc_enc_params *p);
void cc_enc_close(cc_enc *h);

/* Thread-safe and wait-free (one atomic store); takes effect at the next frame boundary.  Accepts
 * bit/s; values are clamped to the reference's legal range 300 000..30 000 000 (src/core/bitrate_control.h:30-31). */
int cc_enc_set_bitrate(cc_enc *h, int bps);
int cc_enc_get_bitrate(const cc_enc *h);

/* One frame in, one access unit out (zero-latency).  = cc_enc_submit + cc_enc_collect. */
int cc_enc_encode(cc_enc *h, const cc_frame *in, unsigned flags, cc_packet *out);
/* Split form, so one host thread can keep several streams' GPU work in flight. */
int cc_enc_submit(cc_enc *h, const cc_frame *in, unsigned flags);
int cc_enc_collect(cc_enc *h, cc_packet *out);

/* Consecutive frames of one stream in one call: for i in [0, n): cc_enc_set_bitrate(h, bitrate_bps[i]) if it changed (what
 * src/gst/encoder_control.c:51-53 does), then cc_enc_encode(frames[i]) — the loop the element's streaming thread runs, so a
 * Python harness can drive many streams without serialising on the interpreter lock.  Sums over the n access units are added
 * to *acc (zero it first); the last access unit stays readable through `last` (may be NULL). */
typedef struct {
    long frames, au_bytes, sym_bytes;
    long n_skip, n_inter, n_intra, n_intra_analysed, n_skip_probed;
    double submit_ms, wait_ms, host_ms, gpu_ms;
    int last_qp;
    int n_idr;           /* IDR frames among the n */
} cc_seq_stats;
int cc_enc_encode_sequence(cc_enc *h, const cc_frame *frames, const int *bitrate_bps, int n, unsigned flags, cc_seq_stats *acc, cc_packet *last);

/* Test hooks. */
int cc_enc_set_next_qp(cc_enc *h, int qp); /* overrides rate control for the next submitted frame only */
int cc_enc_get_recon(cc_enc *h, uint8_t *y, uint8_t *u, uint8_t *v); /* coded size (MB multiple), tight rows */
int cc_enc_mb_w(const cc_enc *h);
int cc_enc_mb_h(const cc_enc *h);
/* per-macroblock motion-search results of the last P frame: 4 int16 per MB (coarse mv x,y in half-res px;
 * final mv x,y in quarter-pel) and one int32 SATD per MB */
int cc_enc_get_me(cc_enc *h, int16_t *mv4, int32_t *satd);
/* 8x8-partition search results of the last P frame: per MB 8 int16 (quarter-pel mv x,y of quadrants 0..3), 4 int32 SATD and
 * one byte: 1 = the search prefers P_8x8 */
int cc_enc_get_me8(cc_enc *h, int16_t *mv8, int32_t *satd4, uint8_t *part);
/* per-macroblock decisions of the last frame: 8 bytes per MB {type, i16_mode | chroma_mode << 2, cbp, transform_size_8x8_flag | P_8x8 << 1,
 * mvd_x(2), mvd_y(2)} */
int cc_enc_get_mbinfo(cc_enc *h, uint8_t *dst8);
size_t cc_enc_device_bytes(const cc_enc *h);
/* per-stage device timing with CUDA events on the encoder's own stream (bench.py's roofline figures) */
enum { CC_STAGE_UPLOAD = 0, CC_STAGE_INGEST, CC_STAGE_ME_COARSE, CC_STAGE_ME_FINE, CC_STAGE_P_CODE, CC_STAGE_MB_ENCODE, CC_STAGE_PACK,
       CC_STAGE_DEBLOCK, CC_STAGE_BORDER, CC_STAGE_HPEL, CC_STAGE_L1REF, CC_STAGE_WAIT_REF, CC_STAGE_COUNT }; /* in execution order; the
       access unit is complete after CC_STAGE_PACK, the later stages run while the host writes the bitstream.  CC_STAGE_WAIT_REF is
       not a kernel: the time the frame's stream waited, between the coarse and the fine search, for the previous frame's
       reference planes (in-loop filter, borders, half-pel planes) to be finished */
int cc_enc_set_profile(cc_enc *h, int on);
int cc_enc_get_stage_ms(const cc_enc *h, double *ms_sum /*[CC_STAGE_COUNT]*/, long *count /*[CC_STAGE_COUNT]*/);
/* Frames handed to cc_enc_submit / cc_enc_encode in pageable host memory (a mapped GstBuffer) are first copied, by the calling
 * thread, into a pinned frame the handle owns, so that the upload itself is asynchronous; page-locked frames (cc_host_alloc) and
 * device frames are used where they lie.  Count of frames that took the staging copy since open: */
long cc_enc_staged_frames(const cc_enc *h);
/* cc_enc_encode_sequence starts the upload of frame i + 1 while frame i is being coded (planar I420; a second set of source planes
 * per handle; pageable frames are copied to a second pinned staging frame by the calling thread first).  Count of frames that
 * arrived that way since open: */
long cc_enc_prefetched_frames(const cc_enc *h);
/* Launch order of the frames coded so far: how many of them ran the in-loop filter NEXT TO the macroblock kernel
 * (programmatic dependent launch) rather than after it.  The default does so while the handle's device has at most 8 handles open
 * in this process (CC_OVERLAP / CC_OVERLAP_MAX_HANDLES / CC_TRAIL_PCT in the environment override it; a handle with stage timing
 * on never does).  Both orders produce the same bytes.  No counterpart in x264enc. */
long cc_enc_overlapped_frames(const cc_enc *h);
/* number of kernel launches issued by this handle since open */
long cc_enc_kernel_launches(const cc_enc *h);

/* pinned host memory for frames (lets cc_enc_submit's upload run asynchronously) */
void *cc_host_alloc(size_t bytes);
void cc_host_free(void *p);
/* page-lock memory the caller already owns (a buffer pool, a shared-memory frame store) so that frames in it upload without the
 * staging copy; undo before freeing it */
int cc_host_register(void *p, size_t bytes);
int cc_host_unregister(void *p);
/* device memory helpers for callers that keep frames resident in HBM (CC_FRAME_DEVICE) */
void *cc_dev_alloc(size_t bytes);
void cc_dev_free(void *p);
int cc_dev_upload(void *dst, const void *src, size_t bytes);

/*
 * Batch primitives — the transform / quant / SAD / SATD / interpolation / deblock building blocks as the
 * kernels compute them, exposed for bit-exact parity tests against the oracle (tests/test_prims_gpu.py).
 * All pointers are HOST memory; n = number of blocks.  They replace libx264's C reference functions
 * named in BASELINE.json's north_star [EXT]: pixel_sad_16x16 / pixel_sad_8x8 / pixel_satd_4x4 /
 * sub4x4_dct / quant_4x4 / dequant_4x4 / add4x4_idct.
 */
int cc_prim_sad(const uint8_t *a, const uint8_t *b, int w, int h, int n, int32_t *out); /* blocks tight w*h */
int cc_prim_satd4x4(const uint8_t *a, const uint8_t *b, int n, int32_t *out);
int cc_prim_satd16x16(const uint8_t *a, const uint8_t *b, int n, int32_t *out);
/* residual -> 4x4 core transform -> quant -> dequant -> inverse -> reconstruct; intra selects the dead zone.
 * levels: n*16 int16 raster order; recon: n*16 u8 */
int cc_prim_tq4x4(const uint8_t *cur, const uint8_t *pred, int n, int qp, int intra, int16_t *levels, uint8_t *recon);
int cc_prim_dct4x4(const uint8_t *cur, const uint8_t *pred, int n, int16_t *coef);
/* 8x8 transform chain (sub8x8_dct8 -> quant_8x8 (inter dead zone) -> dequant_8x8 -> add8x8_idct8 [EXT]) over n_mb 16x16 blocks (tight 256 B each,
 * four 8x8 blocks per macroblock): levels n_mb*4*64 int16 raster, recon n_mb*256, sa8d n_mb = (sum|Hadamard8x8| over the MB + 2) >> 2 */
int cc_prim_tq8x8(const uint8_t *cur, const uint8_t *pred, int n_mb, int qp, int16_t *levels, uint8_t *recon, int32_t *sa8d);
/* intra 4x4 predictors: edge = n*13 bytes {left3,left2,left1,left0,topleft,top0..top7}; out n*16 */
int cc_prim_pred4x4(const uint8_t *edge, int mode, int n, uint8_t *out);
/* quarter-pel luma interpolation of n 16x16 blocks taken at (x,y)[i] + mv[i] from one w*h plane (borders replicated) */
int cc_prim_mc_luma(const uint8_t *plane, int w, int h, const int16_t *xy, const int16_t *mv, int n, uint8_t *out);

/* Allocation-free variant of cc_prim_satd16x16.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same inputs, same outputs, same kernel; the device staging lives in a fixed device array inside the library
 * (516 bytes per block x CC_PRIM_NOALLOC_MAX_BLOCKS = 516 KiB, held by every GPU that loads the library until the process
 * exits) instead of cudaMalloc'ed buffers, and the call makes no cudaMalloc / cudaFree / allocating library call.
 * n > CC_PRIM_NOALLOC_MAX_BLOCKS is rejected with CC_ERR_PARAM (and a line on stderr).  All calls share the one array, which
 * keeps the previous call's contents; calls that would overlap on a GPU race on it, so the entry point is synchronous and
 * serialises concurrent callers with a mutex. */
#define CC_PRIM_NOALLOC_MAX_BLOCKS 1024
int cc_prim_satd16x16_noalloc(const uint8_t *a, const uint8_t *b, int n, int32_t *out);

/* Integer-pipe peak of the device, measured: thread-level G-ops/s of a register-resident loop of independent
 * operations on every SM.  kind 0: VABSDIFF4.U8.ACC (4-pixel abs-diff-accumulate, the SAD searches' op), 1: IADD, 2: LOP3,
 * 3: IDP.4A (the SATD / DCT row dot products).  Used by bench.py for the integer roofline. */
int cc_bench_int_peak(int kind, double *gops);

#ifdef __cplusplus
}
#endif
#endif
