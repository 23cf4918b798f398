/*
 * cc_api.c — the C ABI of libceracuda (include/ceracuda.h): handle life cycle, per-frame orchestration
 * (upload -> kernels -> packed symbols in mapped pinned memory -> host CAVLC -> Annex-B access unit),
 * the bitrate latch written by ceracoder's 20 ms balancer tick
 * (/root/reference/src/gst/encoder_control.c:45-57 -> cc_enc_set_bitrate), and test hooks.
 * Host C; all device work goes through the cck_* launch wrappers.
 */
#include "../../include/ceracuda.h"
#include "cc_host.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <pthread.h>

#define CC_MIN_BPS 300000   /* /root/reference/src/core/bitrate_control.h:30 */
#define CC_MAX_BPS 30000000 /* /root/reference/src/core/bitrate_control.h:31 */
#define CC_N_STAGES CC_STAGE_COUNT
#define CC_N_PEV (CC_N_STAGES + 2) /* per stage its start; [CC_N_STAGES]: end of the tail; [+ 1]: end of the entropy coder */

struct cc_enc {
    cc_enc_params p;
    cc_kparams k;
    cc_seq seq;
    cudaStream_t st;             /* upload, search, coding, entropy coding: what the access unit waits for */
    cudaStream_t st2;            /* in-loop filter and reference finishing of the same frame, concurrent with the entropy coder */
    cudaEvent_t ev0, ev1;
    cudaEvent_t ev_enc, ev_tail; /* ev_enc: st -> st2 after k_mb_encode (after k_deblock when the filter runs next to it: then st2 carries the
                                  * entropy coder); ev_tail: end of the tail (either stream) -> st before the next frame touches what it reads */
    void *dev_alloc[48];
    int n_dev_alloc;
    size_t dev_bytes;
    uint8_t *out_host;   /* mapped pinned: packed symbols */
    uint32_t out_cap;
    uint8_t *au;         /* access unit */
    size_t au_cap;
    uint8_t *rbsp;       /* per-slice scratch */
    size_t rbsp_cap;
    cc_cavlc_ctx cav;
    cc_rc rc;
    volatile int pending_bps;
    int frame_idx, gop_pos, frame_num, idr_pic_id, have_ref;
    int next_qp;         /* test hook */
    int inflight, cur_is_idr, cur_qp;
    int64_t cur_pts;
    long launches;
    int last_was_p;
    int profile;                 /* per-kernel CUDA events on */
    cudaEvent_t pevs[2][CC_N_PEV]; /* two sets, alternating per frame: the stages behind ev1 (reference finishing) are read when the set
                                           * comes round again, two frames later, when they are certainly complete */
    int pevs_used[2][CC_N_PEV];
    int pevs_tail[2];            /* set holds unread tail stages */
    int pb;                      /* set of the frame in flight */
    double stage_ms[CC_N_STAGES];
    long stage_n[CC_N_STAGES];
    uint32_t sym_bytes;
    /* slice-writer pool (cc_enc_params.threads > 1): workers + the collecting thread take slices off job_next */
    int n_workers;               /* extra threads */
    pthread_t workers[16];
    pthread_mutex_t job_mu;
    pthread_cond_t job_cv, done_cv;
    unsigned job_gen;
    int job_stop, job_next, job_left, job_ns, job_is_idr, job_err;
    const cc_mbhost *job_mbh; const cc_rowhdr *job_rh; const uint8_t *const *job_rowp;
    const uint32_t *job_slen; const uint8_t *job_sdata; /* device entropy: per-slice RBSP bytes and the slices' data in the mapped buffer */
    uint8_t **slice_out; size_t *slice_len; size_t slice_cap; /* per slice: its NAL unit */
    uint8_t *worker_rbsp[17];    /* scratch RBSP per thread (index n_workers = the collecting thread) */
    float submit_ms;
    double t_submit0;            /* host time at the start of the last cc_enc_submit */
    float dev_ms_hist[4];        /* device latency (submit -> access unit ready) of the last frames: how long collect may sleep */
    unsigned dev_ms_n;
    int sleep_wait;              /* CERACUDA_WAIT=spin turns the sleep off */
    uint8_t *stage_host;         /* pinned staging frame: pageable input (a mapped GstBuffer) is copied here by the calling thread, then uploaded
                                  * asynchronously.  One frame suffices: a handle has one frame in flight and its upload is over by collect */
    size_t stage_bytes;
    long staged_frames;
    unsigned long long tmaps[5 * 16] __attribute__((aligned(64))); /* five CUtensorMap (k_me_l0's sub-pel windows) */
    int debug_sync;
    /* upload of the NEXT frame while this one is coded (cc_enc_encode_sequence, planar I420 in page-locked host or device memory): a second
     * set of source planes, filled on a stream of its own; cc_enc_submit of that frame swaps the sets instead of copying */
    uint8_t *cur_alt[3];
    uint8_t *stage_host_alt;  /* a second pinned staging frame: a pageable frame that runs ahead is copied here */
    cudaStream_t st_up;
    long overlapped_frames; /* frames whose filter was launched next to k_mb_encode */
    int prev_trail;        /* launch order of the frame before */
    int last_trail_seq;    /* cc_kparams.frame_seq of the last frame whose filter ran next to the encoder (0: none yet) */
    int cav_on_st2;        /* the entropy coder of the last frame submitted ran on the second stream */
    int counted;           /* this handle is in g_open_handles */
    int overlap_max;       /* CC_OVERLAP_MAX_HANDLES: the default policy overlaps while the process has at most that many handles open */
    int trail_pct;         /* CC_TRAIL_PCT: k_deblock is released when k_mb_encode is that far through its rows (percent) */
    int overlap;           /* CC_OVERLAP: 1 the in-loop filter runs next to k_mb_encode, 0 after it, -1 (default) by the number of open handles */
    cudaEvent_t ev_up;
    const uint8_t *pf_plane0; /* plane[0] of the frame sitting in cur_alt, or NULL */
    long prefetched_frames;
    int dbg_dup, dbg_skip; /* CC_DBG_DUP / CC_DBG_SKIP: stage bit masks (1 << CC_STAGE_*) — launch a stage twice / not at all (marginal-cost experiments) */
    int bs_dirty;                /* a frame was abandoned: the slice bit buffers may hold stray bits, clear them before the next frame */
    size_t bs_bytes;
};

static double now_ms(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }

const char *cc_version(void) { return "ceracuda 0.1 (sm_100a)"; }
const char *cc_strerror(int c) {
    switch (c) {
    case CC_OK: return "ok";
    case CC_ERR_PARAM: return "invalid parameter";
    case CC_ERR_CUDA: return "CUDA error or no CUDA device";
    case CC_ERR_NOMEM: return "out of memory";
    case CC_ERR_STATE: return "invalid call order";
    case CC_ERR_OVERFLOW: return "output buffer overflow";
    default: return "unknown error";
    }
}

#define CUCHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "ceracuda: %s failed: %s\n", #x, cudaGetErrorString(e_)); return CC_ERR_CUDA; } } while (0)

void cc_enc_params_default(cc_enc_params *p) {
    memset(p, 0, sizeof *p);
    p->fps_n = 30; p->fps_d = 1; p->format = CC_FMT_I420; p->bitrate_bps = 6000000; /* reference default max, bitrate_control.h:32 */
    p->key_int_max = 60; p->rows_per_slice = 8; p->qp_min = 10; p->qp_max = 51; p->partitions = 1;
}

static int dev_alloc(cc_enc *e, void **p, size_t n) {
    if (e->n_dev_alloc >= 48) return CC_ERR_NOMEM;
    if (cudaMalloc(p, n) != cudaSuccess) { fprintf(stderr, "ceracuda: cudaMalloc(%zu) failed\n", n); return CC_ERR_NOMEM; }
    e->dev_alloc[e->n_dev_alloc++] = *p;
    e->dev_bytes += n;
    return CC_OK;
}
static size_t roundup(size_t v, size_t a) { return (v + a - 1) / a * a; }

/* ---- slice-writer pool ------------------------------------------------------------------------------ */
/* Slices are independent (no prediction, no CAVLC context across slice borders; the per-MB total_coeff array is written
 * by the slice that owns the macroblock), so their NAL units can be written concurrently and concatenated in order. */
static int write_slice_nal(cc_enc *e, int s, uint8_t *rbsp, uint8_t *out, size_t cap, size_t *len_out) {
    const cc_kparams *k = &e->k;
    const int r0 = s * k->rps, r1 = r0 + k->rps < k->mb_h ? r0 + k->rps : k->mb_h, is_idr = e->job_is_idr;
    if (k->dev_entropy) { /* the slice's RBSP is already in host memory (k_cavlc_write): start code, NAL header, emulation prevention */
        const uint32_t len = e->job_slen[s];
        if (len == 0 || len > (size_t)k->bs_slice_words * 4) return CC_ERR_STATE;
        *len_out = cc_write_nal(out, cap, is_idr ? 0x65 : 0x41, e->job_sdata + (size_t)s * k->bs_slice_words * 4, len);
        return *len_out ? CC_OK : CC_ERR_OVERFLOW;
    }
    cc_bw b;
    cc_bw_init(&b, rbsp, e->rbsp_cap);
    cc_write_slice_header(&b, r0 * k->mb_w, is_idr, e->frame_num, e->idr_pic_id, e->cur_qp);
    if (cc_cavlc_write_slice(&e->cav, &b, is_idr, r0, r1, e->job_mbh, e->job_rh, e->job_rowp) != 0) return CC_ERR_STATE;
    const size_t m = cc_bw_finish(&b);
    if (b.overflow) return CC_ERR_OVERFLOW;
    *len_out = cc_write_nal(out, cap, is_idr ? 0x65 : 0x41, rbsp, m);
    return *len_out ? CC_OK : CC_ERR_OVERFLOW;
}
static void pool_take_slices(cc_enc *e, int tid) {
    for (;;) {
        const int s = __atomic_fetch_add(&e->job_next, 1, __ATOMIC_ACQUIRE);
        if (s >= e->job_ns) return;
        const int rc = write_slice_nal(e, s, e->worker_rbsp[tid], e->slice_out[s], e->slice_cap, &e->slice_len[s]);
        if (rc != CC_OK) __atomic_store_n(&e->job_err, rc, __ATOMIC_RELAXED);
        if (__atomic_sub_fetch(&e->job_left, 1, __ATOMIC_ACQ_REL) == 0) {
            pthread_mutex_lock(&e->job_mu); pthread_cond_signal(&e->done_cv); pthread_mutex_unlock(&e->job_mu);
        }
    }
}
typedef struct { cc_enc *e; int tid; } pool_arg;
static void *pool_main(void *a) {
    cc_enc *e = ((pool_arg *)a)->e;
    const int tid = ((pool_arg *)a)->tid;
    free(a);
    unsigned seen = 0;
    for (;;) {
        pthread_mutex_lock(&e->job_mu);
        while (!e->job_stop && e->job_gen == seen) pthread_cond_wait(&e->job_cv, &e->job_mu);
        seen = e->job_gen;
        const int stop = e->job_stop;
        pthread_mutex_unlock(&e->job_mu);
        if (stop) return NULL;
        pool_take_slices(e, tid);
    }
}
static int pool_start(cc_enc *e, int threads) {
    const int ns = (e->k.mb_h + e->k.rps - 1) / e->k.rps;
    if (threads > 16) threads = 16;
    if (threads > ns) threads = ns;
    if (threads < 2) return CC_OK;
    pthread_mutex_init(&e->job_mu, NULL); pthread_cond_init(&e->job_cv, NULL); pthread_cond_init(&e->done_cv, NULL);
    e->slice_cap = e->rbsp_cap + e->rbsp_cap / 2 + 64;
    e->slice_out = (uint8_t **)calloc((size_t)ns, sizeof *e->slice_out);
    e->slice_len = (size_t *)calloc((size_t)ns, sizeof *e->slice_len);
    if (!e->slice_out || !e->slice_len) return CC_ERR_NOMEM;
    for (int s = 0; s < ns; s++) if (!(e->slice_out[s] = (uint8_t *)malloc(e->slice_cap))) return CC_ERR_NOMEM;
    for (int t = 0; t < threads; t++) if (!(e->worker_rbsp[t] = (uint8_t *)malloc(e->rbsp_cap))) return CC_ERR_NOMEM;
    for (int t = 0; t < threads - 1; t++) {
        pool_arg *a = (pool_arg *)malloc(sizeof *a);
        if (!a) return CC_ERR_NOMEM;
        a->e = e; a->tid = t;
        if (pthread_create(&e->workers[t], NULL, pool_main, a) != 0) { free(a); return CC_ERR_NOMEM; }
        e->n_workers = t + 1;
    }
    return CC_OK;
}
static void pool_stop(cc_enc *e) {
    if (e->n_workers) {
        pthread_mutex_lock(&e->job_mu); e->job_stop = 1; pthread_cond_broadcast(&e->job_cv); pthread_mutex_unlock(&e->job_mu);
        for (int t = 0; t < e->n_workers; t++) pthread_join(e->workers[t], NULL);
        pthread_mutex_destroy(&e->job_mu); pthread_cond_destroy(&e->job_cv); pthread_cond_destroy(&e->done_cv);
        e->n_workers = 0;
    }
    if (e->slice_out) { const int ns = (e->k.mb_h + e->k.rps - 1) / e->k.rps; for (int s = 0; s < ns; s++) free(e->slice_out[s]); }
    free(e->slice_out); free(e->slice_len); e->slice_out = NULL; e->slice_len = NULL;
    for (int t = 0; t < 17; t++) { free(e->worker_rbsp[t]); e->worker_rbsp[t] = NULL; }
}

/* handles open per device (process-wide): the default overlap policy of cc_enc_submit reads it */
#define CC_MAX_DEVICES 64
static int g_open_handles[CC_MAX_DEVICES];

int cc_enc_open(cc_enc **h, const cc_enc_params *p) {
    if (!h || !p) return CC_ERR_PARAM;
    *h = NULL;
    if (p->width < 16 || p->height < 16 || (p->width & 1) || (p->height & 1) || p->width > 8192 || p->height > 8192) return CC_ERR_PARAM;
    if (p->format < CC_FMT_I420 || p->format > CC_FMT_BGRX) return CC_ERR_PARAM;
    if (p->fps_n <= 0 || p->fps_d <= 0) return CC_ERR_PARAM;
    /* one warp per macroblock row, at most 8 rows (256 threads) per CTA: slices of 1..8 rows are one CTA; taller slices are whole
     * bands of 8 rows (a multiple of 8) or the whole picture (anything >= its macroblock rows) */
    if (p->rows_per_slice < 0 || (p->rows_per_slice > 8 && p->rows_per_slice % 8 != 0 && p->rows_per_slice < (p->height + 15) / 16)) return CC_ERR_PARAM;
    { /* the SPS must be able to name a level that covers the stream: nothing beyond level 5.2 (Table A-1: 36 864 macroblocks per
       * picture, 2 073 600 per second — 2160p60 fits, 8192 x 8192 does not) */
        const long fs = (long)((p->width + 15) / 16) * ((p->height + 15) / 16);
        if (!cc_level_idc_for(fs, (fs * p->fps_n + p->fps_d - 1) / p->fps_d)) {
            fprintf(stderr, "ceracuda: %dx%d at %d/%d fps is beyond H.264 level 5.2\n", p->width, p->height, p->fps_n, p->fps_d);
            return CC_ERR_PARAM;
        }
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || p->device < 0 || p->device >= ndev) {
        fprintf(stderr, "ceracuda: no usable CUDA device (count %d, asked for %d); there is no CPU fallback\n", ndev, p->device);
        return CC_ERR_CUDA;
    }
    CUCHECK(cudaSetDevice(p->device));
    cc_enc *e = (cc_enc *)calloc(1, sizeof *e);
    if (!e) return CC_ERR_NOMEM;
    e->p = *p;
    if (e->p.key_int_max <= 0) e->p.key_int_max = 60;
    if (e->p.rows_per_slice <= 0) e->p.rows_per_slice = 8;
    if (e->p.qp_min <= 0) e->p.qp_min = 10;
    if (e->p.qp_max <= 0 || e->p.qp_max > 51) e->p.qp_max = 51;
    if (e->p.qp_min < 10) e->p.qp_min = 10;
    if (e->p.bitrate_bps <= 0) e->p.bitrate_bps = 6000000;
    cc_kparams *k = &e->k;
    k->w = p->width; k->h = p->height;
    k->mb_w = (p->width + 15) / 16; k->mb_h = (p->height + 15) / 16;
    if (e->p.rows_per_slice > k->mb_h) e->p.rows_per_slice = k->mb_h > 8 ? (k->mb_h + 7) / 8 * 8 : k->mb_h; /* one slice */
    k->W = k->mb_w * 16; k->H = k->mb_h * 16; k->rps = e->p.rows_per_slice;
    k->band = k->rps < 8 ? k->rps : 8;
    k->sy = (int)roundup((size_t)k->W + 2 * CC_PAD, 128);
    k->sc = (int)roundup((size_t)k->W / 2 + 2 * CC_PADC, 64);
    e->seq.w = k->w; e->seq.h = k->h; e->seq.mb_w = k->mb_w; e->seq.mb_h = k->mb_h; e->seq.fps_n = p->fps_n; e->seq.fps_d = p->fps_d;
    const int nmb = k->mb_w * k->mb_h;
    int rc = CC_OK;
    size_t lsz = (size_t)k->sy * (k->H + 2 * CC_PAD) + 256, csz = (size_t)k->sc * (k->H / 2 + 2 * CC_PADC) + 256;
    uint8_t *lum[6], *chr[8];
    for (int i = 0; i < 6 && rc == CC_OK; i++) { rc = dev_alloc(e, (void **)&lum[i], lsz); if (rc == CC_OK) cudaMemset(lum[i], 0, lsz); }
    for (int i = 0; i < 8 && rc == CC_OK; i++) { rc = dev_alloc(e, (void **)&chr[i], csz); if (rc == CC_OK) cudaMemset(chr[i], 0, csz); }
    if (rc == CC_OK) {
        size_t lo = (size_t)CC_PAD * k->sy + CC_PAD, co = (size_t)CC_PADC * k->sc + CC_PADC;
        k->cur[0] = lum[0] + lo; k->rec[0] = lum[1] + lo; k->ref[0] = lum[2] + lo;
        k->hb = lum[3] + lo; k->hh = lum[4] + lo; k->hj = lum[5] + lo;
        k->cur[1] = chr[0] + co; k->cur[2] = chr[1] + co; k->rec[1] = chr[2] + co; k->rec[2] = chr[3] + co;
        k->ref[1] = chr[4] + co; k->ref[2] = chr[5] + co; k->l1cur = chr[6] + co; k->l1ref = chr[7] + co;
    }
    if (rc == CC_OK && p->format == CC_FMT_I420) {
        uint8_t *alt[3] = {NULL, NULL, NULL};
        rc = dev_alloc(e, (void **)&alt[0], lsz);
        if (rc == CC_OK) rc = dev_alloc(e, (void **)&alt[1], csz);
        if (rc == CC_OK) rc = dev_alloc(e, (void **)&alt[2], csz);
        if (rc == CC_OK) {
            cudaMemset(alt[0], 0, lsz); cudaMemset(alt[1], 0, csz); cudaMemset(alt[2], 0, csz);
            e->cur_alt[0] = alt[0] + (size_t)CC_PAD * k->sy + CC_PAD;
            e->cur_alt[1] = alt[1] + (size_t)CC_PADC * k->sc + CC_PADC; e->cur_alt[2] = alt[2] + (size_t)CC_PADC * k->sc + CC_PADC;
        }
    }
    if (rc == CC_OK && p->format == CC_FMT_NV12) rc = dev_alloc(e, (void **)&k->uvstage, (size_t)k->sy * (k->H / 2 + 1));
    if (rc == CC_OK && (p->format == CC_FMT_YUY2 || p->format == CC_FMT_BGRX)) {
        k->pks = (int)roundup((size_t)k->w * (p->format == CC_FMT_YUY2 ? 2 : 4), 128);
        rc = dev_alloc(e, (void **)&k->pkstage, (size_t)k->pks * k->h);
    }
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->me, sizeof(cc_me) * nmb);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->me8, sizeof(cc_me8) * nmb);
    k->part8 = e->p.partitions != 0;
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->mbd, sizeof(cc_mbdev) * nmb);
    if (rc == CC_OK) rc = cck_me_l1_prepare(k);
    if (rc == CC_OK) {
        uint8_t *const planes[5] = {k->rec[0], k->ref[0], k->hb, k->hh, k->hj};
        rc = cck_me_l0_prepare(k, planes, e->tmaps);
    }
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->nbrec, (size_t)20 * nmb);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->mbh, sizeof(cc_mbhost) * nmb);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->rowbuf, (size_t)nmb * CC_ROW_PAYLOAD_PER_MB);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->pre, sizeof(cc_pre) * nmb);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->symlen, sizeof(uint16_t) * nmb);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->park, (size_t)384 * nmb);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->ubot[0], (size_t)k->mb_h * k->sy);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->ubot[1], (size_t)k->mb_h * k->sc);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->ubot[2], (size_t)k->mb_h * k->sc);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->uright, (size_t)nmb * 32);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->rowhdr, sizeof(cc_rowhdr) * k->mb_h);
    if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->gprog, sizeof(uint32_t) * 3 * k->mb_h);
    if (rc == CC_OK) cudaMemset(k->gprog, 0, sizeof(uint32_t) * 3 * k->mb_h);
    k->dev_entropy = !e->p.host_entropy;
    const int n_slices_open = (k->mb_h + k->rps - 1) / k->rps;
    if (k->dev_entropy) {
        k->bs_slice_words = (uint32_t)roundup((size_t)k->rps * k->mb_w * (CC_BS_BYTES_PER_MB / 4) + 16, 4);
        e->bs_bytes = (size_t)n_slices_open * k->bs_slice_words * 4;
        if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->mbbits, sizeof(uint32_t) * nmb);
        if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->cvrec, (size_t)32 * nmb);
        if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->bs, e->bs_bytes);
        if (rc == CC_OK) rc = dev_alloc(e, (void **)&k->bs_ctl, sizeof(uint32_t) * 2 * n_slices_open);
        if (rc == CC_OK) { cudaMemset(k->bs, 0, e->bs_bytes); cudaMemset(k->bs_ctl, 0, sizeof(uint32_t) * 2 * n_slices_open); }
    }
    if (rc == CC_OK && getenv("CC_DEBUG_CYCLES")) { rc = dev_alloc(e, (void **)&k->cyc, (size_t)k->mb_h * CC_CYC_N * 8); if (rc == CC_OK) cudaMemset(k->cyc, 0, (size_t)k->mb_h * CC_CYC_N * 8); }
    if (rc == CC_OK) { cudaMemset(k->me, 0, sizeof(cc_me) * nmb); cudaMemset(k->me8, 0, sizeof(cc_me8) * nmb); cudaMemset(k->mbd, 0, sizeof(cc_mbdev) * nmb); }
    /* packed symbol buffer: headers + per-MB records + payload (worst case) */
    e->out_cap = (uint32_t)(roundup(sizeof(cc_rowhdr) * k->mb_h, 16) + roundup(sizeof(cc_mbhost) * nmb, 16) + (size_t)nmb * CC_ROW_PAYLOAD_PER_MB + 16 * k->mb_h);
    /* device entropy: [row headers][slice byte counts][slice RBSPs at fixed distances] */
    if (k->dev_entropy) e->out_cap = (uint32_t)(roundup(sizeof(cc_rowhdr) * k->mb_h, 16) + roundup(sizeof(uint32_t) * n_slices_open, 16) + e->bs_bytes);
    if (rc == CC_OK && cudaHostAlloc((void **)&e->out_host, e->out_cap, cudaHostAllocMapped) != cudaSuccess) rc = CC_ERR_NOMEM;
    if (rc == CC_OK && cudaHostGetDevicePointer((void **)&k->out_host_dev, e->out_host, 0) != cudaSuccess) rc = CC_ERR_CUDA;
    k->out_cap = e->out_cap;
    e->au_cap = (size_t)nmb * CC_BS_BYTES_PER_MB * 3 / 2 + 4096; /* every slice's worst case, with emulation prevention */
    e->rbsp_cap = (size_t)k->rps * k->mb_w * CC_BS_BYTES_PER_MB + 4096;
    e->stage_bytes = (size_t)k->w * k->h * (p->format == CC_FMT_BGRX ? 4 : (p->format == CC_FMT_YUY2 ? 2 : 1)) * (p->format <= CC_FMT_NV12 ? 3 : 2) / 2;
    if (rc == CC_OK && cudaHostAlloc((void **)&e->stage_host, e->stage_bytes, cudaHostAllocDefault) != cudaSuccess) rc = CC_ERR_NOMEM;
    if (rc == CC_OK && e->p.format == CC_FMT_I420 && cudaHostAlloc((void **)&e->stage_host_alt, e->stage_bytes, cudaHostAllocDefault) != cudaSuccess) rc = CC_ERR_NOMEM;
    e->au = (uint8_t *)malloc(e->au_cap);
    e->rbsp = (uint8_t *)malloc(e->rbsp_cap);
    e->cav.mb_w = k->mb_w; e->cav.mb_h = k->mb_h; e->cav.rps = k->rps;
    e->cav.nnz = (uint8_t *)calloc((size_t)nmb, 24);
    if (!e->au || !e->rbsp || !e->cav.nnz) rc = CC_ERR_NOMEM;
    if (rc == CC_OK) rc = pool_start(e, p->threads);
    { /* experiment knob CC_PRIO: bit 0 = the coding stream, bit 1 = the reference-finishing stream at the highest priority */
        int lo = 0, hi = 0, pr = getenv("CC_PRIO") ? atoi(getenv("CC_PRIO")) : 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        if (rc == CC_OK && cudaStreamCreateWithPriority(&e->st, cudaStreamNonBlocking, (pr & 1) ? hi : lo) != cudaSuccess) rc = CC_ERR_CUDA;
        if (rc == CC_OK && cudaStreamCreateWithPriority(&e->st2, cudaStreamNonBlocking, (pr & 2) ? hi : lo) != cudaSuccess) rc = CC_ERR_CUDA;
    }
    if (rc == CC_OK && cudaStreamCreateWithFlags(&e->st_up, cudaStreamNonBlocking) != cudaSuccess) rc = CC_ERR_CUDA;
    if (rc == CC_OK && cudaEventCreateWithFlags(&e->ev_up, cudaEventDisableTiming) != cudaSuccess) rc = CC_ERR_CUDA;
    if (rc == CC_OK && (cudaEventCreateWithFlags(&e->ev_enc, cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&e->ev_tail, cudaEventDisableTiming) != cudaSuccess)) rc = CC_ERR_CUDA;
    if (rc == CC_OK && (cudaEventCreate(&e->ev0) != cudaSuccess || cudaEventCreate(&e->ev1) != cudaSuccess)) rc = CC_ERR_CUDA;
    for (int i = 0; i < 2 * CC_N_PEV && rc == CC_OK; i++) if (cudaEventCreate(&e->pevs[i / CC_N_PEV][i % CC_N_PEV]) != cudaSuccess) rc = CC_ERR_CUDA;
    if (rc != CC_OK) { cc_enc_close(e); return rc; }
    e->pending_bps = e->p.bitrate_bps;
    { const char *v = getenv("CERACUDA_WAIT"); e->sleep_wait = !(v && !strcmp(v, "spin")); }
    e->debug_sync = getenv("CC_DEBUG_SYNC") != NULL;
    e->trail_pct = getenv("CC_TRAIL_PCT") ? atoi(getenv("CC_TRAIL_PCT")) : 70;
    if (e->trail_pct < 0) e->trail_pct = 0;
    e->overlap = getenv("CC_OVERLAP") ? atoi(getenv("CC_OVERLAP")) : -1;
    e->overlap_max = getenv("CC_OVERLAP_MAX_HANDLES") ? atoi(getenv("CC_OVERLAP_MAX_HANDLES")) : 8;
    e->dbg_dup = getenv("CC_DBG_DUP") ? atoi(getenv("CC_DBG_DUP")) : 0;
    e->dbg_skip = getenv("CC_DBG_SKIP") ? atoi(getenv("CC_DBG_SKIP")) : 0;
    { const char *v = getenv("CC_DBG_ME"); e->k.dbg = v ? atoi(v) : 0; }
    cc_rc_init(&e->rc, (double)p->fps_n / p->fps_d, e->p.bitrate_bps, e->p.qp_min, e->p.qp_max, e->p.key_int_max);
    cudaStreamSynchronize(e->st);
    cudaDeviceSynchronize();
    if (rc == CC_OK) cudaEventRecord(e->ev_tail, e->st2); /* so that the first frame's wait finds a recorded event */
    __atomic_add_fetch(&g_open_handles[e->p.device & (CC_MAX_DEVICES - 1)], 1, __ATOMIC_RELAXED);
    e->counted = 1;
    *h = e;
    return CC_OK;
}

/* every entry point may be called from a thread whose current device is another GPU (one process driving several devices,
 * or worker threads created after cc_enc_open): bind the handle's device first */
static int bind_device(const cc_enc *e) { return cudaSetDevice(e->p.device) == cudaSuccess ? CC_OK : CC_ERR_CUDA; }

void cc_enc_close(cc_enc *e) {
    if (!e) return;
    if (e->counted) { __atomic_sub_fetch(&g_open_handles[e->p.device & (CC_MAX_DEVICES - 1)], 1, __ATOMIC_RELAXED); e->counted = 0; }
    bind_device(e);
    if (e->st) cudaStreamSynchronize(e->st);
    if (e->st2) cudaStreamSynchronize(e->st2);
    if (e->st_up) cudaStreamSynchronize(e->st_up);
    if (e->k.cyc && e->frame_idx > 0) { /* CC_DEBUG_CYCLES=1: where the wavefront kernel's warps spent their cycles */
        size_t n = (size_t)e->k.mb_h * CC_CYC_N;
        unsigned long long *h = (unsigned long long *)calloc(n, 8), s[CC_CYC_N] = {0};
        if (h && cudaMemcpy(h, e->k.cyc, n * 8, cudaMemcpyDeviceToHost) == cudaSuccess) {
            for (size_t i = 0; i < n; i++) s[i % CC_CYC_N] += h[i];
            double f = 1.0 / ((double)e->frame_idx * e->k.mb_h);
            fprintf(stderr, "ceracuda cycles per MB row per frame: encode wait %.0f fast %.0f (n %.1f) probe %.0f (n %.1f) intra %.0f (n %.1f) | filter wait-enc %.0f wait-up %.0f work %.0f\n",
                    s[0] * f, s[1] * f, s[4] * f, s[2] * f, s[5] * f, s[3] * f, s[6] * f, s[8] * f, s[9] * f, s[10] * f);
            fprintf(stderr, "ceracuda intra path, cycles per MB row per frame: neighbours+i16 analysis %.0f, i4 analysis %.0f (n %.1f), encode+chroma+publish %.0f, emit %.0f\n",
                    s[11] * f, s[12] * f, s[13] * f, s[14] * f, s[15] * f);
            fprintf(stderr, "ceracuda wavefront detail, per MB row per frame: between macroblocks (record / pixel prefetch) %.0f cycles; P_Skip reconstructed in the wavefront: %.1f MBs, %.0f cycles (%.1f of them predicted here); parked P_L0_16x16 pixels restored %.1f; Intra4x4 analyses started %.1f\n",
                    s[16] * f, s[18] * f, s[17] * f, s[19] * f, s[21] * f, s[20] * f);
            fprintf(stderr, "ceracuda P_Skip predicted in the wavefront without a probe, per MB row per frame: at the search vector (speculation lost) %.1f, at the zero vector %.1f (%.1f of them with the candidate probe also passing), other %.1f\n",
                    s[22] * f, s[23] * f, s[25] * f, s[24] * f);
            if (s[28]) fprintf(stderr, "ceracuda intra macroblocks of P frames: %.2f per MB row per frame, %.0f cycles each to code (transform, chroma, publish), %.0f to emit\n",
                               s[28] * f, (double)s[26] / s[28], (double)s[27] / s[28]);
        }
        free(h);
    }
    for (int i = 0; i < e->n_dev_alloc; i++) cudaFree(e->dev_alloc[i]);
    pool_stop(e);
    if (e->out_host) cudaFreeHost(e->out_host);
    if (e->stage_host) cudaFreeHost(e->stage_host);
    if (e->stage_host_alt) cudaFreeHost(e->stage_host_alt);
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    for (int i = 0; i < 2 * CC_N_PEV; i++) if (e->pevs[i / CC_N_PEV][i % CC_N_PEV]) cudaEventDestroy(e->pevs[i / CC_N_PEV][i % CC_N_PEV]);
    if (e->ev_enc) cudaEventDestroy(e->ev_enc);
    if (e->ev_tail) cudaEventDestroy(e->ev_tail);
    if (e->st) cudaStreamDestroy(e->st);
    if (e->st2) cudaStreamDestroy(e->st2);
    if (e->ev_up) cudaEventDestroy(e->ev_up);
    if (e->st_up) cudaStreamDestroy(e->st_up);
    free(e->au); free(e->rbsp); free(e->cav.nnz);
    free(e);
}

int cc_enc_set_bitrate(cc_enc *e, int bps) {
    if (!e) return CC_ERR_PARAM;
    if (bps < CC_MIN_BPS) bps = CC_MIN_BPS;
    if (bps > CC_MAX_BPS) bps = CC_MAX_BPS;
    __atomic_store_n(&e->pending_bps, bps, __ATOMIC_RELAXED);
    return CC_OK;
}
int cc_enc_get_bitrate(const cc_enc *e) { return e ? __atomic_load_n(&e->pending_bps, __ATOMIC_RELAXED) : CC_ERR_PARAM; }
int cc_enc_set_next_qp(cc_enc *e, int qp) { if (!e || qp < 10 || qp > 51) return CC_ERR_PARAM; e->next_qp = qp; return CC_OK; }
int cc_enc_mb_w(const cc_enc *e) { return e ? e->k.mb_w : CC_ERR_PARAM; }
int cc_enc_mb_h(const cc_enc *e) { return e ? e->k.mb_h : CC_ERR_PARAM; }
size_t cc_enc_device_bytes(const cc_enc *e) { return e ? e->dev_bytes : 0; }
long cc_enc_kernel_launches(const cc_enc *e) { return e ? e->launches : 0; }
int cc_enc_set_profile(cc_enc *e, int on) {
    if (!e || e->inflight) return CC_ERR_PARAM;
    e->profile = on != 0;
    memset(e->stage_ms, 0, sizeof e->stage_ms); memset(e->stage_n, 0, sizeof e->stage_n);
    e->pevs_tail[0] = e->pevs_tail[1] = 0;
    return CC_OK;
}
int cc_enc_get_stage_ms(const cc_enc *e, double *ms_sum, long *count) {
    if (!e || !ms_sum || !count) return CC_ERR_PARAM;
    for (int i = 0; i < CC_N_STAGES; i++) { ms_sum[i] = e->stage_ms[i]; count[i] = e->stage_n[i]; }
    return CC_OK;
}

/* page-locked (cudaHostAlloc / cudaHostRegister) host memory? */
static int host_is_pinned(const void *p) {
    struct cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return a.type == cudaMemoryTypeHost;
}
long cc_enc_staged_frames(const cc_enc *e) { return e ? e->staged_frames : 0; }

static void set_quant(cc_kparams *k) {
    static const uint8_t chroma_qp[52] = {0,  1,  2,  3,  4,  5,  6,  7,  8,  9,  10, 11, 12, 13, 14, 15, 16, 17,
                                          18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 29, 30, 31, 32, 32, 33,
                                          34, 34, 35, 35, 36, 36, 37, 37, 37, 38, 38, 38, 39, 39, 39, 39}; /* Table 8-15 */
    cc_quant_tables(k->qp, 1, k->qmf[0], k->qbias[0]);
    cc_quant_tables(k->qp, 0, k->qmf[1], k->qbias[1]);
    cc_quant_tables(chroma_qp[k->qp], 1, k->qmf[2], k->qbias[2]);
    cc_quant_tables(chroma_qp[k->qp], 0, k->qmf[3], k->qbias[3]);
    cc_quant8_tables(k->qp, k->q8mf, k->q8bias);
    /* level = ((bias + |c|) * mf) >> 16 >= n  <=>  |c| >= ceil(n * 65536 / mf) - bias */
    for (int t = 0; t < 4; t++)
        for (int c = 0; c < 3; c++)
            for (int n = 1; n <= 2; n++) {
                const int mf = k->qmf[t][c], thr = (n * 65536 + mf - 1) / mf - k->qbias[t][c];
                k->qthr[t][c][n - 1] = (uint16_t)(thr < 0 ? 0 : (thr > 65535 ? 65535 : thr));
            }
}

int cc_enc_submit(cc_enc *e, const cc_frame *in, unsigned flags) {
    if (!e || !in) return CC_ERR_PARAM;
    if (e->inflight) return CC_ERR_STATE;
    if (bind_device(e) != CC_OK) return CC_ERR_CUDA;
    const double t_submit = now_ms();
    cc_kparams *k = &e->k;
    const int nv12 = e->p.format == CC_FMT_NV12, fmt = e->p.format;
    const int packed = fmt == CC_FMT_YUY2 || fmt == CC_FMT_BGRX;
    if (!in->plane[0]) return CC_ERR_PARAM;
    if (packed) {
        if (in->stride[0] < k->w * (fmt == CC_FMT_YUY2 ? 2 : 4)) return CC_ERR_PARAM;
    } else {
        if (!in->plane[1] || (!nv12 && !in->plane[2])) return CC_ERR_PARAM;
        if (in->stride[0] < k->w || in->stride[1] < (nv12 ? k->w : k->w / 2)) return CC_ERR_PARAM;
    }
    int is_idr = !e->have_ref || (flags & CC_FORCE_IDR) || e->gop_pos >= e->p.key_int_max;
    if (is_idr) { e->gop_pos = 0; e->frame_num = 0; }
    int bps = __atomic_load_n(&e->pending_bps, __ATOMIC_RELAXED);
    int qp = e->p.qp_fixed > 0 ? e->p.qp_fixed : cc_rc_frame_qp(&e->rc, bps, is_idr, k->mb_w * k->mb_h);
    if (e->next_qp) { qp = e->next_qp; e->next_qp = 0; }
    if (qp < 10) qp = 10;
    if (qp > 51) qp = 51;
    k->qp = qp; k->is_idr = is_idr;
    k->frame_num = e->frame_num; k->idr_pic_id = e->idr_pic_id;
    k->frame_seq = (k->frame_seq + 1) & 0xfffff; if (!k->frame_seq) k->frame_seq = 1;
    set_quant(k);
    e->cur_is_idr = is_idr; e->cur_qp = qp; e->cur_pts = in->pts;

    enum cudaMemcpyKind kind = (flags & CC_FRAME_DEVICE) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    e->pb ^= 1;
    cudaEvent_t *pev = e->pevs[e->pb];
    int *pev_used = e->pevs_used[e->pb];
    if (e->pevs_tail[e->pb]) { /* the stages that ran behind the collect of the frame before the previous one */
        for (int i = CC_STAGE_DEBLOCK; i <= CC_STAGE_L1REF; i++) {
            float ms = 0;
            if (pev_used[i] && cudaEventElapsedTime(&ms, pev[i], i == CC_STAGE_L1REF ? pev[CC_N_STAGES] : pev[i + 1]) == cudaSuccess) { e->stage_ms[i] += ms; e->stage_n[i]++; }
        }
        e->pevs_tail[e->pb] = 0;
    }
    if (e->bs_dirty && k->dev_entropy) {
        const int nsl = (k->mb_h + k->rps - 1) / k->rps;
        CUCHECK(cudaMemsetAsync(k->bs, 0, e->bs_bytes, e->st));
        CUCHECK(cudaMemsetAsync(k->bs_ctl, 0, sizeof(uint32_t) * 2 * nsl, e->st));
        e->bs_dirty = 0;
    }
    CUCHECK(cudaEventRecord(e->ev0, e->st));
    if (e->profile) { cudaEventRecord(pev[CC_STAGE_UPLOAD], e->st); }
    /* Host frames that are not page-locked (what gst_video_frame_map hands the element) cannot be copied asynchronously: the
     * runtime would stage them through its own bounce buffer inside the call.  Copy them into the handle's pinned frame here
     * (tight rows), then upload from there like a pinned frame. */
    const uint8_t *pl[3] = {in->plane[0], in->plane[1], in->plane[2]};
    int st[3] = {in->stride[0], in->stride[1], in->stride[2]};
    if (e->pf_plane0 && e->pf_plane0 != in->plane[0]) e->pf_plane0 = NULL; /* another frame than the one announced: the upload is dropped */
    if (!e->pf_plane0 && !(flags & CC_FRAME_DEVICE) && !host_is_pinned(in->plane[0])) {
        uint8_t *d = e->stage_host;
        const int nplanes = packed ? 1 : (nv12 ? 2 : 3);
        for (int i = 0; i < nplanes; i++) {
            const size_t rowb = packed ? (size_t)k->w * (fmt == CC_FMT_YUY2 ? 2 : 4) : (i == 0 ? (size_t)k->w : (nv12 ? (size_t)k->w : (size_t)k->w / 2));
            const size_t rows = i == 0 ? (size_t)k->h : (size_t)k->h / 2;
            if ((size_t)in->stride[i] == rowb) memcpy(d, in->plane[i], rowb * rows);
            else for (size_t r = 0; r < rows; r++) memcpy(d + r * rowb, in->plane[i] + r * (size_t)in->stride[i], rowb);
            pl[i] = d; st[i] = (int)rowb;
            d += rowb * rows;
        }
        e->staged_frames++;
    }
    if (e->pf_plane0 && e->pf_plane0 == in->plane[0] && fmt == CC_FMT_I420) { /* this frame was uploaded while the previous one was coded */
        for (int i = 0; i < 3; i++) { uint8_t *t = k->cur[i]; k->cur[i] = e->cur_alt[i]; e->cur_alt[i] = t; }
        CUCHECK(cudaStreamWaitEvent(e->st, e->ev_up, 0));
        e->pf_plane0 = NULL;
        e->prefetched_frames++;
    } else if (packed) {
        CUCHECK(cudaMemcpy2DAsync(k->pkstage, (size_t)k->pks, pl[0], (size_t)st[0], (size_t)k->w * (fmt == CC_FMT_YUY2 ? 2 : 4), (size_t)k->h, kind, e->st));
    } else {
        CUCHECK(cudaMemcpy2DAsync(k->cur[0], (size_t)k->sy, pl[0], (size_t)st[0], (size_t)k->w, (size_t)k->h, kind, e->st));
        if (nv12) {
            CUCHECK(cudaMemcpy2DAsync(k->uvstage, (size_t)k->sy, pl[1], (size_t)st[1], (size_t)k->w, (size_t)k->h / 2, kind, e->st));
        } else {
            CUCHECK(cudaMemcpy2DAsync(k->cur[1], (size_t)k->sc, pl[1], (size_t)st[1], (size_t)k->w / 2, (size_t)k->h / 2, kind, e->st));
            CUCHECK(cudaMemcpy2DAsync(k->cur[2], (size_t)k->sc, pl[2], (size_t)st[2], (size_t)k->w / 2, (size_t)k->h / 2, kind, e->st));
        }
    }
    /* CC_DEBUG_SYNC=1: wait after every stage and name the one that faulted (compute-sanitizer is not always available) */
#define STAGE(i) do { if (e->debug_sync) { cudaError_t er_ = cudaStreamSynchronize(e->st); if (er_ == cudaSuccess) er_ = cudaStreamSynchronize(e->st2); \
                                           if (er_ != cudaSuccess) { fprintf(stderr, "ceracuda: device fault before stage %d of frame %d: %s\n", (int)(i), e->frame_idx, cudaGetErrorString(er_)); return CC_ERR_CUDA; } } \
                      if (e->profile) { cudaEventRecord(pev[i], e->st); pev_used[i] = 1; } } while (0)
    for (int i = 0; i < CC_N_PEV; i++) pev_used[i] = 0;
#define DUP(i) (e->dbg_dup & (1 << (i)))
    STAGE(CC_STAGE_INGEST);
    e->launches += cck_ingest(k, fmt, e->st);
    if (DUP(CC_STAGE_INGEST)) cck_ingest(k, fmt, e->st);
    /* Upload, ingest and the coarse search touch the source planes and the half-resolution reference, which the coding kernels of the
     * previous frame wrote (unfiltered reconstruction): they run while that frame's tail is still filtering.  Everything after
     * them reads what the tail (in-loop filter, borders, half-pel planes — on st2) produces, or overwrites what it reads (the
     * neighbour records). */
    STAGE(CC_STAGE_ME_COARSE);
    if (!is_idr) { e->launches += cck_l1border(k, e->st); e->launches += cck_me_l1(k, e->st); if (DUP(CC_STAGE_ME_COARSE)) cck_me_l1(k, e->st); }
    STAGE(CC_STAGE_WAIT_REF);
    CUCHECK(cudaStreamWaitEvent(e->st, e->ev_tail, 0));
    STAGE(CC_STAGE_ME_FINE);
    if (!is_idr) { e->launches += cck_me_l0(k, e->st); if (DUP(CC_STAGE_ME_FINE)) cck_me_l0(k, e->st); }
    /* The in-loop filter runs NEXT TO k_mb_encode instead of after it (k->trail): k_deblock is the next kernel in the same stream,
     * launched with programmatic stream serialisation — it begins when every CTA of k_mb_encode has come trail_at macroblocks into
     * its rows — and follows the encoder's progress words macroblock by macroblock.  The tail (filter, half-pel planes) therefore
     * sits on the first stream, behind k_mb_encode, and the entropy coder moves to the second.  This takes most of the filter out
     * of the cycle me_l0 -> p_code -> mb_encode -> deblock -> hpel that bounds ONE stream's frame rate (1 stream: +18 % frames/s;
     * 8: +1.5 %); it buys nothing once enough streams keep every SM busy anyway, where the waiting filter CTAs cost 2 % (16 streams)
     * — hence the default: on while the device has at most overlap_max handles open in this process.  Off while profiling, so that the per-stage
     * times stay times of stages. */
    const int ovl = e->profile || e->debug_sync ? 0 : (e->overlap >= 0 ? e->overlap & 1 : __atomic_load_n(&g_open_handles[e->p.device & (CC_MAX_DEVICES - 1)], __ATOMIC_RELAXED) <= e->overlap_max);
    k->trail = ovl;
    k->trail_at = k->mb_w * e->trail_pct / 100;
    if (k->trail_at > k->mb_w - 1) k->trail_at = k->mb_w - 1;
    /* the entropy coder of the frame before (second stream when the filter trails) reads the symbol slots the pre-pass is about to overwrite */
    if (e->cav_on_st2) CUCHECK(cudaStreamWaitEvent(e->st, e->ev1, 0));
    e->cav_on_st2 = k->trail;
    e->overlapped_frames += k->trail;
    if (k->trail) {
        /* k_mb_encode keeps the progress words of its bands' bottom rows current only in this launch order, and k_deblock takes a word
         * for this frame's when its tag is this frame's (20 bits): a word left by the last frame of an earlier stretch of this order
         * must not carry it — which it could after exactly 2^20 - 1 frames of the other order */
        if (!e->prev_trail && k->frame_seq == e->last_trail_seq) { k->frame_seq = (k->frame_seq + 1) & 0xfffff; if (!k->frame_seq) k->frame_seq = 1; }
        e->last_trail_seq = k->frame_seq;
    }
    e->prev_trail = k->trail;
    STAGE(CC_STAGE_P_CODE);
    if (!is_idr) {
        e->launches += cck_p_code(k, e->st);
        if (DUP(CC_STAGE_P_CODE)) cck_p_code(k, e->st);
    }
    STAGE(CC_STAGE_MB_ENCODE);
    e->launches += cck_encode(k, e->st);
    if (DUP(CC_STAGE_MB_ENCODE) && !k->trail) cck_encode(k, e->st);
#define STAGE2(i) do { if (e->profile) { cudaEventRecord(pev[i], tail_st); pev_used[i] = 1; } } while (0)
    cudaStream_t tail_st, cav_st;
    if (k->trail) { /* st: k_mb_encode, the filter next to it, the half-pel planes; st2: the entropy coder once the filter (hence the encoder) is done */
        tail_st = e->st; cav_st = e->st2;
        if (!(e->dbg_skip & (1 << CC_STAGE_DEBLOCK))) e->launches += cck_deblock(k, e->st, 1);
        CUCHECK(cudaEventRecord(e->ev_enc, e->st));
        CUCHECK(cudaStreamWaitEvent(e->st2, e->ev_enc, 0));
    } else {    /* st: k_mb_encode, the entropy coder; st2: the tail, next to the entropy coder */
        tail_st = e->st2; cav_st = e->st;
        CUCHECK(cudaEventRecord(e->ev_enc, e->st));
        CUCHECK(cudaStreamWaitEvent(e->st2, e->ev_enc, 0));
    }
    if (e->profile) { cudaEventRecord(pev[CC_STAGE_PACK], cav_st); pev_used[CC_STAGE_PACK] = 1; }
    e->launches += k->dev_entropy ? cck_cavlc(k, cav_st) : cck_pack(k, cav_st);
    /* the access unit's bits (or symbols) are in host memory here: cc_enc_collect waits for ev1 only */
    CUCHECK(cudaEventRecord(e->ev1, cav_st));
    if (e->profile) { cudaEventRecord(pev[CC_N_STAGES + 1], cav_st); pev_used[CC_N_STAGES + 1] = 1; }
    /* The kernels that turn the reconstruction into the next frame's reference need nothing from the entropy coder and nothing it
     * writes: they run on the other stream next to it, and then on while the host packages the access unit. */
    STAGE2(CC_STAGE_DEBLOCK);
    if (!k->trail && !(e->dbg_skip & (1 << CC_STAGE_DEBLOCK))) e->launches += cck_deblock(k, tail_st, 0);
    STAGE2(CC_STAGE_BORDER); /* no launch of its own any more: the border CTAs ride behind the half-pel tiles */
    STAGE2(CC_STAGE_HPEL);
    e->launches += cck_hpel_border(k, tail_st);
    if (DUP(CC_STAGE_HPEL)) cck_hpel_border(k, tail_st);
    STAGE2(CC_STAGE_L1REF); /* nothing left here: the half-resolution reference is written with the reconstruction */
    STAGE2(CC_N_STAGES);
#undef STAGE2
#undef DUP
    CUCHECK(cudaEventRecord(e->ev_tail, tail_st));
#undef STAGE
    e->pevs_tail[e->pb] = e->profile;
    CUCHECK(cudaGetLastError());
    /* the new reconstruction becomes the reference of the next frame */
    for (int i = 0; i < 3; i++) { uint8_t *t = k->rec[i]; k->rec[i] = k->ref[i]; k->ref[i] = t; }
    e->have_ref = 1;
    e->last_was_p = !is_idr;
    e->inflight = 1;
    e->submit_ms = (float)(now_ms() - t_submit);
    e->t_submit0 = t_submit;
    return CC_OK;
}

static int is_idr_prof(const cc_enc *e) { return e->cur_is_idr; }

static int collect_inner(cc_enc *e, cc_packet *out);
int cc_enc_collect(cc_enc *e, cc_packet *out) {
    if (!e || !out) return CC_ERR_PARAM;
    if (!e->inflight) return CC_ERR_STATE;
    if (bind_device(e) != CC_OK) return CC_ERR_CUDA;
    const int rc = collect_inner(e, out);
    if (rc != CC_OK) {
        /* The frame's reconstruction already replaced the reference in cc_enc_submit, but no access unit leaves the encoder:
         * a following P frame would predict from a picture the decoder never received.  Restart the stream with an IDR. */
        e->inflight = 0;
        e->have_ref = 0;
        e->bs_dirty = 1;
    }
    return rc;
}
static int collect_inner(cc_enc *e, cc_packet *out) {
    cc_kparams *k = &e->k;
    const double t_wait = now_ms();
    /* The device needs ~2 ms per frame and the driver's wait spins: with one host thread per stream and several streams
     * (or several GPUs' worth of streams) per box the spinning threads take the cores the other streams' bitstream
     * writers need.  Sleep through 80 % of what is left of the shortest recent device latency, then let the driver spin. */
    if (e->sleep_wait && e->dev_ms_n >= 4) {
        float lo = e->dev_ms_hist[0];
        for (int i = 1; i < 4; i++) if (e->dev_ms_hist[i] < lo) lo = e->dev_ms_hist[i];
        const double left = (double)lo - (t_wait - e->t_submit0);
        if (left > 0.3) {
            struct timespec ts = {0, (long)(left * 0.8e6)};
            nanosleep(&ts, NULL);
        }
    }
    CUCHECK(cudaEventSynchronize(e->ev1));
    out->wait_ms = (float)(now_ms() - t_wait); out->submit_ms = e->submit_ms;
    e->inflight = 0;
    float gpu_ms = 0;
    cudaEventElapsedTime(&gpu_ms, e->ev0, e->ev1);
    e->dev_ms_hist[e->dev_ms_n++ & 3] = gpu_ms;
    if (e->profile)
        for (int i = 0; i < CC_STAGE_DEBLOCK; i++) {
            float ms = 0;
            cudaEvent_t end = i == CC_STAGE_PACK ? e->pevs[e->pb][CC_N_STAGES + 1] : (i == CC_STAGE_ME_COARSE ? e->pevs[e->pb][CC_STAGE_WAIT_REF] : e->pevs[e->pb][i + 1]);
            if ((i == CC_STAGE_UPLOAD || e->pevs_used[e->pb][i]) && cudaEventElapsedTime(&ms, e->pevs[e->pb][i], end) == cudaSuccess) {
                if (!is_idr_prof(e) || (i != CC_STAGE_ME_COARSE && i != CC_STAGE_ME_FINE && i != CC_STAGE_P_CODE)) { e->stage_ms[i] += ms; e->stage_n[i]++; }
            }
        }
    if (e->profile && e->pevs_used[e->pb][CC_STAGE_WAIT_REF]) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, e->pevs[e->pb][CC_STAGE_WAIT_REF], e->pevs[e->pb][CC_STAGE_ME_FINE]) == cudaSuccess) { e->stage_ms[CC_STAGE_WAIT_REF] += ms; e->stage_n[CC_STAGE_WAIT_REF]++; }
    }
    double t0 = now_ms();
    const int nmb = k->mb_w * k->mb_h, is_idr = e->cur_is_idr;
    const cc_rowhdr *rh = (const cc_rowhdr *)e->out_host;
    const cc_mbhost *mbh = (const cc_mbhost *)(e->out_host + roundup(sizeof(cc_rowhdr) * k->mb_h, 16));
    const uint8_t *pay = (const uint8_t *)mbh + roundup(sizeof(cc_mbhost) * nmb, 16);
    const uint8_t *rowp[512];
    if (k->mb_h > 512) return CC_ERR_PARAM;
    const int n_slices = (k->mb_h + k->rps - 1) / k->rps;
    int ns = 0, ni = 0, np = 0, ng = 0, npr = 0;
    size_t dev_bytes_out = 0;
    for (int r = 0; r < k->mb_h; r++) {
        ns += (int)rh[r].n_skip; np += (int)rh[r].n_inter; ni += (int)rh[r].n_intra; ng += (int)rh[r].n_gate; npr += (int)rh[r].n_probe;
        if (k->dev_entropy) continue;
        rowp[r] = pay;
        pay += roundup(rh[r].bytes, 16);
        if ((size_t)(pay - e->out_host) > e->out_cap) return CC_ERR_OVERFLOW;
    }
    if (k->dev_entropy) {
        e->job_slen = (const uint32_t *)(e->out_host + roundup(sizeof(cc_rowhdr) * k->mb_h, 16));
        e->job_sdata = (const uint8_t *)e->job_slen + roundup(sizeof(uint32_t) * n_slices, 16);
        dev_bytes_out = roundup(sizeof(cc_rowhdr) * k->mb_h, 16) + roundup(sizeof(uint32_t) * n_slices, 16);
        for (int s = 0; s < n_slices; s++) dev_bytes_out += roundup(e->job_slen[s], 16);
    } else dev_bytes_out = (size_t)(pay - e->out_host);
    size_t n = 0;
    uint8_t hdr[64];
    if (is_idr) {
        size_t m = cc_write_sps(&e->seq, hdr, sizeof hdr);
        size_t w = cc_write_nal(e->au + n, e->au_cap - n, 0x67, hdr, m);
        if (!m || !w) return CC_ERR_OVERFLOW;
        n += w;
        m = cc_write_pps(hdr, sizeof hdr);
        w = cc_write_nal(e->au + n, e->au_cap - n, 0x68, hdr, m);
        if (!m || !w) return CC_ERR_OVERFLOW;
        n += w;
    }
    e->job_is_idr = is_idr; e->job_mbh = mbh; e->job_rh = rh; e->job_rowp = rowp;
    if (e->n_workers) { /* cc_enc_params.threads > 1: the slices are written concurrently, then joined in order */
        pthread_mutex_lock(&e->job_mu);
        e->job_ns = n_slices; e->job_left = n_slices; e->job_err = CC_OK; e->job_gen++;
        __atomic_store_n(&e->job_next, 0, __ATOMIC_RELEASE); /* a worker still leaving the previous frame's loop may pick a slice up at once */
        pthread_cond_broadcast(&e->job_cv);
        pthread_mutex_unlock(&e->job_mu);
        pool_take_slices(e, e->n_workers);
        pthread_mutex_lock(&e->job_mu);
        while (__atomic_load_n(&e->job_left, __ATOMIC_ACQUIRE) > 0) pthread_cond_wait(&e->done_cv, &e->job_mu);
        pthread_mutex_unlock(&e->job_mu);
        if (e->job_err != CC_OK) { if (e->job_err == CC_ERR_STATE) fprintf(stderr, "ceracuda: malformed symbol stream\n"); return e->job_err; }
        for (int s = 0; s < n_slices; s++) {
            if (n + e->slice_len[s] > e->au_cap) return CC_ERR_OVERFLOW;
            memcpy(e->au + n, e->slice_out[s], e->slice_len[s]);
            n += e->slice_len[s];
        }
    } else {
        for (int s = 0; s < n_slices; s++) {
            size_t w = 0;
            const int rc = write_slice_nal(e, s, e->rbsp, e->au + n, e->au_cap - n, &w);
            if (rc != CC_OK) { if (rc == CC_ERR_STATE) fprintf(stderr, "ceracuda: malformed symbol stream\n"); return rc; }
            n += w;
        }
    }
    cc_rc_update(&e->rc, is_idr, e->cur_qp, (long)n * 8);
    e->frame_num = (e->frame_num + 1) & 15;
    if (is_idr) e->idr_pic_id = (e->idr_pic_id + 1) & 0xffff;
    e->gop_pos++;
    e->frame_idx++;
    out->data = e->au; out->size = (int)n; out->is_idr = is_idr; out->qp = e->cur_qp; out->pts = e->cur_pts;
    out->n_skip = ns; out->n_inter = np; out->n_intra = ni; out->n_intra_analysed = ng; out->n_skip_probed = npr;
    out->sym_bytes = (int)dev_bytes_out;
    out->gpu_ms = gpu_ms; out->host_ms = (float)(now_ms() - t0);
    return CC_OK;
}

int cc_enc_encode(cc_enc *e, const cc_frame *in, unsigned flags, cc_packet *out) {
    int rc = cc_enc_submit(e, in, flags);
    return rc != CC_OK ? rc : cc_enc_collect(e, out);
}

/* starts the upload of the frame that will be submitted next (planar I420; any other format is simply uploaded by its own
 * cc_enc_submit).  The set of source planes it goes to was last read by the frame before
 * the one in flight, which has been collected. */
static void enc_prefetch(cc_enc *e, const cc_frame *in, unsigned flags) {
    cc_kparams *k = &e->k;
    if (e->p.format != CC_FMT_I420 || !e->cur_alt[0] || !in->plane[0] || !in->plane[1] || !in->plane[2]) return;
    if (in->stride[0] < k->w || in->stride[1] < k->w / 2 || in->stride[2] < k->w / 2) return;
    const uint8_t *pl[3] = {in->plane[0], in->plane[1], in->plane[2]};
    int st[3] = {in->stride[0], in->stride[1], in->stride[2]};
    if (!(flags & CC_FRAME_DEVICE) && !host_is_pinned(in->plane[0])) {
        /* pageable memory (a mapped GstBuffer): the calling thread copies the frame into the second pinned staging frame now, while the
         * device codes the frame before it; the buffer's previous upload has long finished, but that is waited for, not assumed */
        if (!e->stage_host_alt || cudaEventSynchronize(e->ev_up) != cudaSuccess) { cudaGetLastError(); return; }
        uint8_t *d = e->stage_host_alt;
        for (int i = 0; i < 3; i++) {
            const size_t rowb = i == 0 ? (size_t)k->w : (size_t)k->w / 2, rows = i == 0 ? (size_t)k->h : (size_t)k->h / 2;
            if ((size_t)in->stride[i] == rowb) memcpy(d, in->plane[i], rowb * rows);
            else for (size_t r = 0; r < rows; r++) memcpy(d + r * rowb, in->plane[i] + r * (size_t)in->stride[i], rowb);
            pl[i] = d; st[i] = (int)rowb;
            d += rowb * rows;
        }
        e->staged_frames++;
    }
    const enum cudaMemcpyKind kind = (flags & CC_FRAME_DEVICE) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    if (cudaMemcpy2DAsync(e->cur_alt[0], (size_t)k->sy, pl[0], (size_t)st[0], (size_t)k->w, (size_t)k->h, kind, e->st_up) != cudaSuccess ||
        cudaMemcpy2DAsync(e->cur_alt[1], (size_t)k->sc, pl[1], (size_t)st[1], (size_t)k->w / 2, (size_t)k->h / 2, kind, e->st_up) != cudaSuccess ||
        cudaMemcpy2DAsync(e->cur_alt[2], (size_t)k->sc, pl[2], (size_t)st[2], (size_t)k->w / 2, (size_t)k->h / 2, kind, e->st_up) != cudaSuccess ||
        cudaEventRecord(e->ev_up, e->st_up) != cudaSuccess) { cudaGetLastError(); return; }
    e->pf_plane0 = in->plane[0];
}
long cc_enc_prefetched_frames(const cc_enc *e) { return e ? e->prefetched_frames : 0; }
long cc_enc_overlapped_frames(const cc_enc *e) { return e ? e->overlapped_frames : 0; }

int cc_enc_encode_sequence(cc_enc *e, const cc_frame *frames, const int *bitrate_bps, int n, unsigned flags, cc_seq_stats *acc, cc_packet *last) {
    if (!e || !frames || !acc || n < 0) return CC_ERR_PARAM;
    cc_packet pkt;
    memset(&pkt, 0, sizeof pkt);
    for (int i = 0; i < n; i++) {
        if (bitrate_bps && bitrate_bps[i] != cc_enc_get_bitrate(e)) cc_enc_set_bitrate(e, bitrate_bps[i]);
        int rc = cc_enc_submit(e, &frames[i], flags);
        if (rc == CC_OK && i + 1 < n) enc_prefetch(e, &frames[i + 1], flags); /* the next frame's upload runs next to this frame's kernels */
        if (rc == CC_OK) rc = cc_enc_collect(e, &pkt);
        if (rc != CC_OK) return rc;
        acc->frames++; acc->au_bytes += pkt.size; acc->sym_bytes += pkt.sym_bytes;
        acc->n_skip += pkt.n_skip; acc->n_inter += pkt.n_inter; acc->n_intra += pkt.n_intra;
        acc->n_intra_analysed += pkt.n_intra_analysed; acc->n_skip_probed += pkt.n_skip_probed;
        acc->submit_ms += pkt.submit_ms; acc->wait_ms += pkt.wait_ms; acc->host_ms += pkt.host_ms; acc->gpu_ms += pkt.gpu_ms;
        acc->last_qp = pkt.qp; acc->n_idr += pkt.is_idr;
    }
    if (last) *last = pkt;
    return CC_OK;
}

/* ---- test hooks ----------------------------------------------------------------------------------- */
int cc_enc_get_recon(cc_enc *e, uint8_t *y, uint8_t *u, uint8_t *v) {
    if (!e || !y || !u || !v || e->inflight) return CC_ERR_PARAM;
    if (bind_device(e) != CC_OK) return CC_ERR_CUDA;
    cc_kparams *k = &e->k; /* after submit the newest reconstruction is k->ref */
    CUCHECK(cudaStreamSynchronize(e->st));
    CUCHECK(cudaStreamSynchronize(e->st2));
    CUCHECK(cudaMemcpy2D(y, (size_t)k->W, k->ref[0], (size_t)k->sy, (size_t)k->W, (size_t)k->H, cudaMemcpyDeviceToHost));
    CUCHECK(cudaMemcpy2D(u, (size_t)k->W / 2, k->ref[1], (size_t)k->sc, (size_t)k->W / 2, (size_t)k->H / 2, cudaMemcpyDeviceToHost));
    CUCHECK(cudaMemcpy2D(v, (size_t)k->W / 2, k->ref[2], (size_t)k->sc, (size_t)k->W / 2, (size_t)k->H / 2, cudaMemcpyDeviceToHost));
    return CC_OK;
}
int cc_enc_get_me(cc_enc *e, int16_t *mv4, int32_t *satd) {
    if (!e || !mv4 || !satd || e->inflight) return CC_ERR_PARAM;
    if (bind_device(e) != CC_OK) return CC_ERR_CUDA;
    const int nmb = e->k.mb_w * e->k.mb_h;
    cc_me *tmp = (cc_me *)malloc(sizeof(cc_me) * nmb);
    if (!tmp) return CC_ERR_NOMEM;
    cudaError_t er = cudaMemcpy(tmp, e->k.me, sizeof(cc_me) * nmb, cudaMemcpyDeviceToHost);
    for (int i = 0; i < nmb && er == cudaSuccess; i++) {
        mv4[4 * i] = tmp[i].l1_mv[0]; mv4[4 * i + 1] = tmp[i].l1_mv[1]; mv4[4 * i + 2] = tmp[i].mv[0]; mv4[4 * i + 3] = tmp[i].mv[1];
        satd[i] = tmp[i].satd;
    }
    free(tmp);
    return er == cudaSuccess ? CC_OK : CC_ERR_CUDA;
}
int cc_enc_get_me8(cc_enc *e, int16_t *mv8, int32_t *satd4, uint8_t *part) {
    if (!e || !mv8 || !satd4 || !part || e->inflight) return CC_ERR_PARAM;
    if (bind_device(e) != CC_OK) return CC_ERR_CUDA;
    const int nmb = e->k.mb_w * e->k.mb_h;
    cc_me *t16 = (cc_me *)malloc(sizeof(cc_me) * nmb);
    cc_me8 *t8 = (cc_me8 *)malloc(sizeof(cc_me8) * nmb);
    cudaError_t er = (t16 && t8) ? cudaMemcpy(t16, e->k.me, sizeof(cc_me) * nmb, cudaMemcpyDeviceToHost) : cudaErrorMemoryAllocation;
    if (er == cudaSuccess) er = cudaMemcpy(t8, e->k.me8, sizeof(cc_me8) * nmb, cudaMemcpyDeviceToHost);
    for (int i = 0; i < nmb && er == cudaSuccess; i++) {
        memcpy(mv8 + 8 * i, t8[i].mv, 16); memcpy(satd4 + 4 * i, t8[i].satd, 16);
        part[i] = (uint8_t)t16[i].part;
    }
    free(t16); free(t8);
    return er == cudaSuccess ? CC_OK : CC_ERR_CUDA;
}
int cc_enc_get_mbinfo(cc_enc *e, uint8_t *dst8) {
    if (!e || !dst8 || e->inflight) return CC_ERR_PARAM;
    if (bind_device(e) != CC_OK) return CC_ERR_CUDA;
    CUCHECK(cudaMemcpy(dst8, e->k.mbh, sizeof(cc_mbhost) * e->k.mb_w * e->k.mb_h, cudaMemcpyDeviceToHost));
    return CC_OK;
}

void *cc_host_alloc(size_t bytes) { void *p = NULL; return cudaHostAlloc(&p, bytes, cudaHostAllocDefault) == cudaSuccess ? p : NULL; }
void cc_host_free(void *p) { if (p) cudaFreeHost(p); }
int cc_host_register(void *p, size_t bytes) { return (p && cudaHostRegister(p, bytes, cudaHostRegisterDefault) == cudaSuccess) ? CC_OK : CC_ERR_CUDA; }
int cc_host_unregister(void *p) { return (p && cudaHostUnregister(p) == cudaSuccess) ? CC_OK : CC_ERR_CUDA; }
void *cc_dev_alloc(size_t bytes) { void *p = NULL; return cudaMalloc(&p, bytes) == cudaSuccess ? p : NULL; }
void cc_dev_free(void *p) { if (p) cudaFree(p); }
int cc_dev_upload(void *dst, const void *src, size_t bytes) { return cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice) == cudaSuccess ? CC_OK : CC_ERR_CUDA; }
