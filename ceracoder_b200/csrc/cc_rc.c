/*
 * cc_rc.c — one-pass, zero-lookahead rate control.  Host C, scalar.
 * The target bitrate is whatever ceracoder's balancer last wrote through the element's bitrate property
 * (/root/reference/src/gst/encoder_control.c:53; 20 ms tick, /root/reference/src/core/bitrate_control.h:35);
 * it is sampled at every frame boundary.  Model: bits ~ coef[type] / qscale(qp), qscale = 0.85*2^((qp-12)/6)
 * [EXT: x264 qp2qscale].  Three pieces sit on that model:
 *   - a debt account: (actual - target) bits accumulate and the following frames pay them back over a third of a second;
 *   - a VBV leaky bucket [EXT: x264enc vbv-buf-capacity, default 600 ms of bitrate] that drains at the target rate: no frame
 *     is planned larger than the room left in it — the consumer is an SRT link with a latency budget
 *     (/root/reference/src/core/bitrate_control.h:30-44), so peak-to-mean matters as much as the mean;
 *   - key-frame sizing: what the metric of BASELINE.json looks at is the error of 1 s windows.  When the key-frame interval
 *     is no longer than that window every window holds one key frame and its size cancels; when it is longer (GOP 60 at
 *     30 fps), a key frame of k frame budgets leaves an error of about (k - 1) / GOP in the windows around it whatever the
 *     P frames do, so k is held to 1 + 5 % of the GOP — the key frame is coded coarser and the P frames behind it refine it;
 *   - a balancer step (the schedule drops from 12 to 2 Mbit/s in one tick) moves the QP by the model at once instead of
 *     walking it there a few steps per frame: three frames at the old size are a quarter of a second of the new budget.
 */
#include "cc_host.h"
#include <math.h>
#include <stdlib.h>

static double qp2qscale(double qp) { return 0.85 * pow(2.0, (qp - 12.0) / 6.0); }
static double qscale2qp(double qs) { return 12.0 + 6.0 * log2(qs / 0.85); }

void cc_rc_init(cc_rc *rc, double fps, int bitrate, int qp_min, int qp_max, int key_int) {
    rc->fps = fps > 1 ? fps : 1;
    rc->qp_min = qp_min; rc->qp_max = qp_max; rc->bitrate = bitrate;
    rc->coef[0] = rc->coef[1] = 0; rc->have[0] = rc->have[1] = 0;
    rc->fullness = 0; rc->last_qp_p = -1; rc->frames = 0;
    rc->key_int = key_int > 0 ? key_int : 60;
    rc->vbv_fill = 0; rc->vbv_seconds = 0.6; rc->step = 0; rc->vbv_peak = 0;
    rc->ema = 0.12; rc->max_step = 3; rc->fast = 0; /* measured on the B200, profiles/rc_grid_r02.txt */
    { /* experiment knobs (tools/rc_report.py sweeps them); the defaults above are what ships */
        const char *v;
        if ((v = getenv("CERACUDA_RC_EMA"))) rc->ema = atof(v);
        if ((v = getenv("CERACUDA_RC_STEP"))) rc->max_step = atof(v);
    }
}

static int clampqp(const cc_rc *rc, double q) {
    int qp = (int)floor(q + 0.5);
    return qp < rc->qp_min ? rc->qp_min : (qp > rc->qp_max ? rc->qp_max : qp);
}

int cc_rc_frame_qp(cc_rc *rc, int bitrate_now, int is_idr, int nmb) {
    if (bitrate_now != rc->bitrate) {
        /* a step from the balancer: keep the complexity model, rescale the debt and the bucket to the new budget */
        const double r = (double)bitrate_now / (double)rc->bitrate;
        rc->fullness *= r; rc->vbv_fill *= r;
        if (r < 0.87 || r > 1.15) rc->step = 1; /* more than the ladder's own +-100..300 kbit/s moves */
        rc->bitrate = bitrate_now;
    }
    const double T = rc->bitrate / rc->fps;
    const double room = rc->vbv_seconds * rc->bitrate - rc->vbv_fill + T; /* bits that fit the bucket now (it drains T per frame) */
    double desired = T - rc->fullness / (rc->fps / 3.0);
    if (desired > room) desired = room;
    if (desired < 0.15 * T) desired = 0.15 * T;
    if (desired > 4.0 * T) desired = 4.0 * T;
    const double bpp = T / ((double)nmb * 256.0);
    const double qp0 = 26.0 + 6.0 * log2(0.10 / bpp) * 0.8;
    double q;
    if (!is_idr) {
        if (rc->have[0] > 0) q = qscale2qp(rc->coef[0] / desired);
        else q = rc->last_qp_p >= 0 ? rc->last_qp_p : qp0 + 2;
        if (rc->last_qp_p >= 0 && !rc->step) { /* limit the swing per frame; larger when the debt is far off */
            double lim = fabs(rc->fullness) > 0.25 * rc->bitrate ? 2 * rc->max_step : rc->max_step;
            if (q > rc->last_qp_p + lim) q = rc->last_qp_p + lim;
            if (q < rc->last_qp_p - lim) q = rc->last_qp_p - lim;
        }
        if (rc->step) rc->fast = 4; /* the model is about to be used far from where it was learnt: let the next frames correct it quickly */
        rc->step = 0;
    } else {
        const double win = floor(rc->fps + 0.5);
        double k = rc->key_int <= win ? 0.25 * rc->fps : 1.0 + 0.05 * rc->key_int;
        if (k > 0.25 * rc->fps) k = 0.25 * rc->fps;
        if (k < 2.0) k = 2.0;
        double cap = k * T;
        if (cap > room) cap = room;
        if (cap < T) cap = T;
        q = rc->last_qp_p >= 0 ? rc->last_qp_p - 3 : qp0 - 1;
        if (rc->have[1] > 0) {
            double pred = rc->coef[1] / qp2qscale(q);
            if (pred > cap) q = qscale2qp(rc->coef[1] / cap);
        }
    }
    return clampqp(rc, q);
}

void cc_rc_update(cc_rc *rc, int is_idr, int qp, long bits) {
    const double T = rc->bitrate / rc->fps, v = (double)bits * qp2qscale(qp);
    int t = is_idr ? 1 : 0;
    /* P frames: a slow average.  Content with grain has a cliff in bits over QP, and a frame coded finer than its reference pays
     * for the reference's coarseness as well: following the last frame's size makes QP and size alternate (2.4 T / 0.27 T at
     * QP 21 / 24 was measured at 2160p30).  After a balancer step the model adapts fast for a few frames. */
    const double a = is_idr ? 0.5 : (rc->fast > 0 ? 0.5 : rc->ema);
    if (!is_idr && rc->fast > 0) rc->fast--;
    rc->coef[t] = rc->have[t] > 0 ? (1.0 - a) * rc->coef[t] + a * v : v;
    rc->have[t] = 1;
    rc->fullness += (double)bits - T;
    if (rc->fullness > rc->bitrate) rc->fullness = rc->bitrate;
    if (rc->fullness < -0.5 * rc->bitrate) rc->fullness = -0.5 * rc->bitrate;
    rc->vbv_fill += (double)bits - T;
    if (rc->vbv_fill < 0) rc->vbv_fill = 0;
    if (rc->vbv_fill > rc->vbv_peak * rc->bitrate) rc->vbv_peak = rc->vbv_fill / rc->bitrate;
    if (rc->vbv_fill > 2.0 * rc->vbv_seconds * rc->bitrate) rc->vbv_fill = 2.0 * rc->vbv_seconds * rc->bitrate;
    if (!is_idr) rc->last_qp_p = qp;
    else if (rc->last_qp_p < 0) rc->last_qp_p = qp + 3;
    rc->frames++;
}
