/*
 * cc_rc.c — one-pass, zero-lookahead rate control.  Host C, scalar.
 * The target bitrate is whatever ceracoder's balancer last wrote through the element's bitrate property
 * (/root/reference/src/gst/encoder_control.c:53; 20 ms tick, /root/reference/src/core/bitrate_control.h:35);
 * it is sampled at every frame boundary.  Model: bits ~ coef[type] / qscale(qp), qscale = 0.85*2^((qp-12)/6)
 * [EXT: x264 qp2qscale]; a virtual buffer accumulates (actual - target) bits and is paid back over ~0.5 s.
 */
#include "cc_host.h"
#include <math.h>

static double qp2qscale(double qp) { return 0.85 * pow(2.0, (qp - 12.0) / 6.0); }
static double qscale2qp(double qs) { return 12.0 + 6.0 * log2(qs / 0.85); }

void cc_rc_init(cc_rc *rc, double fps, int bitrate, int qp_min, int qp_max) {
    rc->fps = fps > 1 ? fps : 1;
    rc->qp_min = qp_min; rc->qp_max = qp_max; rc->bitrate = bitrate;
    rc->coef[0] = rc->coef[1] = 0; rc->have[0] = rc->have[1] = 0;
    rc->fullness = 0; rc->last_qp_p = -1; rc->frames = 0;
}

static int clampqp(const cc_rc *rc, double q) {
    int qp = (int)floor(q + 0.5);
    return qp < rc->qp_min ? rc->qp_min : (qp > rc->qp_max ? rc->qp_max : qp);
}

int cc_rc_frame_qp(cc_rc *rc, int bitrate_now, int is_idr, int nmb) {
    if (bitrate_now != rc->bitrate) {
        /* a step from the balancer: keep the complexity model, rescale the debt to the new budget */
        rc->fullness *= (double)bitrate_now / (double)rc->bitrate;
        rc->bitrate = bitrate_now;
    }
    const double T = rc->bitrate / rc->fps;
    double desired = T - rc->fullness / (0.5 * rc->fps);
    if (desired < 0.15 * T) desired = 0.15 * T;
    if (desired > 4.0 * T) desired = 4.0 * T;
    const double bpp = T / ((double)nmb * 256.0);
    const double qp0 = 26.0 + 6.0 * log2(0.10 / bpp) * 0.8;
    double q;
    if (!is_idr) {
        if (rc->have[0] > 0) q = qscale2qp(rc->coef[0] / desired);
        else q = rc->last_qp_p >= 0 ? rc->last_qp_p : qp0 + 2;
        if (rc->last_qp_p >= 0) { /* limit the swing per frame; larger when the buffer is far off */
            double lim = fabs(rc->fullness) > 0.5 * rc->bitrate * 0.5 ? 6 : 3;
            if (q > rc->last_qp_p + lim) q = rc->last_qp_p + lim;
            if (q < rc->last_qp_p - lim) q = rc->last_qp_p - lim;
        }
    } else {
        q = rc->last_qp_p >= 0 ? rc->last_qp_p - 3 : qp0 - 1;
        if (rc->have[1] > 0) { /* keep a key frame below a quarter second of bits */
            double cap = 0.25 * rc->bitrate, pred = rc->coef[1] / qp2qscale(q);
            if (pred > cap) q = qscale2qp(rc->coef[1] / cap);
        }
    }
    return clampqp(rc, q);
}

void cc_rc_update(cc_rc *rc, int is_idr, int qp, long bits) {
    const double T = rc->bitrate / rc->fps, v = (double)bits * qp2qscale(qp);
    int t = is_idr ? 1 : 0;
    rc->coef[t] = rc->have[t] > 0 ? 0.5 * rc->coef[t] + 0.5 * v : v;
    rc->have[t] = 1;
    rc->fullness += (double)bits - T;
    if (rc->fullness > rc->bitrate) rc->fullness = rc->bitrate;
    if (rc->fullness < -0.5 * rc->bitrate) rc->fullness = -0.5 * rc->bitrate;
    if (!is_idr) rc->last_qp_p = qp;
    else if (rc->last_qp_p < 0) rc->last_qp_p = qp + 3;
    rc->frames++;
}
