/*
 * oracle/ref_convert.c — TEST INFRASTRUCTURE ONLY (never linked or loaded by the product).
 *
 * CPU statement of the colour conversion the ingest kernel fuses for packed input: what the `videoconvert`
 * element in front of x264enc does in the reference's pipeline text
 * (/root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:3-5, `… ! videoconvert ! … ! x264enc`).
 *
 * PARITY UNPINNED: videoconvert is GStreamer code (libgstvideo, un-vendored, absent from this image) and the reference holds no
 * golden frame for it, so the arithmetic below is this project's definition, not a restatement of GStreamer's:
 *   YUY2 -> I420   luma copied; chroma of a 2x2 block = rounded mean of the block's two rows ((a + b + 1) >> 1)
 *   BGRx -> I420   8-bit limited-range matrix, BT.709 when the picture has more than 576 lines, else BT.601 (GStreamer's
 *                  default colorimetry for YUV video of that size); luma per pixel, chroma from the rounded 2x2 mean of R, G, B
 *                    Y  = (cyr R + cyg G + cyb B +  4224) >> 8       (4224  = 128 + 16 << 8)
 *                    Cb = (cbr R + cbg G + cbb B + 32896) >> 8       (32896 = 128 + 128 << 8)
 *                    Cr = (crr R + crg G + crb B + 32896) >> 8
 * The CUDA kernel (ceracoder_b200/csrc/cc_k_frame.cu k_ingest_packed) must produce the same planes bit for bit
 * (tests/test_convert.py).
 */
#include <stdint.h>

static const int kMat[2][9] = {
    /* BT.601 */ {66, 129, 25, -38, -74, 112, 112, -94, -18},
    /* BT.709 */ {47, 157, 16, -26, -87, 112, 112, -102, -10},
};

/* w, h even.  y: w x h (stride sy); u, v: w/2 x h/2 (stride sc) */
void ref_convert_yuy2(const uint8_t *src, int stride, int w, int h, uint8_t *y, int sy, uint8_t *u, uint8_t *v, int sc) {
    for (int r = 0; r < h; r += 2)
        for (int p = 0; p < w / 2; p++) {
            const uint8_t *a = src + (long)r * stride + 4 * p, *b = a + stride;
            y[(long)r * sy + 2 * p] = a[0]; y[(long)r * sy + 2 * p + 1] = a[2];
            y[(long)(r + 1) * sy + 2 * p] = b[0]; y[(long)(r + 1) * sy + 2 * p + 1] = b[2];
            u[(long)(r / 2) * sc + p] = (uint8_t)((a[1] + b[1] + 1) >> 1);
            v[(long)(r / 2) * sc + p] = (uint8_t)((a[3] + b[3] + 1) >> 1);
        }
}

void ref_convert_bgrx(const uint8_t *src, int stride, int w, int h, uint8_t *y, int sy, uint8_t *u, uint8_t *v, int sc) {
    const int *m = kMat[h > 576];
    for (int r = 0; r < h; r += 2)
        for (int p = 0; p < w / 2; p++) {
            int sr = 0, sg = 0, sb = 0;
            for (int dy = 0; dy < 2; dy++)
                for (int dx = 0; dx < 2; dx++) {
                    const uint8_t *q = src + (long)(r + dy) * stride + 4 * (2 * p + dx);
                    const int B = q[0], G = q[1], R = q[2];
                    y[(long)(r + dy) * sy + 2 * p + dx] = (uint8_t)((m[0] * R + m[1] * G + m[2] * B + 4224) >> 8);
                    sr += R; sg += G; sb += B;
                }
            const int R = (sr + 2) >> 2, G = (sg + 2) >> 2, B = (sb + 2) >> 2;
            u[(long)(r / 2) * sc + p] = (uint8_t)((m[3] * R + m[4] * G + m[5] * B + 32896) >> 8);
            v[(long)(r / 2) * sc + p] = (uint8_t)((m[6] * R + m[7] * G + m[8] * B + 32896) >> 8);
        }
}
