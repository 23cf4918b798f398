/*
 * gstcudah264enc.c — GStreamer element `cudah264enc`: the outer drop-in boundary (SURVEY.md §8b, B-outer).
 *
 * It stands where ceracoder's pipeline files name `x264enc`
 * (/root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6: "x264enc speed-preset=3 key-int-max=60
 * name=venc_kbps"), is instantiated by gst_parse_launch (/root/reference/src/io/pipeline_loader.c:59), found by
 * NAME (/root/reference/src/gst/encoder_control.c:29-32) and driven by g_object_set(element, "bps", ...)
 * (/root/reference/src/gst/encoder_control.c:53) from the 20 ms balancer tick (/root/reference/src/ceracoder.c:263).
 * Downstream is `h264parse config-interval=-1 ! queue ! mpegtsmux` (same file, lines 6-10), so byte-stream access
 * units with SPS/PPS in-band at every IDR are what it must produce.
 *
 * All codec work is in libceracuda.so (include/ceracuda.h); this file is glue: caps, properties, buffers.
 * Built where GStreamer development files exist (`make gst`, gated on pkg-config gstreamer-video-1.0).  Neither this
 * container nor the GPU box has them, so here the file is held to the GStreamer API by compiling it against the
 * declaration-only header set under tests/shim/gst (tests/test_element_source.py: gcc -fsyntax-only -Wall -Werror), and its
 * property table is checked against every pipeline/generic/cuda_h264_* line by the same test; the property semantics are
 * exercised through the ctypes mirror (ceracoder_b200/encoder.py).
 *
 * Threads: set_property runs on the GLib main-loop thread (the balancer tick), set_format / handle_frame on
 * the streaming thread, stop / finalize on the state-change thread.  `self->enc` is therefore only read or swapped under
 * GST_OBJECT_LOCK; the streaming thread, the only one that swaps it while data flows, may read its own pointer unlocked.
 */
#include <gst/gst.h>
#include <gst/video/video.h>
#include <gst/video/gstvideoencoder.h>
#include <string.h>
#include "../../include/ceracuda.h"

#define GST_TYPE_CUDA_H264_ENC (gst_cuda_h264_enc_get_type())
G_DECLARE_FINAL_TYPE(GstCudaH264Enc, gst_cuda_h264_enc, GST, CUDA_H264_ENC, GstVideoEncoder)

struct _GstCudaH264Enc {
    GstVideoEncoder parent;
    cc_enc *enc;                 /* NULL until set_format; guarded by GST_OBJECT_LOCK (see "Threads" above) */
    GstVideoCodecState *input_state;
    /* properties; `bitrate_bps` is written from the GLib main-loop thread while handle_frame runs on a streaming thread */
    volatile gint bitrate_bps;
    guint key_int_max, threads, rows_per_slice;
    gint speed_preset;           /* GstCudaH264EncPreset */
    guint tune;                  /* GstCudaH264EncTune */
    gint device;
    gboolean byte_stream;
};

G_DEFINE_TYPE(GstCudaH264Enc, gst_cuda_h264_enc, GST_TYPE_VIDEO_ENCODER)
GST_DEBUG_CATEGORY_STATIC(cudah264enc_debug);
#define GST_CAT_DEFAULT cudah264enc_debug

enum { PROP_0, PROP_BITRATE, PROP_BPS, PROP_SPEED_PRESET, PROP_KEY_INT_MAX, PROP_TUNE, PROP_BYTE_STREAM, PROP_THREADS, PROP_DEVICE, PROP_ROWS_PER_SLICE };

/* x264enc's `speed-preset` enum and `tune` flags, value for value and nick for nick [EXT: GstX264EncPreset, GstX264EncTune], so
 * that both spellings in use parse: `speed-preset=3` (the reference's pipeline files) and `speed-preset=veryfast
 * tune=zerolatency` (BASELINE.json).  This encoder has one speed and is always zero-latency; the values are kept and
 * reported back, nothing more. */
#define GST_TYPE_CUDA_H264_ENC_PRESET (gst_cuda_h264_enc_preset_get_type())
static GType gst_cuda_h264_enc_preset_get_type(void) {
    static gsize type_id = 0;
    static const GEnumValue values[] = {
        {0, "No preset", "None"},      {1, "ultrafast", "ultrafast"}, {2, "superfast", "superfast"}, {3, "veryfast", "veryfast"},
        {4, "faster", "faster"},       {5, "fast", "fast"},           {6, "medium", "medium"},       {7, "slow", "slow"},
        {8, "slower", "slower"},       {9, "veryslow", "veryslow"},   {10, "placebo", "placebo"},    {0, NULL, NULL}};
    if (g_once_init_enter(&type_id)) g_once_init_leave(&type_id, (gsize)g_enum_register_static("GstCudaH264EncPreset", values));
    return (GType)type_id;
}
#define GST_TYPE_CUDA_H264_ENC_TUNE (gst_cuda_h264_enc_tune_get_type())
static GType gst_cuda_h264_enc_tune_get_type(void) {
    static gsize type_id = 0;
    static const GFlagsValue values[] = {
        {0x1, "Still image", "stillimage"}, {0x2, "Fast decode", "fastdecode"}, {0x4, "Zero latency", "zerolatency"}, {0, NULL, NULL}};
    if (g_once_init_enter(&type_id)) g_once_init_leave(&type_id, (gsize)g_flags_register_static("GstCudaH264EncTune", values));
    return (GType)type_id;
}

/* The stream is High profile (profile_idc 100: the 8x8 transform is on; entropy coding is CAVLC) — cc_bitstream.c cc_write_sps */
static GstStaticPadTemplate sink_tmpl = GST_STATIC_PAD_TEMPLATE("sink", GST_PAD_SINK, GST_PAD_ALWAYS,
    GST_STATIC_CAPS("video/x-raw, format=(string){ I420, NV12, YUY2, BGRx }, width=(int)[16,8192], height=(int)[16,8192], framerate=(fraction)[0/1,MAX]"));
static GstStaticPadTemplate src_tmpl = GST_STATIC_PAD_TEMPLATE("src", GST_PAD_SRC, GST_PAD_ALWAYS,
    GST_STATIC_CAPS("video/x-h264, stream-format=(string)byte-stream, alignment=(string)au, profile=(string)high"));

/* The reference writes `bitrate_bps / div` to "bps": kbit/s when the element is named venc_kbps, bit/s when named
 * venc_bps (/root/reference/src/gst/encoder_control.c:29-32,53).  Disambiguate by our own object name, with the
 * magnitude fallback of SURVEY §8b (legal range 300 k..30 M bps, /root/reference/src/core/bitrate_control.h:30-31). */
static gint bps_from_property(GstCudaH264Enc *self, gint v) {
    gchar *name = gst_object_get_name(GST_OBJECT(self));
    gboolean kbps = name && g_str_has_prefix(name, "venc_kbps");
    gboolean bps = name && g_str_has_prefix(name, "venc_bps");
    g_free(name);
    if (kbps || (!bps && v <= 30000)) return v * 1000;
    return v;
}

/* main-loop thread, every 20 ms at most: must never block on the encoder.  The lock is held for one atomic store. */
static void set_bitrate(GstCudaH264Enc *self, gint bps) {
    g_atomic_int_set(&self->bitrate_bps, bps);
    GST_OBJECT_LOCK(self);
    if (self->enc) cc_enc_set_bitrate(self->enc, bps); /* wait-free atomic latch; applied at the next frame boundary */
    GST_OBJECT_UNLOCK(self);
}

static void gst_cuda_h264_enc_set_property(GObject *obj, guint id, const GValue *val, GParamSpec *ps) {
    GstCudaH264Enc *self = GST_CUDA_H264_ENC(obj);
    switch (id) {
    case PROP_BITRATE: set_bitrate(self, (gint)g_value_get_uint(val) * 1000); break; /* x264enc-compatible: kbit/s */
    case PROP_BPS: set_bitrate(self, bps_from_property(self, g_value_get_int(val))); break;
    case PROP_SPEED_PRESET: self->speed_preset = g_value_get_enum(val); break; /* accepted for pipeline-text compatibility */
    case PROP_KEY_INT_MAX: self->key_int_max = g_value_get_uint(val); break;
    case PROP_TUNE: self->tune = g_value_get_flags(val); break;
    case PROP_BYTE_STREAM: self->byte_stream = g_value_get_boolean(val); break;
    case PROP_THREADS: self->threads = g_value_get_uint(val); break;
    case PROP_DEVICE: self->device = g_value_get_int(val); break;
    case PROP_ROWS_PER_SLICE: self->rows_per_slice = g_value_get_uint(val); break;
    default: G_OBJECT_WARN_INVALID_PROPERTY_ID(obj, id, ps);
    }
}
static void gst_cuda_h264_enc_get_property(GObject *obj, guint id, GValue *val, GParamSpec *ps) {
    GstCudaH264Enc *self = GST_CUDA_H264_ENC(obj);
    gint bps = g_atomic_int_get(&self->bitrate_bps);
    switch (id) {
    case PROP_BITRATE: g_value_set_uint(val, (guint)(bps / 1000)); break;
    case PROP_BPS: {
        gchar *name = gst_object_get_name(GST_OBJECT(self));
        g_value_set_int(val, (name && g_str_has_prefix(name, "venc_bps")) ? bps : bps / 1000);
        g_free(name);
        break;
    }
    case PROP_SPEED_PRESET: g_value_set_enum(val, self->speed_preset); break;
    case PROP_KEY_INT_MAX: g_value_set_uint(val, self->key_int_max); break;
    case PROP_TUNE: g_value_set_flags(val, self->tune); break;
    case PROP_BYTE_STREAM: g_value_set_boolean(val, self->byte_stream); break;
    case PROP_THREADS: g_value_set_uint(val, self->threads); break;
    case PROP_DEVICE: g_value_set_int(val, self->device); break;
    case PROP_ROWS_PER_SLICE: g_value_set_uint(val, self->rows_per_slice); break;
    default: G_OBJECT_WARN_INVALID_PROPERTY_ID(obj, id, ps);
    }
}

/* detach the handle under the lock, close it outside: cc_enc_close waits for the device, and set_bitrate on the main-loop
 * thread must never wait behind that */
static void close_codec(GstCudaH264Enc *self) {
    cc_enc *old;
    GST_OBJECT_LOCK(self);
    old = self->enc;
    self->enc = NULL;
    GST_OBJECT_UNLOCK(self);
    if (old) cc_enc_close(old);
    if (self->input_state) { gst_video_codec_state_unref(self->input_state); self->input_state = NULL; }
}
static gboolean gst_cuda_h264_enc_stop(GstVideoEncoder *venc) { close_codec(GST_CUDA_H264_ENC(venc)); return TRUE; }

/* (re)open the codec for the negotiated format — x264enc does the same in its set_format [EXT] */
static gboolean gst_cuda_h264_enc_set_format(GstVideoEncoder *venc, GstVideoCodecState *state) {
    GstCudaH264Enc *self = GST_CUDA_H264_ENC(venc);
    GstVideoInfo *info = &state->info;
    cc_enc_params p;
    cc_enc *fresh = NULL;
    close_codec(self);
    cc_enc_params_default(&p);
    p.device = self->device;
    p.width = GST_VIDEO_INFO_WIDTH(info); p.height = GST_VIDEO_INFO_HEIGHT(info);
    p.fps_n = GST_VIDEO_INFO_FPS_N(info) > 0 ? GST_VIDEO_INFO_FPS_N(info) : 30;
    p.fps_d = GST_VIDEO_INFO_FPS_D(info) > 0 ? GST_VIDEO_INFO_FPS_D(info) : 1;
    switch (GST_VIDEO_INFO_FORMAT(info)) { /* packed formats are converted inside the ingest kernel: no videoconvert upstream */
    case GST_VIDEO_FORMAT_NV12: p.format = CC_FMT_NV12; break;
    case GST_VIDEO_FORMAT_YUY2: p.format = CC_FMT_YUY2; break;
    case GST_VIDEO_FORMAT_BGRx: p.format = CC_FMT_BGRX; break;
    default: p.format = CC_FMT_I420; break;
    }
    p.bitrate_bps = g_atomic_int_get(&self->bitrate_bps);
    p.key_int_max = (int)self->key_int_max;
    p.rows_per_slice = (int)self->rows_per_slice;
    p.threads = (int)self->threads;
    int rc = cc_enc_open(&fresh, &p);
    if (rc != CC_OK) {
        GST_ELEMENT_ERROR(self, LIBRARY, INIT, ("cc_enc_open failed: %s", cc_strerror(rc)), ("there is no CPU fallback"));
        return FALSE;
    }
    GST_OBJECT_LOCK(self);
    self->enc = fresh;
    cc_enc_set_bitrate(fresh, g_atomic_int_get(&self->bitrate_bps)); /* a tick that fell between open and this swap is not lost */
    GST_OBJECT_UNLOCK(self);
    self->input_state = gst_video_codec_state_ref(state);
    GstCaps *caps = gst_caps_new_simple("video/x-h264", "stream-format", G_TYPE_STRING, "byte-stream", "alignment", G_TYPE_STRING, "au",
                                        "profile", G_TYPE_STRING, "high", NULL);
    GstVideoCodecState *out = gst_video_encoder_set_output_state(venc, caps, state);
    gst_video_codec_state_unref(out);
    gst_video_encoder_set_latency(venc, 0, 0); /* one frame in, one access unit out, synchronously (zerolatency) */
    return gst_video_encoder_negotiate(venc);
}

/* Every exit hands the frame back to the base class: gst_video_encoder_finish_frame with no output buffer drops it from the
 * base class's pending list (a bare gst_video_codec_frame_unref would leave it there until stop()). */
static GstFlowReturn gst_cuda_h264_enc_handle_frame(GstVideoEncoder *venc, GstVideoCodecFrame *frame) {
    GstCudaH264Enc *self = GST_CUDA_H264_ENC(venc);
    GstVideoFrame vf;
    cc_frame in;
    cc_packet pkt;
    cc_enc *enc = self->enc; /* streaming thread: the only thread that swaps the pointer while data flows */
    if (!enc || !self->input_state) { /* data before caps: nothing to encode with, and nothing was produced */
        GST_ELEMENT_ERROR(self, CORE, NEGOTIATION, ("frame received before the input format was set"), (NULL));
        (void)gst_video_encoder_finish_frame(venc, frame);
        return GST_FLOW_NOT_NEGOTIATED;
    }
    if (!gst_video_frame_map(&vf, &self->input_state->info, frame->input_buffer, GST_MAP_READ)) {
        GST_ELEMENT_ERROR(self, RESOURCE, READ, ("cannot map input buffer"), (NULL));
        (void)gst_video_encoder_finish_frame(venc, frame);
        return GST_FLOW_ERROR;
    }
    memset(&in, 0, sizeof in);
    for (guint i = 0; i < GST_VIDEO_FRAME_N_PLANES(&vf) && i < 3; i++) { /* GStreamer strides are not tight */
        in.plane[i] = GST_VIDEO_FRAME_PLANE_DATA(&vf, i);
        in.stride[i] = GST_VIDEO_FRAME_PLANE_STRIDE(&vf, i);
    }
    in.pts = (int64_t)frame->pts;
    unsigned flags = GST_VIDEO_CODEC_FRAME_IS_FORCE_KEYFRAME(frame) ? CC_FORCE_IDR : 0; /* upstream force-key-unit */
    /* the mapped GstBuffer is pageable memory: cc_enc_submit stages it through the handle's pinned ring (ceracuda.h) */
    int rc = cc_enc_encode(enc, &in, flags, &pkt);
    gst_video_frame_unmap(&vf);
    if (rc != CC_OK) {
        GST_ELEMENT_ERROR(self, STREAM, ENCODE, ("cc_enc_encode failed: %s", cc_strerror(rc)), (NULL));
        (void)gst_video_encoder_finish_frame(venc, frame);
        return GST_FLOW_ERROR; /* ceracoder turns any bus ERROR into stop() (/root/reference/src/ceracoder.c:425-430) */
    }
    frame->output_buffer = gst_buffer_new_allocate(NULL, (gsize)pkt.size, NULL);
    gst_buffer_fill(frame->output_buffer, 0, pkt.data, (gsize)pkt.size); /* pkt.data is owned by the handle */
    if (pkt.is_idr) GST_VIDEO_CODEC_FRAME_SET_SYNC_POINT(frame); else GST_VIDEO_CODEC_FRAME_UNSET_SYNC_POINT(frame);
    frame->dts = frame->pts; /* no B frames */
    return gst_video_encoder_finish_frame(venc, frame);
}

static void gst_cuda_h264_enc_finalize(GObject *obj) {
    close_codec(GST_CUDA_H264_ENC(obj));
    G_OBJECT_CLASS(gst_cuda_h264_enc_parent_class)->finalize(obj);
}

static void gst_cuda_h264_enc_class_init(GstCudaH264EncClass *klass) {
    GObjectClass *oc = G_OBJECT_CLASS(klass);
    GstElementClass *ec = GST_ELEMENT_CLASS(klass);
    GstVideoEncoderClass *vc = GST_VIDEO_ENCODER_CLASS(klass);
    const GParamFlags rw = (GParamFlags)(G_PARAM_READWRITE | G_PARAM_STATIC_STRINGS);
    const GParamFlags rwp = (GParamFlags)(rw | GST_PARAM_MUTABLE_PLAYING);
    oc->set_property = gst_cuda_h264_enc_set_property;
    oc->get_property = gst_cuda_h264_enc_get_property;
    oc->finalize = gst_cuda_h264_enc_finalize;
    /* `bitrate`: what the docs promise (/root/reference/README.md:256) and x264enc exposes [EXT]; `bps`: what the code writes */
    g_object_class_install_property(oc, PROP_BITRATE, g_param_spec_uint("bitrate", "Bitrate", "Target bitrate in kbit/s (runtime-writable)", 300, 30000, 6000, rwp));
    g_object_class_install_property(oc, PROP_BPS, g_param_spec_int("bps", "Bitrate (ceracoder)", "Target bitrate as ceracoder's encoder_control writes it: kbit/s for an element named venc_kbps, bit/s for venc_bps", 1, G_MAXINT, 6000, rwp));
    g_object_class_install_property(oc, PROP_SPEED_PRESET, g_param_spec_enum("speed-preset", "Speed preset", "x264enc speed-preset (value or nick), accepted for pipeline compatibility", GST_TYPE_CUDA_H264_ENC_PRESET, 3, rw));
    g_object_class_install_property(oc, PROP_KEY_INT_MAX, g_param_spec_uint("key-int-max", "Key-frame interval", "Maximum distance between IDR frames (0 = 60)", 0, G_MAXINT, 60, rw));
    g_object_class_install_property(oc, PROP_TUNE, g_param_spec_flags("tune", "Tune", "x264enc tune flags, accepted for compatibility (zerolatency is always in effect)", GST_TYPE_CUDA_H264_ENC_TUNE, 0x4, rw));
    g_object_class_install_property(oc, PROP_BYTE_STREAM, g_param_spec_boolean("byte-stream", "Byte stream", "Always Annex-B byte-stream", TRUE, rw));
    g_object_class_install_property(oc, PROP_THREADS, g_param_spec_uint("threads", "Threads", "Host threads packaging the slices of one access unit (0 = the streaming thread alone)", 0, 1024, 0, rw));
    g_object_class_install_property(oc, PROP_DEVICE, g_param_spec_int("device", "CUDA device", "CUDA device ordinal", 0, 63, 0, rw));
    g_object_class_install_property(oc, PROP_ROWS_PER_SLICE, g_param_spec_uint("rows-per-slice", "MB rows per slice", "Macroblock rows per slice: 1..8, a multiple of 8, or >= the picture height in macroblocks for one slice per picture (0 = 8)", 0, 1024, 0, rw));
    gst_element_class_set_static_metadata(ec, "CUDA H.264 encoder (ceracuda)", "Codec/Encoder/Video/Hardware",
                                          "H.264 encoder on sm_100a CUDA kernels; drop-in for x264enc in ceracoder pipelines", "ceracuda");
    gst_element_class_add_static_pad_template(ec, &sink_tmpl);
    gst_element_class_add_static_pad_template(ec, &src_tmpl);
    vc->set_format = gst_cuda_h264_enc_set_format;
    vc->handle_frame = gst_cuda_h264_enc_handle_frame;
    vc->stop = gst_cuda_h264_enc_stop;
}
static void gst_cuda_h264_enc_init(GstCudaH264Enc *self) {
    self->bitrate_bps = 6000000; /* /root/reference/src/core/bitrate_control.h:32 */
    self->key_int_max = 60; self->speed_preset = 3; self->tune = 0x4; self->byte_stream = TRUE;
}

static gboolean plugin_init(GstPlugin *plugin) {
    GST_DEBUG_CATEGORY_INIT(cudah264enc_debug, "cudah264enc", 0, "ceracuda H.264 encoder");
    return gst_element_register(plugin, "cudah264enc", GST_RANK_NONE, GST_TYPE_CUDA_H264_ENC);
}
#ifndef PACKAGE
#define PACKAGE "ceracuda"
#endif
GST_PLUGIN_DEFINE(GST_VERSION_MAJOR, GST_VERSION_MINOR, cudah264enc, "ceracuda H.264 encoder", plugin_init, "0.1", "LGPL", "ceracuda", "https://github.com/CERALIVE/ceracoder")
