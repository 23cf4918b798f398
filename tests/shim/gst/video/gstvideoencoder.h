/*
 * tests/shim/gst/video/gstvideoencoder.h — DECLARATION-ONLY stand-in for GstVideoEncoder (see tests/shim/gst/gst.h).
 * TEST INFRASTRUCTURE: no behaviour.
 */
#ifndef SHIM_GST_VIDEO_ENCODER_H
#define SHIM_GST_VIDEO_ENCODER_H
#include <gst/video/video.h>
typedef struct _GstVideoCodecState { gint ref_count; GstVideoInfo info; GstCaps *caps; GstBuffer *codec_data; } GstVideoCodecState;
typedef struct _GstVideoCodecFrame {
    gint ref_count;
    guint flags;
    guint system_frame_number;
    GstClockTime dts, pts, duration;
    GstBuffer *input_buffer, *output_buffer;
} GstVideoCodecFrame;
enum { GST_VIDEO_CODEC_FRAME_FLAG_SYNC_POINT = 1 << 1, GST_VIDEO_CODEC_FRAME_FLAG_FORCE_KEYFRAME = 1 << 2 };
#define GST_VIDEO_CODEC_FRAME_IS_FORCE_KEYFRAME(frame) (((frame)->flags & GST_VIDEO_CODEC_FRAME_FLAG_FORCE_KEYFRAME) != 0)
#define GST_VIDEO_CODEC_FRAME_SET_SYNC_POINT(frame) ((frame)->flags |= GST_VIDEO_CODEC_FRAME_FLAG_SYNC_POINT)
#define GST_VIDEO_CODEC_FRAME_UNSET_SYNC_POINT(frame) ((frame)->flags &= ~(guint)GST_VIDEO_CODEC_FRAME_FLAG_SYNC_POINT)
typedef struct _GstVideoEncoder { GstElement element; } GstVideoEncoder;
typedef struct _GstVideoEncoderClass {
    GstElementClass element_class;
    gboolean (*open)(GstVideoEncoder *encoder);
    gboolean (*close)(GstVideoEncoder *encoder);
    gboolean (*start)(GstVideoEncoder *encoder);
    gboolean (*stop)(GstVideoEncoder *encoder);
    gboolean (*set_format)(GstVideoEncoder *encoder, GstVideoCodecState *state);
    GstFlowReturn (*handle_frame)(GstVideoEncoder *encoder, GstVideoCodecFrame *frame);
    gboolean (*reset)(GstVideoEncoder *encoder, gboolean hard);
    GstFlowReturn (*finish)(GstVideoEncoder *encoder);
} GstVideoEncoderClass;
GType gst_video_encoder_get_type(void);
#define GST_TYPE_VIDEO_ENCODER (gst_video_encoder_get_type())
#define GST_VIDEO_ENCODER_CLASS(klass) ((GstVideoEncoderClass *)(klass))
GstVideoCodecState *gst_video_codec_state_ref(GstVideoCodecState *state);
void gst_video_codec_state_unref(GstVideoCodecState *state);
void gst_video_codec_frame_unref(GstVideoCodecFrame *frame);
GstVideoCodecState *gst_video_encoder_set_output_state(GstVideoEncoder *encoder, GstCaps *caps, GstVideoCodecState *reference);
void gst_video_encoder_set_latency(GstVideoEncoder *encoder, GstClockTime min_latency, GstClockTime max_latency);
gboolean gst_video_encoder_negotiate(GstVideoEncoder *encoder);
GstFlowReturn gst_video_encoder_finish_frame(GstVideoEncoder *encoder, GstVideoCodecFrame *frame);
#endif
